"""Multi-GPU check of the expert-sharded dMoE FFN (launched under torchrun, one rank per GPU,
NCCL): forward and backward through sputnik_b200.moe.ExpertShardedFFN with ragged routing.

Checked on every rank:
  1. dispatch: the local expert rows are exactly the rows the source ranks hold (bit-exact);
  2. the local chain, product by product, against dense fp64 matmuls per expert fed with the SAME
     16-bit intermediates the device produced (max relative error <= 1e-2, tau guard of
     tests/util.py);
  3. combine: the rows this rank gets back are exactly the owners' output rows (bit-exact);
  4. the fused return exchange (bsmm_dsd_scatter storing into the source ranks' symmetric-memory
     buffers over NVLink, no all-to-all): y and dx bit-identical to the NCCL path, twice in a row.
Prints "moe_check ok" on rank 0."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from sputnik_b200 import moe  # noqa: E402


def rel_err(got, ref):
    ref = ref.double()
    tau = 1e-2 * ref.pow(2).mean().sqrt()
    return float(((got.double() - ref).abs() / torch.maximum(ref.abs(), tau)).max())


def device_routed_equals(layer, plan, counts, x, dy, y, dx, rank, world, dev):
    """DeviceRoutedFFN on the same weights and routing: y, dx, dw1, dw2 bit-identical to the
    host-planned layer's, with the whole step (route + forward + backward) run under
    torch.cuda.set_sync_debug_mode("error") — any host synchronisation raises."""
    experts = counts.shape[1]
    per_local = counts.reshape(world, world, plan.local_experts).sum(axis=0)   # [owner, e_local]
    cap = int(max(sum(-(-int(v) // 256) * 2 for v in per_local[g]) for g in range(world))) + 2
    cap += cap % 2
    drl = moe.DeviceRoutedFFN(world, rank, experts, plan.hidden, plan.ffn, plan.max_rows_in, cap,
                              torch.bfloat16, dev, group=layer.group, weights_from=layer)
    my_counts = torch.from_numpy(counts[rank].astype(np.int32)).to(dev)
    n = plan.rows_in
    drl.x_in[:n].copy_(x)
    drl.dy_in[:n].copy_(dy)
    dw1_ref, dw2_ref = layer.dw1.clone(), layer.dw2.clone()
    torch.cuda.synchronize()
    same = True
    for it in range(2):
        torch.cuda.set_sync_debug_mode("error")
        try:
            drl.route(my_counts)
            yd = drl.forward(n)
            dxd = drl.backward(n)
        finally:
            torch.cuda.set_sync_debug_mode("default")
        torch.cuda.synchronize()
        ok = (bool(torch.equal(yd, y)) and bool(torch.equal(dxd, dx))
              and bool(torch.equal(drl.dw1, dw1_ref)) and bool(torch.equal(drl.dw2, dw2_ref)))
        if not ok:
            print(f"rank {rank}: device-routed layer differs from the host-planned one (iteration {it}): "
                  f"y {bool(torch.equal(yd, y))} dx {bool(torch.equal(dxd, dx))} "
                  f"dw1 {bool(torch.equal(drl.dw1, dw1_ref))} dw2 {bool(torch.equal(drl.dw2, dw2_ref))}",
                  flush=True)
        same &= ok
    if world > 1:
        dist.barrier()
    del drl
    return same


def check(rank, world, dev, experts_per_rank=2, hidden=264, ffn=384, seed=5):
    """Runs the comparison on an initialised process group (or world == 1 without one), once per
    storage layout of w1 (moe.py: "fh" and "hf" — between them every product kind the layer can
    issue).  Returns {"checked", "max_rel_err", "exchange_bit_exact", "fused_equals_nccl", "ok"} —
    the same on every rank (all-reduced); worst case over the layouts."""
    res = None
    for layout in ("fh", "hf"):
        r = _check_one(rank, world, dev, experts_per_rank, hidden, ffn, seed, layout)
        if res is None:
            res = r
        else:
            res["max_rel_err"] = max(res["max_rel_err"], r["max_rel_err"])
            for k in ("exchange_bit_exact", "fused_equals_nccl", "device_routed_equals_host_routed", "ok"):
                res[k] = res[k] and r[k]
    res["w1_layouts"] = ["fh", "hf"]
    return res


def _check_one(rank, world, dev, experts_per_rank, hidden, ffn, seed, layout):
    experts = experts_per_rank * world
    fc = ffn // 128
    rng = np.random.default_rng(seed)
    counts = rng.integers(0, 300, size=(world, experts))
    counts[0, 1] = 0
    plan = moe.ExpertShardPlan(world, rank, counts, hidden, ffn)
    layer = moe.ExpertShardedFFN(plan, torch.bfloat16, dev, seed=rank, fused_combine=True,
                                 w1_layout=layout)
    g = torch.Generator(device=dev).manual_seed(100 + rank)
    x = torch.empty((plan.rows_in, hidden), device=dev).uniform_(-1, 1, generator=g).bfloat16()
    dy = torch.empty((plan.rows_in, hidden), device=dev).uniform_(-1, 1, generator=g).bfloat16()
    y = layer.forward(x, fused=False)
    dx = layer.backward(dy, fused=False)
    torch.cuda.synchronize()

    sizes = [int(counts[s].sum()) for s in range(world)]

    def gather_rows(t, ns):
        if world == 1:
            return [t[:ns[0]]]
        cap = max(ns)
        padded = torch.zeros((cap, t.shape[1]), dtype=t.dtype, device=dev)
        padded[:t.shape[0]] = t
        outs = [torch.empty_like(padded) for _ in range(world)]
        dist.all_gather(outs, padded)
        return [o[:n] for o, n in zip(outs, ns)]
    xs, dys = gather_rows(x, sizes), gather_rows(dy, sizes)
    src_rows = lambda ts, e: torch.cat(
        [ts[s][int(counts[s, :e].sum()):int(counts[s, :e + 1].sum())] for s in range(world)])

    exact, errs = True, {}
    row0 = 0
    for el, nblk in enumerate(plan.block_rows_per_expert):
        e = plan.expert_lo + el
        xe, dye = src_rows(xs, e), src_rows(dys, e)
        n = xe.shape[0]
        exact &= bool(torch.equal(layer.x[row0:row0 + n], xe))          # 1. dispatch
        exact &= bool(torch.equal(layer.dy[row0:row0 + n], dye))
        if n == 0:
            row0 += nblk * 128
            continue
        w1e = layer.w1_hf[:, el * ffn:(el + 1) * ffn].double()
        w2e = layer.w2[el * ffn:(el + 1) * ffn].double()
        # the expert's sparse blocks as a dense [rows, ffn] matrix
        blocks = lambda t: t.data[(row0 // 128) * fc:(row0 // 128 + nblk) * fc] \
            .reshape(nblk, fc, 128, 128).permute(0, 2, 1, 3).reshape(nblk * 128, ffn)[:n]
        h, dh = blocks(layer.h), blocks(layer.dh)
        xe, dye = xe.double(), dye.double()
        errs[f"h e{e}"] = rel_err(h, xe @ w1e)                           # 2. the chain
        errs[f"y e{e}"] = rel_err(layer.y[row0:row0 + n], h.double() @ w2e)
        errs[f"dh e{e}"] = rel_err(dh, dye @ w2e.T)
        errs[f"dw2 e{e}"] = rel_err(layer.dw2[el * ffn:(el + 1) * ffn], h.double().T @ dye)
        errs[f"dx e{e}"] = rel_err(layer.dx[row0:row0 + n], dh.double() @ w1e.T)
        errs[f"dw1 e{e}"] = rel_err(layer.dw1_hf[:, el * ffn:(el + 1) * ffn], xe.T @ dh.double())
        row0 += nblk * 128

    # 3. combine: my rows of y / dx == the owners' rows
    local_sizes = []
    for gk in range(world):
        lo = gk * plan.local_experts
        local_sizes.append(int(sum(-(-int(v) // 128) * 128
                                   for v in counts[:, lo:lo + plan.local_experts].sum(axis=0))))
    ys, dxs = gather_rows(layer.y, local_sizes), gather_rows(layer.dx, local_sizes)
    pos = 0
    for e in range(experts):
        owner, el = divmod(e, plan.local_experts)
        lo = owner * plan.local_experts
        start = sum(-(-int(v) // 128) * 128 for v in counts[:, lo:lo + el].sum(axis=0))
        start += int(counts[:rank, e].sum())
        n = int(counts[rank, e])
        exact &= bool(torch.equal(y[pos:pos + n], ys[owner][start:start + n]))
        exact &= bool(torch.equal(dx[pos:pos + n], dxs[owner][start:start + n]))
        pos += n

    # 4. fused exchanges on the SAME layer: peer-memory row gather in, DSD scatter epilogue out
    y, dx = y.clone(), dx.clone()
    fused_same = True
    for it in range(2):
        yf = layer.forward(x, fused=True).clone()
        dxf = layer.backward(dy, fused=True).clone()
        torch.cuda.synchronize()
        same = bool(torch.equal(yf, y)) and bool(torch.equal(dxf, dx))
        if not same:
            print(f"rank {rank}: fused exchange differs from NCCL path (iteration {it})", flush=True)
        fused_same &= same
    if world > 1:
        dist.barrier()

    # 5. device-routed layer (routing counts only on the device: bsmm_moe_exchange +
    #    bsmm_moe_topology, no host synchronisation in the step) == the host-planned layer
    dev_same = device_routed_equals(layer, plan, counts, x, dy, y, dx, rank, world, dev)

    worst = max(errs.values()) if errs else 0.0
    bad = {k: round(v, 4) for k, v in errs.items() if v > 1e-2}
    if bad or not exact:
        print(f"rank {rank}: exchange exact={exact}; over tolerance: {bad}", flush=True)
    t = torch.tensor([worst, 0.0 if exact else 1.0, 0.0 if fused_same else 1.0,
                      0.0 if dev_same else 1.0], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    res = {"checked": True, "max_rel_err": float(t[0]), "tolerance": 1e-2,
           "exchange_bit_exact": float(t[1]) == 0.0, "fused_equals_nccl": float(t[2]) == 0.0,
           "device_routed_equals_host_routed": float(t[3]) == 0.0,
           "what": f"ragged routing, {experts} experts, hidden {hidden}, ffn {ffn}: dispatch / combine "
                   "bit-exact vs the source ranks' rows; 6 products per local expert vs fp64 matmuls of "
                   "the same 16-bit operands; fused (peer-memory gather + DSD scatter epilogue) y / dx "
                   "bit-equal to the NCCL all-to-all path, twice"}
    res["ok"] = (res["max_rel_err"] <= 1e-2 and res["exchange_bit_exact"] and res["fused_equals_nccl"]
                 and res["device_routed_equals_host_routed"])
    del layer
    torch.cuda.empty_cache()
    return res


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local_rank = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local_rank)
    dev = torch.device(f"cuda:{local_rank}")
    dist.init_process_group("nccl", device_id=dev)
    res = check(rank, world, dev)
    if rank == 0:
        print(f"world {world}: exchange bit-exact={res['exchange_bit_exact']}, fused == nccl: "
              f"{res['fused_equals_nccl']}, device-routed == host-routed: "
              f"{res['device_routed_equals_host_routed']}, chain max rel err {res['max_rel_err']:.3e}")
        print("moe_check ok" if res["ok"] else "moe_check FAILED")
    dist.destroy_process_group()
    sys.exit(0 if res["ok"] else 1)


if __name__ == "__main__":
    main()
