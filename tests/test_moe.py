"""Expert-sharded dMoE FFN (sputnik_b200/moe.py).

CPU: the shard plan and the token all-to-all (world_size 2, gloo).  GPU: the local SDD -> DSD
chain forward + backward against the oracle (1 GPU), and the sharded layer at world_size 2 over
NCCL against a dense fp64 computation (needs 2 GPUs; skipped otherwise)."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from sputnik_b200 import moe

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_uniform_plan_matches_baseline_config():
    # BASELINE.json configs[4]: 64 experts, hidden 4096, ffn 14336, 65536 tokens, on 8 GPUs
    p = moe.ExpertShardPlan.uniform(8, 3, 64, 65536, 4096, 14336)
    assert p.local_experts == 8 and (p.expert_lo, p.expert_hi) == (24, 32)
    assert p.rows_in == 8192 and p.rows_local == 8192
    assert p.send_splits == [1024] * 8 and p.recv_splits == [1024] * 8
    t = p.local_topology()
    assert (t.block_rows, t.block_cols, t.nnz_blocks) == (64, 8 * 112, 64 * 112)
    assert p.flops(backward=False) == 2 * 2.0 * 57344 * 128 * 128 * 4096
    # one GPU: the whole SURVEY 8d topology
    p1 = moe.ExpertShardPlan.uniform(1, 0, 64, 65536, 4096, 14336)
    t1 = p1.local_topology()
    assert (t1.block_rows, t1.block_cols, t1.nnz_blocks) == (512, 7168, 57344)
    assert p1.is_identity


def test_ragged_plan_pads_each_expert_to_blocks():
    counts = np.array([[100, 0, 300, 28], [60, 5, 0, 100]])
    p0 = moe.ExpertShardPlan(2, 0, counts, 64, 256)   # owns experts 0, 1
    p1 = moe.ExpertShardPlan(2, 1, counts, 64, 256)   # owns experts 2, 3
    assert p0.send_splits == [100, 328] and p0.recv_splits == [100, 65]
    assert p1.send_splits == [65, 100] and p1.recv_splits == [328, 100]
    assert p0.block_rows_per_expert == [2, 1] and p0.rows_local == 384
    assert p1.block_rows_per_expert == [3, 1] and p1.rows_local == 512
    # rank 0, recv order [src0: e0 x100, e1 x0][src1: e0 x60, e1 x5] -> e0 rows 0..159, e1 at 256
    want = np.concatenate([np.arange(100), 100 + np.arange(60), 256 + np.arange(5)])
    assert np.array_equal(p0.recv_to_local, want)
    t = p1.local_topology()
    assert t.block_rows == 4 and t.block_cols == 4
    assert list(t.offsets) == [0, 2, 4, 6, 8] and list(t.indices) == [0, 1, 0, 1, 0, 1, 2, 3]


def _gloo_worker(rank, world, port, counts, hidden):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        plan = moe.ExpertShardPlan(world, rank, counts, hidden, 128)
        # every token row carries (source rank, global expert, ordinal) so placement is checkable
        rows = []
        for e in range(counts.shape[1]):
            for i in range(counts[rank, e]):
                rows.append([rank, e, i] + [0] * (hidden - 3))
        x = torch.tensor(rows, dtype=torch.float32).reshape(-1, hidden)
        local = moe.dispatch(x, plan)
        assert local.shape == (plan.rows_local, hidden)
        pos = 0
        for el, nblk in enumerate(plan.block_rows_per_expert):
            e = plan.expert_lo + el
            for s in range(world):
                for i in range(counts[s, e]):
                    assert local[pos, :3].tolist() == [s, e, i], (rank, pos)
                    pos += 1
            pad = nblk * 128 - int(counts[:, e].sum())
            assert float(local[pos:pos + pad].abs().sum()) == 0.0
            pos += pad
        back = moe.combine(local * 2.0, plan)
        assert torch.equal(back, x * 2.0)
    finally:
        dist.destroy_process_group()


def test_dispatch_combine_world2_gloo():
    counts = np.array([[100, 0, 300, 28], [60, 5, 0, 100]])
    mp.spawn(_gloo_worker, args=(2, 29533, counts, 8), nprocs=2, join=True)


def test_dispatch_combine_uniform_world2_gloo():
    counts = np.full((2, 4), 128)
    mp.spawn(_gloo_worker, args=(2, 29534, counts, 8), nprocs=2, join=True)


# ------------------------------------------------------------------------------------- GPU
@pytest.mark.parametrize("world,experts", [(1, 4), (2, 4), (4, 8)])
def test_return_map_inverts_dispatch(world, experts):
    """Simulate the token all-to-all for every rank in one process: each local row's
    (source rank, source row) label must be what return_map reports; pad rows map to (-1, -1)."""
    rng = np.random.default_rng(world)
    counts = rng.integers(0, 300, size=(world, experts))
    plans = [moe.ExpertShardPlan(world, r, counts, 64, 128) for r in range(world)]
    for g, pl in enumerate(plans):
        recv = []
        for s in range(world):
            off = sum(plans[s].send_splits[:g])
            recv += [(s, off + i) for i in range(plans[s].send_splits[g])]
        local = [(-1, -1)] * pl.rows_local
        for i, l in enumerate(pl.recv_to_local):
            local[l] = recv[i]
        dst_rank, dst_row = pl.return_map()
        assert [(int(a), int(b)) for a, b in zip(dst_rank, dst_row)] == local
        assert pl.max_rows_in == int(counts.sum(axis=1).max())


@pytest.mark.gpu
@pytest.mark.parametrize("counts", [np.full((1, 4), 256), np.array([[200, 0, 130, 384]])],
                         ids=["uniform", "ragged"])
@pytest.mark.parametrize("layout", ["fh", "hf"])
def test_fused_return_exchange_equals_combine(counts, layout):
    """world = 1: the scatter epilogue (un-pad + un-permute inside y = dsd and dx = dsd) gives
    the same bits as DSD + combine.  (world = 2 over NVLink: tools/moe_check.py.)"""
    plan = moe.ExpertShardPlan(1, 0, counts, 264, 384)
    plain = moe.ExpertShardedFFN(plan, torch.bfloat16, "cuda:0", seed=1, w1_layout=layout)
    fused = moe.ExpertShardedFFN(plan, torch.bfloat16, "cuda:0", seed=1, fused_combine=True,
                                 w1_layout=layout)
    assert torch.equal(plain.w1, fused.w1)
    g = torch.Generator(device="cuda:0").manual_seed(7)
    x = torch.empty((plan.rows_in, 264), device="cuda:0").uniform_(-1, 1, generator=g).bfloat16()
    dy = torch.empty((plan.rows_in, 264), device="cuda:0").uniform_(-1, 1, generator=g).bfloat16()
    y, dx = plain.forward(x), plain.backward(dy)
    yf, dxf = fused.forward(x), fused.backward(dy)
    torch.cuda.synchronize()
    assert torch.equal(yf, y) and torch.equal(dxf, dx)
    assert torch.equal(fused.dw1, plain.dw1) and torch.equal(fused.dw2, plain.dw2)


@pytest.mark.gpu
@pytest.mark.parametrize("hidden,ffn", [(136, 256), (264, 384)])
@pytest.mark.parametrize("counts", [np.full((1, 4), 256), np.array([[200, 0, 130, 384]])],
                         ids=["uniform", "ragged"])
@pytest.mark.parametrize("layout", ["fh", "hf"])
def test_local_chain_matches_oracle(counts, hidden, ffn, layout):
    """Both storage layouts of w1 ([E*ffn, hidden]: sdd nt / dsd nn / dsd tn; [hidden, E*ffn]:
    sdd nn / dsd nt / dds tn) against the same oracle products."""
    from oracle import oracle
    from util import assert_close
    plan = moe.ExpertShardPlan(1, 0, counts, hidden, ffn)
    layer = moe.ExpertShardedFFN(plan, torch.bfloat16, "cuda:0", seed=1, w1_layout=layout)
    assert layer.w1.shape == ((plan.local_experts * ffn, hidden) if layout == "fh"
                              else (hidden, plan.local_experts * ffn))
    g = torch.Generator(device="cuda:0").manual_seed(7)
    x = torch.empty((plan.rows_in, hidden), device="cuda:0").uniform_(-1, 1, generator=g).bfloat16()
    dy = torch.empty((plan.rows_in, hidden), device="cuda:0").uniform_(-1, 1, generator=g).bfloat16()
    y = layer.forward(x)
    dx = layer.backward(dy)
    torch.cuda.synchronize()
    f = lambda t: t.float().cpu().numpy()
    topo = plan.local_topology()
    xl, dyl = f(layer.x), f(layer.dy)
    w1, w2 = f(layer.w1_hf), f(layer.w2)
    h_ref = oracle.sdd(xl, False, w1, False, topo)
    assert_close(f(layer.h.data), h_ref, "bf16", "h = sdd(x, w1)")
    h16 = f(layer.h.data)                       # downstream products see the rounded h
    assert_close(f(layer.y), oracle.dsd(h16, topo, False, w2, False), "bf16", "y = dsd(h, w2)")
    dh_ref = oracle.sdd(dyl, False, w2, True, topo)
    assert_close(f(layer.dh.data), dh_ref, "bf16", "dh = sdd(dy, w2^T)")
    dh16 = f(layer.dh.data)
    assert_close(f(layer.dw2), oracle.dsd(h16, topo, True, dyl, False), "bf16", "dw2")
    assert_close(f(layer.dx), oracle.dsd(dh16, topo, False, w1, True), "bf16", "dx")
    assert_close(f(layer.dw1_hf), oracle.dds(xl, True, dh16, topo, False), "bf16", "dw1")
    # the caller's view: rows in, rows out, pad rows dropped
    assert y.shape == x.shape and dx.shape == x.shape
    idx = torch.from_numpy(plan.recv_to_local).cuda()
    assert torch.equal(y, layer.y[idx]) and torch.equal(dx, layer.dx[idx])


@pytest.mark.gpu
def test_sharded_layer_two_gpus_nccl():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    out = subprocess.run(
        [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
         "--master-addr", "127.0.0.1", "--master-port", "29541",
         os.path.join(ROOT, "tools", "moe_check.py")],
        capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert "moe_check ok" in out.stdout


@pytest.mark.gpu
def test_device_routed_layer_matches_host_planned_layer_and_tables():
    """bsmm_moe_exchange: the gather / scatter tables built on the device equal the host plan's
    return_map (world 1 here; world 2 / 8 through tools/moe_check.py in bench.py's parity gate),
    and the device-routed layer reproduces the host-planned one bit for bit with no host
    synchronisation in the step."""
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    import moe_check
    dev = torch.device("cuda:0")
    res = moe_check.check(0, 1, dev, experts_per_rank=5, hidden=264, ffn=512, seed=11)
    assert res["ok"], res
    # the tables themselves, against the pure-host plan (simulated world of 4, this rank = 2)
    import sputnik_b200 as sb
    world, rank, experts = 4, 2, 8
    rng = np.random.default_rng(3)
    counts = rng.integers(0, 400, size=(world, experts))
    counts[1, 4] = 0
    plan = moe.ExpertShardPlan(world, rank, counts, 64, 256)
    le = plan.local_experts
    per = counts[:, rank * le:(rank + 1) * le].sum(axis=0)
    cap = int(sum(-(-int(v) // 256) * 2 for v in per)) + 2
    tab = torch.full((cap * 128,), -1, dtype=torch.int64, device=dev)
    tok = torch.zeros(le, dtype=torch.int32, device=dev)
    bases = [[(s + 1) << 40 for s in range(world)]]
    pitch = 64 * 2
    sb.moe_exchange(torch.from_numpy(counts.astype(np.int32)).to(dev), rank, cap, pitch, bases, [tab], tok)
    torch.cuda.synchronize()
    assert np.array_equal(tok.cpu().numpy(), per.astype(np.int32))
    got = tab.cpu().numpy()
    # expected from the plan: local rows are padded per expert to 256 here (128 in the host plan)
    dst_rank, dst_row = plan.return_map()
    exp = np.zeros(cap * 128, np.int64)
    src = 0
    dstp = 0
    for e in range(le):
        n = int(per[e])
        rows128 = -(-n // 128) * 128
        seg_rank, seg_row = dst_rank[src:src + n], dst_row[src:src + n]
        exp[dstp:dstp + n] = ((seg_rank + 1) << 40) + seg_row * pitch
        src += rows128
        dstp += -(-n // 256) * 256
    assert np.array_equal(got, exp)
