"""Multi-GPU check of the expert-sharded dMoE FFN (launched under torchrun, one rank per GPU,
NCCL): forward and backward through sputnik_b200.moe.ExpertShardedFFN with ragged routing, each
rank verifying ITS token rows against a dense fp64 computation with the gathered expert weights.
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


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local_rank = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local_rank)
    dev = torch.device(f"cuda:{local_rank}")
    dist.init_process_group("nccl", device_id=dev)
    experts, hidden, ffn = 2 * world, 264, 384
    rng = np.random.default_rng(5)
    counts = rng.integers(0, 300, size=(world, experts))
    counts[0, 1] = 0
    plan = moe.ExpertShardPlan(world, rank, counts, hidden, ffn)
    layer = moe.ExpertShardedFFN(plan, torch.bfloat16, dev, seed=rank)
    g = torch.Generator(device=dev).manual_seed(100 + rank)
    x = torch.empty((plan.rows_in, hidden), device=dev).uniform_(-1, 1, generator=g).bfloat16()
    dy = torch.empty((plan.rows_in, hidden), device=dev).uniform_(-1, 1, generator=g).bfloat16()
    y = layer.forward(x)
    dx = layer.backward(dy)
    torch.cuda.synchronize()

    # gather every rank's expert weights: w1 [H, E_l*F] -> all experts
    w1s = [torch.empty_like(layer.w1) for _ in range(world)]
    w2s = [torch.empty_like(layer.w2) for _ in range(world)]
    dist.all_gather(w1s, layer.w1)
    dist.all_gather(w2s, layer.w2)
    w1 = torch.cat(w1s, dim=1).double()
    w2 = torch.cat(w2s, dim=0).double()
    worst = 0.0
    pos = 0
    for e in range(experts):
        n = int(counts[rank, e])
        if n == 0:
            continue
        xe, dye = x[pos:pos + n].double(), dy[pos:pos + n].double()
        w1e, w2e = w1[:, e * ffn:(e + 1) * ffn], w2[e * ffn:(e + 1) * ffn]
        h = (xe @ w1e).bfloat16().double()
        worst = max(worst, rel_err(y[pos:pos + n], h @ w2e))
        dh = (dye @ w2e.T).bfloat16().double()
        worst = max(worst, rel_err(dx[pos:pos + n], dh @ w1e.T))
        pos += n
    # weight gradients of the local experts need every rank's rows: gather x and dy
    sizes = [int(counts[s].sum()) for s in range(world)]
    cap = max(sizes)

    def gather_rows(t):
        padded = torch.zeros((cap, hidden), dtype=t.dtype, device=dev)
        padded[:t.shape[0]] = t
        outs = [torch.empty_like(padded) for _ in range(world)]
        dist.all_gather(outs, padded)
        return [o[:n] for o, n in zip(outs, sizes)]
    xs, dys = gather_rows(x), gather_rows(dy)
    for el in range(plan.local_experts):
        e = plan.expert_lo + el
        xe = torch.cat([xs[s][int(counts[s, :e].sum()):int(counts[s, :e + 1].sum())] for s in range(world)]).double()
        dye = torch.cat([dys[s][int(counts[s, :e].sum()):int(counts[s, :e + 1].sum())] for s in range(world)]).double()
        if xe.shape[0] == 0:
            continue
        w1e, w2e = w1[:, e * ffn:(e + 1) * ffn], w2[e * ffn:(e + 1) * ffn]
        h = (xe @ w1e).bfloat16().double()
        dh = (dye @ w2e.T).bfloat16().double()
        worst = max(worst, rel_err(layer.dw2[el * ffn:(el + 1) * ffn], h.T @ dye))
        worst = max(worst, rel_err(layer.dw1[:, el * ffn:(el + 1) * ffn], xe.T @ dh))
    t = torch.tensor([worst], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ok = float(t) <= 1e-2
    if rank == 0:
        print(f"world {world}: max rel err {float(t):.3e}")
        print("moe_check ok" if ok else "moe_check FAILED")
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
