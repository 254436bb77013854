"""Per-phase device time of one fused / NCCL dMoE layer step (torchrun, one rank per GPU).
Developer tool: CUDA events around every call of ExpertShardedFFN.forward/backward."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from sputnik_b200 import block as sb  # noqa: E402
from sputnik_b200 import moe  # noqa: E402


def main():
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG", "WARN")
        dist.init_process_group("nccl", device_id=dev)
    fused = "--nccl" not in sys.argv
    plan = moe.ExpertShardPlan.uniform(world, rank, 64, 65536, 4096, 14336)
    layer = moe.ExpertShardedFFN(plan, torch.bfloat16, dev, seed=rank, fused_combine=fused)
    x = torch.randn((plan.rows_in, 4096), device=dev).bfloat16()
    dy = torch.randn((plan.rows_in, 4096), device=dev).bfloat16()
    if fused:
        layer.x_in[:plan.rows_in].copy_(x)
        layer.dy_in[:plan.rows_in].copy_(dy)
        x, dy = layer.x_in[:plan.rows_in], layer.dy_in[:plan.rows_in]
    if "--series" in sys.argv:
        # step time of each of N consecutive steps (one event per step boundary, nothing else in
        # the stream): does a step slow down as the run goes on (power cap), or is it flat?
        n = int(sys.argv[sys.argv.index("--series") + 1])
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(n + 1)]
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        evs[0].record()
        for i in range(n):
            layer.forward(x)
            layer.backward(dy)
            evs[i + 1].record()
        torch.cuda.synchronize()
        t = torch.tensor([evs[i].elapsed_time(evs[i + 1]) for i in range(n)], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            print(f"world {world} {'fused' if fused else 'nccl'}: per-step ms (max over ranks): "
                  + " ".join(f"{v:.2f}" for v in t.tolist()))
            print(f"  first 5 mean {float(t[:5].mean()):.3f}  last 5 mean {float(t[-5:].mean()):.3f}  "
                  f"all {float(t.mean()):.3f}")
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return
    marks = []

    def wrap(obj, name, label=None):
        fn = getattr(obj, name)

        def timed(*a, **k):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            out = fn(*a, **k)
            e1.record()
            tag = label or name
            if name == "matmul_ex":
                kinds = ["S" if isinstance(v, sb.BlockMatrix) else "D" for v in (a[0], a[2], a[4])]
                tag = "".join(kinds).lower().replace("dds", "sdd-out" if False else "dds") + \
                    ("_t" if a[1] else "_n") + ("t" if a[3] else "n")
            marks.append((tag, e0, e1))
            return out
        setattr(obj, name, timed)
    wrap(layer.sb, "matmul_ex")
    wrap(layer.sb, "dsd_scatter")
    wrap(layer.sb, "gather_rows")
    wrap(layer, "_peer_barrier", "barrier")
    wrap(moe, "dispatch")
    wrap(moe, "combine")
    for it in range(6):
        marks.clear()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        layer.forward(x)
        layer.backward(dy)
        t1.record()
        torch.cuda.synchronize()
    if rank == 0:
        total = t0.elapsed_time(t1)
        print(f"world {world} {'fused' if fused else 'nccl'}: step {total:.3f} ms")
        acc = 0.0
        for tag, e0, e1 in marks:
            ms = e0.elapsed_time(e1)
            acc += ms
            print(f"  {tag:14s} {ms * 1e3:9.1f} us")
        print(f"  sum of phases {acc:.3f} ms, gaps {total - acc:.3f} ms")
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
