"""What the fused epilogues buy, on one B200:

    python tools/epilogue_bench.py [--iters 20]

 * weight-gradient accumulation  dw += dsd(h^T, dy)  /  dw += dds(x^T, dh):  BSMM_FLAG_ACCUMULATE
   against the unfused sequence (product into a temporary, then `dw.add_(tmp)`);
 * dMoE MLP activation  h = gelu(sdd(x, w1)):  BSMM_FLAG_GELU against product + F.gelu(tanh).
CUDA events around `iters` back-to-back repetitions after warm-up; median of 5."""
import argparse
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sputnik_b200 as sb  # noqa: E402
from sputnik_b200 import harness, workloads  # noqa: E402


def timed(fn, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    out = []
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        out.append(e0.elapsed_time(e1) * 1e3 / iters)
    return float(np.median(out))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=20)
    a = ap.parse_args()
    print(f"{'workload':<16}{'epilogue':<12}{'fused us':>10}{'unfused us':>12}{'x':>7}   (plain product us)")
    for name, kind in [("c3a", "acc"), ("c3b", "acc"), ("c4_dsd_50", "acc"), ("c5_dsd_tn_e8", "acc"),
                       ("c5_dds_tn_e8", "acc"), ("c4_sdd_50", "gelu"), ("c5_sdd_e8", "gelu")]:
        wl = workloads.get(name)
        prob = harness.DeviceProblem(wl, "cuda:0", sets=1)
        s = prob.sets[0]
        out = s.output()
        out.zero_()
        args = ((s.a, wl.ta, s.b, wl.tb, s.sparse) if wl.op == "sdd" else
                (s.sparse, wl.ta, s.b, wl.tb, s.c) if wl.op == "dsd" else
                (s.a, wl.ta, s.sparse, wl.tb, s.c))
        plain = timed(lambda: sb.matmul_ex(*args), a.iters)
        if kind == "acc":
            acc = torch.zeros_like(out)
            acc_m = sb.Matrix.of(acc)
            fargs = args[:4] + (acc_m,)
            fused = timed(lambda: sb.matmul_ex(*fargs, accumulate=True), a.iters)

            def unfused():
                sb.matmul_ex(*args)
                acc.add_(out)
            unf = timed(unfused, a.iters)
        else:
            fused = timed(lambda: sb.matmul_ex(*args, activation="gelu"), a.iters)
            act = torch.empty_like(out)

            def unfused():
                sb.matmul_ex(*args)
                torch.nn.functional.gelu(out, approximate="tanh", out=act) if False else \
                    act.copy_(torch.nn.functional.gelu(out, approximate="tanh"))
            unf = timed(unfused, a.iters)
        print(f"{name:<16}{kind:<12}{fused:10.1f}{unf:12.1f}{unf / fused:7.2f}   ({plain:.1f})", flush=True)
        del prob, s, out
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
