"""Developer tool: A/B environment variants of the library in ONE process (the switches below are
read on every call): us per call, CUDA-graph replay of 100 calls after 10 warm-ups, rotating
operand sets.   python tools/ab.py "BSMM_EPI=staged;BSMM_EPI=direct;BSMM_PREFETCH=0" c2 c4_dsd_25"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from sputnik_b200 import harness, workloads  # noqa: E402
import bench  # noqa: E402

variants = [dict(kv.split("=") for kv in v.split(",") if kv) for v in sys.argv[1].split(";")]
keys = sorted({k for v in variants for k in v})
print(f"{'workload':16s}" + "".join(f"{','.join(f'{k[5:]}={v}' for k, v in var.items()) or 'default':>26s}"
                                    for var in variants))
for rep in range(2):
    for name in sys.argv[2:]:
        wl = workloads.get(name)
        nsets = max(2, min(4, -(-2 * bench.L2_BYTES // wl.algorithmic_bytes)))
        prob = harness.DeviceProblem(wl, "cuda:0", sets=nsets, seed=7)
        row = []
        for var in variants:
            for k in keys:
                os.environ.pop(k, None)
            os.environ.update(var)
            row.append(bench.time_calls(torch, lambda i: prob.step(i)) * 1e3)
        for k in keys:
            os.environ.pop(k, None)
        print(f"{name:16s}" + "".join(f"{v:26.2f}" for v in row), flush=True)
        del prob
        torch.cuda.empty_cache()
