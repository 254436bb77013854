"""Developer tool (trace build only): time workloads with the epilogue experiments of
g_bsmm_dbg — bit 0 = no global stores, bit 1 = no TMEM reads, bit 2 = no conversion / staging into
shared memory (TMA-store path) — to see what the epilogue is bound by.  python tools/dbg_ab.py c4_dsd_10 c4_dsd_25 c2"""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.environ["BSMM_LIB"] = os.path.join(ROOT, "tools", "_trace", "libbsmm_trace.so")
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import sputnik_b200 as sb  # noqa: E402
from sputnik_b200 import harness, workloads  # noqa: E402
import bench  # noqa: E402

lib = sb._lib.lib
lib.bsmm_set_dbg.argtypes = [ctypes.c_int]
print(f"{'workload':14s}" + "".join(f"{'dbg=' + str(d):>10s}" for d in range(8)) + "   (us per call)")
for name in sys.argv[1:]:
    wl = workloads.get(name)
    prob = harness.DeviceProblem(wl, "cuda:0", sets=3, seed=7)
    row = []
    for d in range(8):
        lib.bsmm_set_dbg(d)
        row.append(bench.time_calls(torch, lambda i: prob.step(i)) * 1e3)
    lib.bsmm_set_dbg(0)
    print(f"{name:14s}" + "".join(f"{v:10.2f}" for v in row), flush=True)
    del prob
