"""Developer tool: in-kernel timeline of one launch.  Builds libbsmm with -DBSMM_TRACE (globaltimer
stamps written per CTA, see BSMM_STAMP in csrc/), runs a workload and prints, per stamp, the
min / median / max over CTAs relative to the earliest CTA start, plus the event-timed duration.

    python tools/trace.py build            (in the container: nvcc)
    python tools/trace.py run c2 [...]     (on the GPU box)
"""
import ctypes
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "tools", "_trace", "libbsmm_trace.so")
NAMES = ["cta start", "setup done", "producer: offsets read", "producer: indices in smem",
         "mma: first stage full", "mma: tile issued (last)", "epi: first tile ready",
         "epi: first tile stored", "epi: last tile stored", "teardown sync passed"]

if sys.argv[1] == "build":
    os.makedirs(os.path.dirname(LIB), exist_ok=True)
    subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3",
                           "-std=c++17", "-shared", "-Xcompiler", "-fPIC", "-DBSMM_TRACE",
                           "-I", os.path.join(ROOT, "include"), "-o", LIB,
                           os.path.join(ROOT, "sputnik_b200", "csrc", "bsmm_api.cu")])
    sys.exit(0)

os.environ["BSMM_LIB"] = LIB
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import sputnik_b200 as sb  # noqa: E402
from sputnik_b200 import harness, workloads  # noqa: E402

lib = sb._lib.lib
lib.bsmm_set_trace.argtypes = [ctypes.c_void_p]
if os.environ.get("BSMM_DBG"):   # experiments: 1 = epilogues skip global stores, 2 = skip TMEM reads
    lib.bsmm_set_dbg.argtypes = [ctypes.c_int]
    lib.bsmm_set_dbg(int(os.environ["BSMM_DBG"]))
    print("BSMM_DBG =", os.environ["BSMM_DBG"])
for name in sys.argv[2:]:
    wl = workloads.get(name)
    prob = harness.DeviceProblem(wl, "cuda:0", sets=2)
    for i in range(5):
        prob.step(i)
    buf = torch.zeros(148 * 16, dtype=torch.int64, device="cuda:0")
    torch.cuda.synchronize()
    lib.bsmm_set_trace(ctypes.c_void_p(buf.data_ptr()))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    prob.step(5)
    e1.record()
    torch.cuda.synchronize()
    lib.bsmm_set_trace(None)
    t = buf.cpu().numpy().reshape(148, 16).astype(np.float64)
    used = t[:, 0] > 0
    t0 = t[used, 0].min()
    print(f"== {wl.describe()}: event time {e0.elapsed_time(e1) * 1e3:.1f} us, {int(used.sum())} CTAs")
    for s, nm in enumerate(NAMES):
        col = t[used, s]
        col = col[col > 0] - t0
        if col.size:
            print(f"  {nm:28s} n={col.size:3d}  min {col.min() / 1e3:7.2f}  med {np.median(col) / 1e3:7.2f}"
                  f"  max {col.max() / 1e3:7.2f} us")
    del prob
