"""Developer tool: y = dsd(h, w) with scattered output rows (bsmm_dsd_scatter, identity row table)
against the plain product, per epilogue variant.  python tools/scatter_ab.py c5_dsd_e8 c5_dsd_nt_e8"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import sputnik_b200 as sb  # noqa: E402
from sputnik_b200 import harness, workloads  # noqa: E402
import bench  # noqa: E402

for rep in range(2):
    for name in sys.argv[1:]:
        wl = workloads.get(name)
        prob = harness.DeviceProblem(wl, "cuda:0", sets=2, seed=7)
        ptrs = [sb.row_pointers(s.c.data) for s in prob.sets]

        def scat(i):
            s = prob.sets[i % 2]
            sb.dsd_scatter(s.sparse, wl.ta, s.b, wl.tb, ptrs[i % 2], wl.n)
        row = [bench.time_calls(torch, lambda i: prob.step(i), 5, 20) * 1e3]
        for mode in ("staged", "direct"):
            os.environ["BSMM_SCATTER_EPI"] = mode
            row.append(bench.time_calls(torch, scat, 5, 20) * 1e3)
        os.environ.pop("BSMM_SCATTER_EPI")
        print(f"{name:16s} plain (TMA store) {row[0]:9.1f}   scatter staged {row[1]:9.1f}   scatter direct {row[2]:9.1f} us",
              flush=True)
        del prob
