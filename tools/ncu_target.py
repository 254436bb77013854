"""Tiny ncu target: 6 calls of one named workload (sputnik_b200/workloads.py) on rotating operand
sets, or — with several names — ONE call of each (one ncu invocation then captures a launch of
every named workload:  ncu --set full -k regex:bsmm_ -c <n> ... python tools/ncu_target.py a b c).
    ncu --set full -k regex:bsmm_ -s 4 -c 1 ... python tools/ncu_target.py c5_dsd_e8"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from sputnik_b200 import harness, workloads  # noqa: E402

names = sys.argv[1:]
for name in names:
    wl = workloads.get(name)
    prob = harness.DeviceProblem(wl, "cuda:0", sets=2 if len(names) == 1 else 1, seed=3)
    for i in range(6 if len(names) == 1 else 1):
        prob.step(i)
    torch.cuda.synchronize()
    print("ok", wl.describe())
    del prob
    torch.cuda.empty_cache()
