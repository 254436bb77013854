"""Tiny ncu target: 6 calls of one named workload (sputnik_b200/workloads.py) on rotating operand
sets.   ncu --set full -k regex:bsmm_ -s 4 -c 1 ... python tools/ncu_target.py c5_dsd_e8"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from sputnik_b200 import harness, workloads  # noqa: E402

wl = workloads.get(sys.argv[1])
prob = harness.DeviceProblem(wl, "cuda:0", sets=2, seed=3)
for i in range(6):
    prob.step(i)
torch.cuda.synchronize()
print("ok", wl.describe())
