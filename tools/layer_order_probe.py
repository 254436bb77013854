"""Developer tool: the C5 layer at N = 1, fused (identity exchange) and general (dispatch / combine
copies) paths timed alternately on one box, to separate path cost from measurement-order effects
under the power cap.   python tools/layer_order_probe.py [steps] [rounds]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from sputnik_b200 import moe  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
rounds = int(sys.argv[2]) if len(sys.argv) > 2 else 4
dev = torch.device("cuda:0")
E, T, H, F = 64, 65536, 4096, 14336
plan = moe.ExpertShardPlan.uniform(1, 0, E, T, H, F)
layer = moe.ExpertShardedFFN(plan, torch.bfloat16, dev, seed=0, fused_combine=True)
g = torch.Generator(device=dev).manual_seed(0)
mk = lambda: torch.empty((plan.rows_in, H), device=dev).uniform_(-1, 1, generator=g).bfloat16()
layer.x_in[:plan.rows_in].copy_(mk())
layer.dy_in[:plan.rows_in].copy_(mk())
x_f, dy_f = layer.x_in[:plan.rows_in], layer.dy_in[:plan.rows_in]
x_n, dy_n = x_f.clone(), dy_f.clone()
y, dx = torch.empty_like(x_n), torch.empty_like(x_n)
flops = 6 * 2.0 * T * H * F


def fused():
    layer.forward(x_f, fused=True)
    layer.backward(dy_f, fused=True)


def general():
    layer.forward(x_n, out=y, fused=False)
    layer.backward(dy_n, out=dx, fused=False)


def timed(step):
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


for r in range(rounds):
    a = timed(fused)
    layer.timers = {}          # bench.py's per-product CUDA events around every product
    c = timed(fused)
    layer.timers = None
    b = timed(general)
    print(f"round {r}: fused/identity {a:7.2f} ms ({flops / a / 1e9:6.0f} TF)   with per-product events "
          f"{c:7.2f} ms ({flops / c / 1e9:6.0f} TF)   general {b:7.2f} ms ({flops / b / 1e9:6.0f} TF)",
          flush=True)
