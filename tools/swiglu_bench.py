"""Developer tool: fused SwiGLU SDD (BSMM_FLAG_SWIGLU, 1-SM kernel) against the unfused sequence
(two SDD products on the 2-SM kernel + an elementwise silu(g) * u) on a dMoE shape.
    python tools/swiglu_bench.py [experts] [nn|nt]      (nt: weights stored [2 E F, hidden])"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402
import sputnik_b200 as sb  # noqa: E402
from sputnik_b200 import synth  # noqa: E402
import bench  # noqa: E402

E = int(sys.argv[1]) if len(sys.argv) > 1 else 8
NT = len(sys.argv) > 2 and sys.argv[2] == "nt"
H, FF, R = 4096, 14336, 8
dev = "cuda:0"
topo = synth.moe_topology(E, R, FF // 128)
g = torch.Generator(device=dev).manual_seed(0)
rnd = lambda *s: torch.empty(s, device=dev).uniform_(-1, 1, generator=g).bfloat16()
x = rnd(topo.rows, H)
w_gu = rnd(H, 2 * topo.cols) * 0.02           # gate | up interleaved by block column
w_g = w_gu.view(H, topo.cols // 128, 2, 128)[:, :, 0].reshape(H, topo.cols).contiguous()
w_u = w_gu.view(H, topo.cols // 128, 2, 128)[:, :, 1].reshape(H, topo.cols).contiguous()


def blocks():
    m = sb.BlockMatrix(topo.rows, topo.cols, 128, topo.nonzeros,
                       torch.empty((topo.nnz_blocks, 128, 128), dtype=torch.bfloat16, device=dev),
                       torch.from_numpy(topo.offsets).to(dev), torch.from_numpy(topo.indices).to(dev))
    sb.allocate_row_indices_buffer(m)
    sb.row_indices(m)
    return m


if NT:   # K-major storage of op(B): [n][k]
    w_gu, w_g, w_u = (t.t().contiguous() for t in (w_gu, w_g, w_u))
h, hg, hu = blocks(), blocks(), blocks()
for m in (h, hg, hu):
    m.__dict__.update(topo.hints())
M = sb.Matrix.of
out = torch.empty_like(h.data)


def fused(i):
    sb.matmul_ex(M(x), False, M(w_gu), NT, h, activation="swiglu")


def unfused(i):
    sb.matmul_ex(M(x), False, M(w_g), NT, hg)
    sb.matmul_ex(M(x), False, M(w_u), NT, hu)
    torch.mul(F.silu(hg.data), hu.data, out=out)


fl = 2 * 2.0 * topo.nnz_blocks * 128 * 128 * H
a = bench.time_calls(torch, fused, 3, 10, graph=False)
b = bench.time_calls(torch, unfused, 3, 10, graph=False)
fused(0); unfused(0); torch.cuda.synchronize()
err = float((h.data.float() - out.float()).abs().max() / out.float().abs().max())
print(f"experts {E} {'nt' if NT else 'nn'}: fused swiglu SDD {a * 1e3:8.1f} us ({fl / a / 1e9:6.0f} TF)   2 x SDD (2-SM) + silu*mul "
      f"{b * 1e3:8.1f} us ({fl / b / 1e9:6.0f} TF)   fused/unfused {a / b:.2f}   max diff / max {err:.2e}")
