"""Experiment: does the 2.2x DRAM over-read of h = sdd(x, w1) at 64 experts (w1 row pitch
7 * 2^18 B) come from power-of-two-pitch aliasing?  Times SDD nn and DSD nt (the two products that
walk w1 [hidden, E*F] by column panels) with `pad` extra EMPTY block columns appended to the
topology / w1 (pitch + 256 B per pad).  python tools/pitch_probe.py [experts]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402,F401
import torch  # noqa: E402
import sputnik_b200 as sb  # noqa: E402
from sputnik_b200 import synth  # noqa: E402
import bench  # noqa: E402

E = int(sys.argv[1]) if len(sys.argv) > 1 else 64
H, F, R = 4096, 14336, 8
dev = "cuda:0"
for pad in (0, 1, 2, 5):
    t0 = synth.moe_topology(E, R, F // 128)
    t = synth.Topology(t0.block_rows, t0.block_cols + pad, t0.offsets, t0.indices)
    g = torch.Generator(device=dev).manual_seed(0)
    x = torch.empty((t.rows, H), device=dev).uniform_(-1, 1, generator=g).bfloat16()
    w1 = torch.empty((H, t.cols), device=dev, dtype=torch.bfloat16)
    w1.uniform_(-0.02, 0.02, generator=g)
    vals = torch.empty((t.nnz_blocks, 128, 128), dtype=torch.bfloat16, device=dev)
    h = sb.BlockMatrix(t.rows, t.cols, 128, t.nonzeros, vals, torch.from_numpy(t.offsets).to(dev),
                       torch.from_numpy(t.indices).to(dev))
    sb.allocate_row_indices_buffer(h)
    sb.row_indices(h)
    sb.allocate_transpose_buffers(h)
    sb.transpose(h)
    h.__dict__.update(t0.hints())
    dx = torch.empty_like(x)
    M = sb.Matrix.of
    ms_sdd = bench.time_calls(torch, lambda i: sb.matmul_ex(M(x), False, M(w1), False, h), 3, 10, graph=False)
    ms_dsd = bench.time_calls(torch, lambda i: sb.matmul_ex(h, False, M(w1), True, M(dx)), 3, 10, graph=False)
    fl = 2.0 * t.nnz_blocks * 128 * 128 * H
    print(f"experts {E} pad {pad}: w1 pitch {t.cols * 2} B  sdd nn {ms_sdd * 1e3:8.1f} us "
          f"({fl / ms_sdd / 1e9:6.0f} TF)   dsd nt {ms_dsd * 1e3:8.1f} us ({fl / ms_dsd / 1e9:6.0f} TF)", flush=True)
    del x, w1, vals, h, dx
    torch.cuda.empty_cache()
