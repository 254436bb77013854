"""bsmm_dsd_scatter: the DSD epilogue writes output row i at a caller-given address (the fused
un-permute / un-pad / return exchange of a dMoE layer).  Against the plain DSD on the same
operands: permuted rows bit-identical, dropped rows and foreign rows of the destination
untouched — 1-SM and 2-SM kernels, every transpose pair, both dtypes, ragged widths."""
import numpy as np
import pytest
import torch

import sputnik_b200 as sb
from oracle import oracle
from sputnik_b200 import synth
from util import dev_block_matrix, to_dev, torch_dtype

pytestmark = pytest.mark.gpu
DEV = "cuda:0"

TOPOS = {"random": lambda: synth.random_topology(5, 6, 13, seed=3, unordered=True),   # 1-SM kernel
         "moe": lambda: synth.moe_topology(3, 4, 2),                                  # 2-SM kernel
         "empty_rows": lambda: synth.Topology(4, 3, np.array([0, 0, 2, 2, 3], dtype=np.int32),
                                              np.array([0, 2, 1], dtype=np.int16))}


def _problem(topo, ta, tb, n, dtype, seed):
    r = lambda shape, s: oracle.round_to(synth.uniform_values(shape, seed * 131 + s), dtype)
    sp = dev_block_matrix(topo, r((topo.nnz_blocks, 128, 128), 1), dtype)
    sb.allocate_transpose_buffers(sp)
    sb.transpose(sp)
    sp.twin_rows, sp.twin_cols = topo.twin_rows(), topo.twin_cols()
    kdim, m = (topo.rows, topo.cols) if ta else (topo.cols, topo.rows)
    b = to_dev(r((n, kdim) if tb else (kdim, n), 2), dtype)
    return sp, b, m


@pytest.mark.parametrize("dtype", ["bf16", "f16"])
@pytest.mark.parametrize("ta,tb", [(0, 0), (0, 1), (1, 0), (1, 1)])
@pytest.mark.parametrize("kind", list(TOPOS))
def test_scatter_equals_dsd_then_permute(kind, ta, tb, dtype):
    topo = TOPOS[kind]()
    n = 264
    sp, b, m = _problem(topo, ta, tb, n, dtype, seed=5)
    tdt = torch_dtype(dtype)
    y = torch.empty((m, n), device=DEV, dtype=tdt)
    sb.matmul_ex(sp, ta, sb.Matrix.of(b), tb, sb.Matrix.of(y))
    # destination: 1.5 m rows; a random injective map, every 7th source row dropped
    g = torch.Generator().manual_seed(11)
    dest_rows = m + m // 2
    perm = torch.randperm(dest_rows, generator=g)[:m]
    perm[::7] = -1
    dest = torch.full((dest_rows, n), 3.0, device=DEV, dtype=tdt)
    ptrs = sb.row_pointers(dest, perm.to(DEV))
    sb.dsd_scatter(sp, ta, sb.Matrix.of(b), tb, ptrs, n)
    torch.cuda.synchronize()
    keep = perm >= 0
    assert torch.equal(dest[perm[keep].to(DEV)], y[keep.to(DEV)])
    untouched = torch.ones(dest_rows, dtype=torch.bool)
    untouched[perm[keep]] = False
    assert bool((dest[untouched.to(DEV)] == 3.0).all())


def test_scatter_wide_identity_map():
    """n = 1032 (four full 256-wide tiles + 8 columns), identity row map == plain DSD."""
    topo = synth.moe_topology(2, 4, 4)
    sp, b, m = _problem(topo, 0, 0, 1032, "bf16", seed=9)
    y = torch.empty((m, 1032), device=DEV, dtype=torch.bfloat16)
    sb.matmul_ex(sp, False, sb.Matrix.of(b), False, sb.Matrix.of(y))
    dest = torch.full_like(y, float("nan"))
    sb.dsd_scatter(sp, False, sb.Matrix.of(b), False, sb.row_pointers(dest), 1032)
    torch.cuda.synchronize()
    assert torch.equal(dest, y)


def test_scatter_without_blocks_writes_zeros():
    topo = synth.Topology(2, 2, np.zeros(3, dtype=np.int32), np.zeros(0, dtype=np.int16))
    vals = torch.empty((0, 128, 128), device=DEV, dtype=torch.bfloat16)
    sp = sb.BlockMatrix(256, 256, 128, 0, vals, torch.zeros(3, dtype=torch.int32, device=DEV),
                        torch.zeros(0, dtype=torch.int16, device=DEV))
    b = torch.ones((256, 64), device=DEV, dtype=torch.bfloat16)
    dest = torch.full((300, 64), 5.0, device=DEV, dtype=torch.bfloat16)
    idx = torch.arange(256, device=DEV) + 10
    idx[3] = -1
    sb.dsd_scatter(sp, False, sb.Matrix.of(b), False, sb.row_pointers(dest, idx), 64,
                   dtype=torch.bfloat16)
    torch.cuda.synchronize()
    assert not dest[10:13].any() and not dest[14:266].any()
    assert bool((dest[:10] == 5.0).all()) and bool((dest[266:] == 5.0).all())
    assert bool((dest[13] == 5.0).all())


def test_scatter_rejects_bad_arguments():
    topo = synth.moe_topology(2, 2, 2)
    sp, b, m = _problem(topo, 0, 0, 256, "bf16", seed=1)
    dest = torch.empty((m, 256), device=DEV, dtype=torch.bfloat16)
    with pytest.raises(sb.BsmmError):      # wrong number of rows for this product
        sb.dsd_scatter(sp, False, sb.Matrix.of(b), False, sb.row_pointers(dest)[:-128].contiguous(), 256)
    with pytest.raises(sb.BsmmError):      # output width does not match the dense operand
        sb.dsd_scatter(sp, False, sb.Matrix.of(b), False, sb.row_pointers(dest), 128)


@pytest.mark.parametrize("cols", [8, 136, 4096])
def test_gather_rows_permutes_pads_and_zero_fills(cols):
    """bsmm_gather_rows against torch indexing: random source rows, every 5th row null."""
    g = torch.Generator().manual_seed(cols)
    src = torch.empty((700, cols), device=DEV).uniform_(-1, 1).bfloat16()
    idx = torch.randint(0, 700, (517,), generator=g)
    idx[::5] = -1
    dst = torch.full((517, cols), 9.0, device=DEV, dtype=torch.bfloat16)
    sb.gather_rows(sb.row_pointers(src, idx.to(DEV)), dst)
    torch.cuda.synchronize()
    want = src[idx.clamp(min=0).to(DEV)]
    want[(idx < 0).to(DEV)] = 0
    assert torch.equal(dst, want)


def test_gather_and_scatter_stay_inside_their_rows():
    """Guard columns / rows around the destinations (no compute-sanitizer on the GPU pool): the
    gather writes exactly rows x cols elements, the scatter exactly the addressed rows."""
    G = 64
    src = torch.empty((300, 136), device=DEV).uniform_(-1, 1).bfloat16()
    base = torch.full((G + 200 * 136 + G,), 9.0, device=DEV, dtype=torch.bfloat16)
    dst = base[G:G + 200 * 136].view(200, 136)
    idx = torch.arange(200, device=DEV) + 50
    sb.gather_rows(sb.row_pointers(src, idx), dst)
    torch.cuda.synchronize()
    assert torch.equal(dst, src[50:250])
    assert bool((base[:G] == 9.0).all()) and bool((base[-G:] == 9.0).all())
    # scatter into every second row of a wider, guarded destination
    topo = synth.moe_topology(2, 2, 2)
    sp, b, m = _problem(topo, 0, 0, 264, "bf16", seed=2)
    y = torch.empty((m, 264), device=DEV, dtype=torch.bfloat16)
    sb.matmul_ex(sp, False, sb.Matrix.of(b), False, sb.Matrix.of(y))
    wide = torch.full((2 * m + 2, 264), 4.0, device=DEV, dtype=torch.bfloat16)
    sb.dsd_scatter(sp, False, sb.Matrix.of(b), False,
                   sb.row_pointers(wide, 2 * torch.arange(m, device=DEV) + 1), 264)
    torch.cuda.synchronize()
    assert torch.equal(wide[1:2 * m:2], y)
    assert bool((wide[0:2 * m + 2:2] == 4.0).all())
