"""GPU checks at sizes the CPU oracle cannot reach, through size-independent properties
(SURVEY.md 8c): block-wise agreement with a dense fp32 product of the same 16-bit operands,
run-to-run determinism, and addressing beyond the reference's 32-bit byte-offset limit
(> 65535 blocks = 2 GiB of values, block_tile_access_iterator.h:223-229)."""
import numpy as np
import pytest
import torch

import sputnik_b200 as sb
from sputnik_b200 import harness, synth, workloads

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _rel(got, ref):
    ref = ref.double()
    tau = 1e-2 * ref.pow(2).mean().sqrt()
    return float(((got.double() - ref).abs() / torch.maximum(ref.abs(), tau)).max())


def test_more_than_65535_blocks():
    br, bc, k = 264, 250, 8                       # 66 000 blocks: 2.16 GB of bf16 values
    topo = synth.random_topology(br, bc, br * bc, seed=0)
    g = torch.Generator(device=DEV).manual_seed(1)
    a = torch.empty((br * 128, k), device=DEV).uniform_(-1, 1, generator=g).bfloat16()
    b = torch.empty((k, bc * 128), device=DEV).uniform_(-1, 1, generator=g).bfloat16()
    vals = torch.full((topo.nnz_blocks, 128, 128), float("nan"), dtype=torch.bfloat16, device=DEV)
    c = sb.BlockMatrix(topo.rows, topo.cols, 128, topo.nonzeros, vals,
                       torch.from_numpy(topo.offsets).to(DEV), torch.from_numpy(topo.indices).to(DEV))
    sb.allocate_row_indices_buffer(c)
    sb.row_indices(c)
    sb.matmul(sb.Matrix.of(a), False, sb.Matrix.of(b), False, c)          # SDD
    torch.cuda.synchronize()
    for blk in (0, 1, 65535, 65536, 65537, topo.nnz_blocks - 1, 40000, 65999):
        r, cc = blk // bc, blk % bc
        ref = a[r * 128:(r + 1) * 128].float() @ b[:, cc * 128:(cc + 1) * 128].float()
        assert _rel(vals[blk].float(), ref) <= 1e-2, blk
    assert bool(torch.isfinite(vals.view(-1)[::4099].float()).all())
    # the same > 65535-block matrix as the sparse operand of a DSD and a transposed DSD
    sb.allocate_transpose_buffers(c)
    x = torch.empty((bc * 128, 16), device=DEV).uniform_(-1, 1, generator=g).bfloat16()
    y = torch.full((br * 128, 16), float("nan"), dtype=torch.bfloat16, device=DEV)
    sb.matmul(c, False, sb.Matrix.of(x), False, sb.Matrix.of(y))
    xt = torch.empty((br * 128, 16), device=DEV).uniform_(-1, 1, generator=g).bfloat16()
    yt = torch.full((bc * 128, 16), float("nan"), dtype=torch.bfloat16, device=DEV)
    sb.matmul(c, True, sb.Matrix.of(xt), False, sb.Matrix.of(yt))           # builds metadata
    torch.cuda.synchronize()
    for r in (0, 131, br - 1):                      # block rows incl. blocks beyond 65535
        dense_rows = vals[r * bc:(r + 1) * bc].float().permute(1, 0, 2).reshape(128, bc * 128)
        assert _rel(y[r * 128:(r + 1) * 128].float(), dense_rows @ x.float()) <= 1e-2, r
    for col in (0, 249):
        blocks = vals[col::bc].float()              # [br, 128, 128]: block (r, col)
        dense_cols = blocks.reshape(br * 128, 128)  # rows of the column panel
        assert _rel(yt[col * 128:(col + 1) * 128].float(), dense_cols.t() @ xt.float()) <= 1e-2, col


@pytest.mark.parametrize("name", ["c2", "c5_sdd_e8", "c5_dsd_e8", "c5_dds_tn_e8", "c3a", "c3b", "c4_dsd_50",
                                  "c4_sdd_50", "c4_dds_50", "c1"])
def test_full_size_blockwise_agreement_and_determinism(name):
    wl = workloads.get(name)
    prob = harness.DeviceProblem(wl, DEV, sets=1, seed=5)
    s = prob.sets[0]
    prob.call(s)
    first = s.output().clone()
    prob.call(s)
    torch.cuda.synchronize()
    assert torch.equal(first, s.output()), "not deterministic"
    t = wl.topo
    vals = s.sparse.data
    rows = torch.from_numpy(t.row_indices().astype(np.int64))
    cols = torch.from_numpy(t.indices.astype(np.int64))

    def densify(r0, r1, c0, c1):
        """fp32 dense sub-matrix of the sparse operand: block rows [r0,r1) x block cols [c0,c1)."""
        out = torch.zeros(((r1 - r0) * 128, (c1 - c0) * 128), device=DEV)
        sel = ((rows >= r0) & (rows < r1) & (cols >= c0) & (cols < c1)).nonzero().flatten().tolist()
        for j in sel:
            rr, cc = int(rows[j]) - r0, int(cols[j]) - c0
            out[rr * 128:(rr + 1) * 128, cc * 128:(cc + 1) * 128] = vals[j].float()
        return out

    opa = (lambda m: m.data.float().t()) if wl.ta else (lambda m: m.data.float())
    opb = (lambda m: m.data.float().t()) if wl.tb else (lambda m: m.data.float())
    if wl.op == "sdd":
        A, B = opa(s.a), opb(s.b)
        for j in (0, t.nnz_blocks // 2 + 1, t.nnz_blocks - 1):
            r, c = int(rows[j]), int(cols[j])
            ref = A[r * 128:(r + 1) * 128] @ B[:, c * 128:(c + 1) * 128]
            assert _rel(vals[j].float(), ref) <= 1e-2, j
    elif wl.op == "dsd":
        B = opb(s.b)
        nbr = t.block_cols if wl.ta else t.block_rows
        for r in (0, nbr // 2, nbr - 1):            # output block rows
            if wl.ta:
                sp = densify(0, t.block_rows, r, r + 1).t()
            else:
                sp = densify(r, r + 1, 0, t.block_cols)
            assert _rel(s.c.data[r * 128:(r + 1) * 128].float(), sp @ B) <= 1e-2, r
    else:
        A = opa(s.a)
        nbc = t.block_rows if wl.tb else t.block_cols
        for c in (0, nbc // 2, nbc - 1):            # output block columns
            if wl.tb:
                sp = densify(c, c + 1, 0, t.block_cols).t()
            else:
                sp = densify(0, t.block_rows, c, c + 1)
            assert _rel(s.c.data[:, c * 128:(c + 1) * 128].float(), A @ sp) <= 1e-2, c
