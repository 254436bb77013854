"""GPU parity tests (run on the B200 box): the CUDA path, reached through the C-ABI, against the
CPU oracle on the reference's own problem lists, against the reference's own kernels
(oracle/_ref, fp16) on the same inputs, and bit-exact topology handling."""
import os

import numpy as np
import pytest

from oracle import oracle
from problems import DDS_PROBLEMS, DSD_PROBLEMS, SDD_PROBLEMS
from sputnik_b200 import synth
from util import (Case, assert_close, dev_block_matrix, errors, load_ref_gpu, run_ours,
                  run_ref_gpu, to_dev)

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "host_matmul.npz")


def _id(p):
    m, k, n, nnz, ta, tb, un = p
    return f"{m}x{k}x{n}-nb{nnz // 16384}-{'t' if ta else 'n'}{'t' if tb else 'n'}{'-unord' if un else ''}"


def _run_list(op, problems, dtype):
    worst = (0.0, 0.0)
    for i, p in enumerate(problems):
        c = Case(op, *p, dtype=dtype, seed=i % 3)  # seeds {0,1,2}
        got = run_ours(c)
        rel, ab = assert_close(got, c.expected(), dtype, f"{op} {_id(p)} {dtype}")
        worst = (max(worst[0], rel), max(worst[1], ab))
    return worst


# The three problem lists of the reference's test-suite, fp16 (the reference's type) and bf16.
@pytest.mark.parametrize("dtype", ["f16", "bf16"])
def test_sdd_reference_problem_list(dtype):
    _run_list("sdd", SDD_PROBLEMS, dtype)


@pytest.mark.parametrize("dtype", ["f16", "bf16"])
def test_dsd_reference_problem_list(dtype):
    _run_list("dsd", DSD_PROBLEMS, dtype)


@pytest.mark.parametrize("dtype", ["f16", "bf16"])
def test_dds_reference_problem_list(dtype):
    _run_list("dds", DDS_PROBLEMS, dtype)


def test_golden_fixtures():
    """Expectations computed by the reference's own host code (oracle/make_golden.py)."""
    z = np.load(GOLDEN)
    for i in range(int(z["n"])):
        s = list(z[f"spec{i}"])
        c = Case(s[0], int(s[1]), int(s[2]), int(s[3]), int(s[4]), s[5] == "True",
                 s[6] == "True", s[7] == "True", s[8], int(s[9]))
        exp = z[f"expected{i}"]
        if c.op == "sdd":
            exp = oracle.sample(exp, c.topo.offsets, c.topo.indices, c.topo.block_rows,
                                c.topo.block_cols)
        assert_close(run_ours(c), exp, c.dtype, f"golden {i} {s}")


@pytest.mark.parametrize("op,problems", [("sdd", SDD_PROBLEMS), ("dsd", DSD_PROBLEMS),
                                         ("dds", DDS_PROBLEMS)])
def test_against_reference_kernels(op, problems):
    """Same inputs through the reference's own CUTLASS kernels (compiled for sm_100 from
    /root/reference into oracle/_ref) and through ours: both within tolerance of the oracle and
    within 2 fp16 ulp-at-scale of each other."""
    lib = load_ref_gpu()
    if lib is None:
        pytest.skip("oracle/_ref/libsputnik_ref.so not built")
    for i, p in enumerate(problems[::3]):
        c = Case(op, *p, dtype="f16", seed=i % 3)
        exp = c.expected()
        ours, ref = run_ours(c), run_ref_gpu(lib, c)
        assert_close(ref, exp, "f16", f"reference kernel {op} {_id(p)}")
        assert_close(ours, exp, "f16", f"ours {op} {_id(p)}")
        scale = max(1.0, float(np.abs(exp).max()))
        assert float(np.abs(ours - ref).max()) <= 2 * scale * 2.0 ** -10, f"{op} {_id(p)}"


@pytest.mark.parametrize("br,bc,nb,unordered,seed", [
    (4, 4, 8, False, 0), (8, 8, 32, True, 1), (64, 256, 2048, False, 2), (40, 3, 100, True, 3),
    (3, 5, 1, False, 4), (1, 1, 1, False, 6), (16, 16, 256, True, 7), (32, 32, 102, False, 8),
    (512, 7168, 57344, False, 9), (300, 2, 600, True, 10), (2, 300, 450, True, 11)])
def test_transpose_metadata_bit_exact(br, bc, nb, unordered, seed):
    """Device-side transpose == the oracle's stable sort == the reference's host Transpose()."""
    import torch
    import sputnik_b200 as sb
    t = synth.moe_topology(64, 8, 112) if nb == 57344 else \
        synth.random_topology(br, bc, nb, seed, unordered)
    a = dev_block_matrix(t, None, "f16", fill=0.0)
    sb.allocate_transpose_buffers(a)
    a.offsets_t.fill_(-1), a.indices_t.fill_(-1), a.block_offsets.fill_(-1)
    sb.transpose(a)
    torch.cuda.synchronize()
    ot, it, bo = oracle.transpose_metadata(t.offsets, t.indices, t.block_rows, t.block_cols)
    assert np.array_equal(a.offsets_t.cpu().numpy(), ot)
    assert np.array_equal(a.indices_t.cpu().numpy(), it)
    assert np.array_equal(a.block_offsets.cpu().numpy(), bo)
    lib = load_ref_gpu()
    if lib is not None:
        import ctypes
        p = lambda x: ctypes.c_void_p(x.data_ptr())
        rot, rit, rbo = (torch.empty_like(a.offsets_t), torch.empty_like(a.indices_t),
                         torch.empty_like(a.block_offsets))
        assert lib.ref_transpose(t.rows, t.cols, t.nonzeros, p(a.offsets), p(a.indices), p(rot),
                                 p(rit), p(rbo), None) == 0
        torch.cuda.synchronize()
        assert torch.equal(rot, a.offsets_t) and torch.equal(rit, a.indices_t)
        assert torch.equal(rbo, a.block_offsets)


@pytest.mark.parametrize("br,bc,nb,seed", [(4, 4, 8, 0), (64, 64, 1000, 1), (5, 3, 0, 2),
                                           (512, 7168, 57344, 3), (100, 1, 37, 4)])
def test_row_indices_bit_exact(br, bc, nb, seed):
    import torch
    import sputnik_b200 as sb
    t = synth.moe_topology(64, 8, 112) if nb == 57344 else synth.random_topology(br, bc, nb, seed)
    a = dev_block_matrix(t, None, "f16", fill=0.0)
    sb.allocate_row_indices_buffer(a)
    sb.row_indices(a)
    torch.cuda.synchronize()
    assert np.array_equal(a.row_indices.cpu().numpy(), oracle.row_indices(t.offsets, br if nb != 57344 else 512))


@pytest.mark.parametrize("op", ["dsd", "dds"])
def test_matmul_ex_equals_matmul(op):
    """MatmulEx (prebuilt metadata, dsd.cu:20-27) and Matmul (builds it) give identical bits."""
    p = (512, 512, 256, 6 * 16384, op == "dsd", False, True)
    c = Case(op, *p, dtype="bf16", seed=1)
    assert np.array_equal(run_ours(c, ex=False), run_ours(c, ex=True))


def test_empty_rows_and_empty_matrix_are_zero_filled():
    import torch
    import sputnik_b200 as sb
    topo = synth.Topology(3, 2, np.array([0, 2, 2, 3], np.int32), np.array([1, 0, 1], np.int16))
    c = Case("dsd", 384, 256, 72, 3 * 16384, False, False, False, "bf16", 0)
    c.topo = topo
    got = run_ours(c)
    assert np.all(got[128:256] == 0)
    assert_close(got, c.expected(), "bf16", "empty row")
    # nnz == 0: the dense output is all zeros
    t0 = synth.Topology(2, 2, np.zeros(3, np.int32), np.zeros(0, np.int16))
    a = dev_block_matrix(t0, np.zeros((0, 128, 128), np.float32), "f16")
    sb.allocate_transpose_buffers(a)
    out = torch.randn(256, 64, device="cuda:0").half()
    sb.matmul(a, False, sb.Matrix.of(to_dev(np.ones((256, 64), np.float32), "f16")), False,
              sb.Matrix.of(out))
    torch.cuda.synchronize()
    assert float(out.abs().max()) == 0.0


@pytest.mark.gpu
@pytest.mark.parametrize("ta,tb", [(0, 0), (0, 1), (1, 0), (1, 1)])
@pytest.mark.parametrize("unordered", [False, True])
def test_sdd_more_blocks_than_sms(ta, tb, unordered):
    """> 148 blocks: tiles are PAIRS of consecutive blocks (N = 256 when they share a block row,
    two 128-wide passes when they do not) handed out by the dynamic scheduler."""
    case = Case("sdd", 6 * 128, 72, 90 * 128, 333 * 16384, ta, tb, unordered, "bf16", seed=17)
    assert case.topo.nnz_blocks == 333
    assert_close(run_ours(case), case.expected(), "bf16", f"sdd {ta}{tb} 333 blocks")


def _skewed_topology(block_rows, block_cols, full_cols, extra, seed):
    """Columns [0, full_cols) hold a block in EVERY row; `extra` more blocks at random positions
    of the remaining columns; per-row column order shuffled."""
    rng = np.random.default_rng(seed)
    rest = block_cols - full_cols
    pick = np.sort(rng.permutation(block_rows * rest)[:extra]) if rest > 0 else np.zeros(0, np.int64)
    er, ec = pick // max(rest, 1), pick % max(rest, 1) + full_cols
    rows = np.concatenate([np.repeat(np.arange(block_rows), full_cols), er])
    cols = np.concatenate([np.tile(np.arange(full_cols), block_rows), ec])
    order = np.lexsort((rng.random(rows.size), rows))      # by row, random order inside a row
    rows, cols = rows[order], cols[order]
    offsets = np.zeros(block_rows + 1, np.int32)
    np.cumsum(np.bincount(rows, minlength=block_rows), out=offsets[1:])
    return synth.Topology(block_rows, block_cols, offsets, cols.astype(np.int16))


@pytest.mark.parametrize("kind", ["rounds", "streaming", "unaligned", "single_cta", "max_blocks"])
def test_transpose_metadata_multi_cta_paths(kind, monkeypatch):
    """The multi-CTA transposed-metadata kernel on the shapes that take its less common paths:
    an owned column range that overflows the shared-memory list (rounds), one column longer than
    the list (streamed), an `indices` pointer that is not 16-byte aligned (scalar loads), the
    largest block count an `int nonzeros` allows — and the one-CTA kernel it replaced
    (BSMM_TRANSPOSE=single), which stays as the fallback."""
    import torch
    import sputnik_b200 as sb
    if kind == "rounds":
        t = _skewed_topology(12000, 300, 3, 500, 1)         # 3 x 12 000 blocks in one CTA's range
    elif kind == "streaming":
        t = _skewed_topology(30000, 2, 1, 9000, 2)          # one column of 30 000 blocks
    elif kind == "max_blocks":
        t = synth.random_topology(362, 362, 131031, 3, True)
    else:
        t = synth.random_topology(96, 80, 2500, 4, True)
    # metadata only: no values needed (Transpose never touches `data`)
    a = sb.BlockMatrix(t.rows, t.cols, 128, t.nonzeros, None,
                       torch.from_numpy(t.offsets).cuda(), torch.from_numpy(t.indices).cuda())
    if kind == "unaligned":
        pad = torch.zeros(t.nnz_blocks + 1, dtype=torch.int16, device="cuda")
        pad[1:] = a.indices
        a.indices = pad[1:]
        assert a.indices.data_ptr() % 16 != 0
    if kind == "single_cta":
        monkeypatch.setenv("BSMM_TRANSPOSE", "single")
    sb.allocate_transpose_buffers(a)
    a.offsets_t.fill_(-1), a.indices_t.fill_(-1), a.block_offsets.fill_(-1)
    sb.transpose(a)
    torch.cuda.synchronize()
    ot, it, bo = oracle.transpose_metadata(t.offsets, t.indices, t.block_rows, t.block_cols)
    assert np.array_equal(a.offsets_t.cpu().numpy(), ot)
    assert np.array_equal(a.block_offsets.cpu().numpy(), bo)
    assert np.array_equal(a.indices_t.cpu().numpy(), it)


@pytest.mark.parametrize("switch", ["BSMM_TAIL_SPLIT=1", "BSMM_WARM=80", "BSMM_EPI=staged", "BSMM_EPI=direct",
                                    "BSMM_PREFETCH=0", "BSMM_RASTER=1"])
def test_experiment_switches_keep_the_numbers(monkeypatch, switch):
    """The opt-in experiment / A-B switches (read on every call) change schedules and epilogues,
    never results: 1-SM DSD with more tiles than SMs (tail-wave half tiles need 2 r <= SMs:
    13 x 12 = 156 tiles -> r = 8), its transposed-output twin (DDS) and a dMoE product (2-SM)."""
    from test_gpu_baseline_sizes import WorkloadCase
    k, v = switch.split("=")
    monkeypatch.setenv(k, v)
    case = Case("dsd", 13 * 128, 1024, 12 * 256, 40 * 16384, False, False, True, "bf16", 9)
    assert_close(run_ours(case, ex=True), case.expected(), "bf16", f"dsd {switch}")
    case = Case("dds", 12 * 256, 1024, 13 * 128, 40 * 16384, False, False, True, "bf16", 10)
    assert_close(run_ours(case, ex=True), case.expected(), "bf16", f"dds {switch}")
    moe = WorkloadCase("m_dsd_2_2_4_256")
    assert_close(run_ours(moe, ex=True, hints=True), moe.expected(), "bf16", f"moe dsd {switch}")
