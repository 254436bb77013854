"""GPU parity tests of the SSD / SDS / DSS siblings and the device bitmask:
the CUDA path, reached through the C-ABI, against the CPU oracle on the reference's own problem
lists (ssd_test.cu, sds_test.cu, dss_test.cu), against the reference-generated golden fixtures,
and on the edge cases the domain has (empty rows, empty intersections, empty operands)."""
import numpy as np
import pytest

from oracle import oracle
from problems import DSS_PROBLEMS, SDS_PROBLEMS, SSD_PROBLEMS
from sputnik_b200 import synth
from test_siblings import golden_cases
from util import SibCase, assert_close, dev_block_matrix, run_sibling, to_dev

pytestmark = pytest.mark.gpu

LISTS = {"ssd": SSD_PROBLEMS, "sds": SDS_PROBLEMS, "dss": DSS_PROBLEMS}


def _id(p):
    m, k, n, n1, n2, ta, tb, un = p
    return (f"{m}x{k}x{n}-nb{n1 // 16384}/{n2 // 16384}-{'t' if ta else 'n'}{'t' if tb else 'n'}"
            f"{'-unord' if un else ''}")


@pytest.mark.parametrize("dtype", ["f16", "bf16"])
@pytest.mark.parametrize("op", ["ssd", "sds", "dss"])
def test_reference_problem_list(op, dtype):
    for i, p in enumerate(LISTS[op]):
        c = SibCase(op, *p, dtype=dtype, seed=i % 3)
        assert_close(run_sibling(c), c.expected(), dtype, f"{op} {_id(p)} {dtype}")


@pytest.mark.parametrize("op", ["ssd", "sds", "dss"])
def test_matmul_ex_takes_metadata_as_valid(op):
    """MatmulEx (ssd.cu:17-24, sds.cu:17-24, dss.cu:18-27): metadata prebuilt by Transpose()."""
    for i, p in enumerate(LISTS[op][::7]):
        c = SibCase(op, *p, dtype="bf16", seed=i % 3)
        assert_close(run_sibling(c, ex=True), c.expected(), "bf16", f"{op} ex {_id(p)}")


def test_golden_fixtures():
    """Expectations computed by the reference's own host code (oracle/make_golden_siblings.py)."""
    for spec, expected in golden_cases():
        c = SibCase(*spec)
        exp = expected
        if c.topo_c is not None:
            t = c.topo_c
            exp = oracle.sample(expected, t.offsets, t.indices, t.block_rows, t.block_cols)
        assert_close(run_sibling(c), exp, c.dtype, f"golden {spec}")


@pytest.mark.parametrize("op,shape", [
    ("ssd", (2048, 1024, 2048, 40, 200)),   # > 148 output blocks: dynamic scheduler, pair tiles
    ("sds", (2048, 1024, 2048, 40, 200)),
    ("dss", (2048, 2048, 2048, 90, 110)),   # 256 output tiles, sparse intersections
    ("dss", (512, 8192, 512, 100, 100)),    # long walks (> 32 entries per chunk), 64 K blocks
    ("ssd", (512, 8192, 512, 150, 16)),
])
@pytest.mark.parametrize("ta,tb", [(False, False), (True, True), (True, False), (False, True)])
def test_larger_problems(op, shape, ta, tb):
    m, k, n, nb1, nb2 = shape
    c = SibCase(op, m, k, n, nb1 * 16384, nb2 * 16384, ta, tb, True, "bf16", 1)
    assert_close(run_sibling(c), c.expected(), "bf16", f"{op} {shape} {ta} {tb}")


def test_dss_disjoint_topologies_give_zeros():
    """Every K block of A's rows misses B's columns: the whole (pre-filled) output is zeros."""
    import torch
    import sputnik_b200 as sb
    ta_ = synth.Topology(2, 4, np.array([0, 2, 4], np.int32), np.array([0, 1, 1, 0], np.int16))
    tb_ = synth.Topology(4, 2, np.array([0, 0, 0, 2, 4], np.int32), np.array([0, 1, 1, 0], np.int16))
    va = synth.uniform_values((4, 128, 128), 1)
    vb = synth.uniform_values((4, 128, 128), 2)
    a, b = dev_block_matrix(ta_, va, "bf16"), dev_block_matrix(tb_, vb, "bf16")
    sb.allocate_transpose_buffers(b)
    out = torch.randn(256, 256, device="cuda:0").bfloat16()
    sb.matmul(a, False, b, False, sb.Matrix.of(out))
    torch.cuda.synchronize()
    assert not out.any().item()


@pytest.mark.parametrize("op", ["ssd", "sds", "dss"])
def test_empty_sparse_input_gives_zeros(op):
    c = SibCase(op, 256, 256, 256, 0, 2 * 16384, False, True, False, "f16", 0)
    got = run_sibling(c)
    assert got.shape == c.expected().shape and not got.any()


@pytest.mark.parametrize("br,bc,nb,unordered", [(3, 70, 40, True), (64, 64, 1000, False),
                                                (5, 130, 0, False), (300, 9, 700, True)])
def test_bitmask_bit_exact(br, bc, nb, unordered):
    """bsmm_bitmask vs the restatement of Bitmask() + BitMatrix (bitmask.cu:7-45): both views."""
    import torch
    import sputnik_b200 as sb
    t = synth.random_topology(br, bc, nb, 5, unordered)
    m = dev_block_matrix(t, None, "f16", fill=0.0)
    got = sb.bitmask(m, sb.allocate_bitmask_buffers(m))
    torch.cuda.synchronize()
    exp = oracle.bitmask(t.offsets, t.indices, br, bc)
    assert np.array_equal(got.cpu().numpy().view(np.uint64), exp)
    if nb == 0:
        return  # zero-length index tensors have no device address to hand to Transpose()
    # with the transposed buffers present the reference walks the transposed view (bitmask.cu:8)
    sb.allocate_transpose_buffers(m)
    sb.transpose(m)
    got = sb.bitmask(m, sb.allocate_bitmask_buffers(m))
    torch.cuda.synchronize()
    ot, it, _ = oracle.transpose_metadata(t.offsets, t.indices, br, bc)
    assert np.array_equal(got.cpu().numpy().view(np.uint64), oracle.bitmask(ot, it, bc, br))
