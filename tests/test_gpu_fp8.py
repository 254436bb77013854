"""FP8 (e4m3) operands: BSMM_F8E4M3 products — e4m3 x e4m3, fp32 accumulate on
tcgen05.mma.kind::f8f6f4, bf16 output (no reference counterpart: the reference
instantiates cutlass::half_t only).

Expectation: the oracle's host matmul (sputnik/matrix_utils.h:376-391: float product, double
accumulate) on the e4m3-ROUNDED operands (oracle.round_e4m3, pinned against torch.float8_e4m3fn in
tests/test_oracle.py).  Products of two e4m3 values are exact in fp32, so the only differences are
the fp32 summation order and the one bf16 rounding of the output: same tolerance as the 16-bit
types (max relative error <= 1e-2 with the tau guard, tests/util.assert_close).

Problems: the reference's three lists; their dense extents (8, 24, 40, ... — multiples of 8, the
`align8` kernels) are rounded up to multiples of 16, which 1-byte operands need (16 B rows)."""
import numpy as np
import pytest
import torch

import sputnik_b200 as sb
from problems import DDS_PROBLEMS, DSD_PROBLEMS, SDD_PROBLEMS
from sputnik_b200 import _lib
from util import Case, assert_close, dev_block_matrix, run_ours, to_dev

pytestmark = pytest.mark.gpu


def _up16(x):
    return (x + 15) // 16 * 16


def _fp8_problem(op, p):
    m, k, n, nnz, ta, tb, un = p
    if op == "sdd":
        k = _up16(k)
    elif op == "dsd":
        n = _up16(n)
    else:
        m = _up16(m)
    return (m, k, n, nnz, ta, tb, un)


@pytest.mark.parametrize("op,problems", [("sdd", SDD_PROBLEMS), ("dsd", DSD_PROBLEMS),
                                         ("dds", DDS_PROBLEMS)])
@pytest.mark.parametrize("hints", [False, True])
def test_reference_problem_lists_fp8(op, problems, hints):
    """Every transpose pair, ordered and unordered indices, ragged dense extents; `hints`: pass the
    topology's twin / even hints so the 2-SM kernels run where the topology allows."""
    for i, p in enumerate(problems):
        c = Case(op, *_fp8_problem(op, p), dtype="f8", seed=i % 3)
        if hints and not any(c.topo.hints().values()):
            continue
        got = run_ours(c, hints=hints)
        assert_close(got, c.expected(), "bf16", f"{op} fp8 {p} hints={hints}")


@pytest.mark.parametrize("name", ["m_dsd_4_2_6_272", "m_dsdt_4_2_6_272", "m_sdd_4_2_6_272"])
def test_fp8_twin_topologies_pair_kernels(name):
    """dMoE-shaped topologies (twin rows / columns): 2-SM kernels at fp8 == 1-SM kernels, bit for
    bit (same products, same K order), and both match the oracle."""
    from test_gpu_baseline_sizes import WorkloadCase
    c = WorkloadCase(name)
    c.dtype = "f8"
    from oracle import oracle
    for attr in ("values", "a", "b"):
        v = getattr(c, attr)
        if v is not None:
            setattr(c, attr, oracle.round_to(v, "f8"))
    ref = c.expected()
    one = run_ours(c, ex=True, hints=False)
    two = run_ours(c, ex=True, hints=True)
    assert_close(one, ref, "bf16", f"{name} 1-SM fp8")
    assert np.array_equal(one, two), f"{name}: 2-SM fp8 kernel differs from the 1-SM one"


def test_fp8_validation():
    """Mixed operand types and misaligned fp8 extents are refused, not reinterpreted."""
    c = Case("dsd", 256, 256, 64, 2 * 16384, False, False, False, "f8", 1)
    sp = dev_block_matrix(c.topo, c.values, "f8")
    b8 = sb.Matrix.of(to_dev(c.b, "f8"))
    b16 = sb.Matrix.of(to_dev(c.b, "bf16"))
    out = sb.Matrix.of(torch.zeros(256, 64, dtype=torch.bfloat16, device="cuda"))
    with pytest.raises(sb.BsmmError) as ei:
        sb.matmul(sp, False, b16, False, out)              # e4m3 x bf16
    assert ei.value.status == _lib.BSMM_ERR_DTYPE
    out16 = sb.Matrix.of(torch.zeros(256, 64, dtype=torch.float16, device="cuda"))
    with pytest.raises(sb.BsmmError):
        sb.matmul(sp, False, b8, False, out16)             # fp8 products store bf16
    b_odd = sb.Matrix.of(torch.zeros(256, 24, dtype=torch.float8_e4m3fn, device="cuda"))
    out_odd = sb.Matrix.of(torch.zeros(256, 24, dtype=torch.bfloat16, device="cuda"))
    with pytest.raises(sb.BsmmError) as ei:
        sb.matmul(sp, False, b_odd, False, out_odd)        # 24 bytes per row
    assert ei.value.status == _lib.BSMM_ERR_MISALIGNED
    sb.matmul(sp, False, b8, False, out)
    torch.cuda.synchronize()
    assert_close(out.data.float().cpu().numpy(), c.expected(), "bf16", "fp8 dsd")
