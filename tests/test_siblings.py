"""CPU tests of the SSD / SDS / DSS siblings and the bitmask: the oracle
against fixtures produced by the reference's own host code (oracle/make_golden_siblings.py),
the bitmask restatement against a plain-Python statement of BitMatrix::Set, and the C-ABI
validation (no compute call: no GPU here)."""
import ctypes
import os

import numpy as np
import pytest

from oracle import oracle
from problems import DSS_PROBLEMS, SDS_PROBLEMS, SSD_PROBLEMS
from sputnik_b200 import _lib, synth
from sputnik_b200._lib import CBlockMatrix, CMatrix, lib
from util import SibCase

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "host_matmul_siblings.npz")
FAKE = ctypes.c_void_p(0x10000)


def golden_cases():
    z = np.load(GOLDEN)
    for i in range(int(z["n"])):
        s = list(z[f"spec{i}"])
        spec = (s[0], int(s[1]), int(s[2]), int(s[3]), int(s[4]), int(s[5]), s[6] == "True",
                s[7] == "True", s[8] == "True", s[9], int(s[10]))
        yield spec, z[f"expected{i}"]


@pytest.mark.parametrize("spec,expected", list(golden_cases()),
                         ids=lambda v: "-".join(map(str, v)) if isinstance(v, tuple) else "exp")
def test_oracle_matches_reference_golden(spec, expected):
    """Fixtures come from the reference's Matrix::operator* / T / Mul on densified operands;
    the restatement (which skips zero blocks) must reproduce them."""
    c = SibCase(*spec)
    got = c.expected()
    if c.topo_c is not None:
        t = c.topo_c
        exp = oracle.sample(expected, t.offsets, t.indices, t.block_rows, t.block_cols)
        # nothing outside the output topology survives the reference's mask
        assert np.array_equal(oracle.densify(exp, t.offsets, t.indices, t.block_rows, t.block_cols),
                              expected)
    else:
        exp = expected
    assert np.array_equal(got, exp)  # numeric equality (-0.0 == 0.0 allowed)


def test_problem_lists_are_the_reference_ones():
    # ssd_test.cu / sds_test.cu: 80 typed problems each; dss_test.cu: 92
    assert (len(SSD_PROBLEMS), len(SDS_PROBLEMS), len(DSS_PROBLEMS)) == (80, 80, 92)
    for m, k, n, n1, n2, ta, tb, un in SSD_PROBLEMS + SDS_PROBLEMS + DSS_PROBLEMS:
        assert m % 128 == 0 and k % 128 == 0 and n % 128 == 0
        assert n1 % 16384 == 0 and n2 % 16384 == 0


@pytest.mark.parametrize("br,bc,nb,unordered", [(1, 1, 1, False), (3, 70, 40, True),
                                                (5, 64, 64, False), (4, 130, 0, False)])
def test_bitmask_oracle(br, bc, nb, unordered):
    t = synth.random_topology(br, bc, nb, 3, unordered)
    got = oracle.bitmask(t.offsets, t.indices, br, bc)
    words = (bc + 63) // 64
    exp = [0] * (br * words)
    for i in range(br):  # BitMatrix::Set (bit_matrix.h:38-43)
        for o in range(int(t.offsets[i]), int(t.offsets[i + 1])):
            j = int(t.indices[o])
            exp[i * words + j // 64] |= 1 << (j % 64)
    assert [int(x) for x in got] == exp
    # the transposed view is the bitmask of the transposed topology
    ot, it, _ = oracle.transpose_metadata(t.offsets, t.indices, br, bc)
    gt = oracle.bitmask(ot, it, bc, br)
    wt = (br + 63) // 64
    for i in range(br):
        for j in range(bc):
            a = (int(got[i * words + j // 64]) >> (j % 64)) & 1
            b = (int(gt[j * wt + i // 64]) >> (i % 64)) & 1
            assert a == b


def _bm(rows, cols, nb, block=128, meta_t=False, row_indices=None):
    t = FAKE if meta_t else None
    return CBlockMatrix(rows, cols, nb * 16384, block, FAKE, FAKE, FAKE, t, t, t, row_indices)


def _r(x):
    return ctypes.byref(x)


def test_ssd_validation():
    a, b = _bm(256, 384, 3), CMatrix(384, 256, FAKE)
    c = _bm(256, 256, 2, row_indices=FAKE)
    assert lib.bsmm_ssd_check(_r(a), 0, _r(b), 0, _r(c), 0, 0) == _lib.BSMM_OK
    assert lib.bsmm_ssd_check(_r(a), 0, _r(CMatrix(256, 384, FAKE)), 1, _r(c), 1, 0) == _lib.BSMM_OK
    # sparse output needs row_indices (ssd_mixed_b128_128x128x32x5_tn_align8.cu:88)
    assert lib.bsmm_ssd_check(_r(a), 0, _r(b), 0, _r(_bm(256, 256, 2)), 0, 0) == \
        _lib.BSMM_ERR_MISSING_META
    # transposed A needs the transposed metadata (:75-77)
    at = _bm(384, 256, 3)
    assert lib.bsmm_ssd_check(_r(at), 1, _r(b), 0, _r(c), 0, 0) == _lib.BSMM_ERR_MISSING_META
    at = _bm(384, 256, 3, meta_t=True)
    assert lib.bsmm_ssd_check(_r(at), 1, _r(b), 0, _r(c), 0, 0) == _lib.BSMM_OK
    # block size (ssd.cu:11-14) and shapes
    c16 = _bm(256, 256, 2, block=16, row_indices=FAKE)
    assert lib.bsmm_ssd_check(_r(a), 0, _r(b), 0, _r(c16), 0, 0) == _lib.BSMM_ERR_BLOCK_SIZE
    assert lib.bsmm_ssd_check(_r(a), 0, _r(CMatrix(256, 256, FAKE)), 0, _r(c), 0, 0) == \
        _lib.BSMM_ERR_SHAPE


def test_sds_validation():
    a, c = CMatrix(256, 384, FAKE), _bm(256, 256, 2, row_indices=FAKE)
    b = _bm(384, 256, 3)
    # NON-transposed B is walked by block column: transposed metadata required
    assert lib.bsmm_sds_check(_r(a), 0, _r(b), 0, _r(c), 0, 0) == _lib.BSMM_ERR_MISSING_META
    b2 = _bm(384, 256, 3, meta_t=True)
    assert lib.bsmm_sds_check(_r(a), 0, _r(b2), 0, _r(c), 0, 0) == _lib.BSMM_OK
    assert lib.bsmm_sds_check(_r(a), 0, _r(_bm(256, 384, 3)), 1, _r(c), 0, 0) == _lib.BSMM_OK
    assert lib.bsmm_sds_check(_r(a), 0, _r(b2), 0, _r(_bm(256, 256, 2)), 0, 0) == \
        _lib.BSMM_ERR_MISSING_META
    assert lib.bsmm_sds_check(_r(a), 1, _r(b2), 0, _r(c), 0, 0) == _lib.BSMM_ERR_SHAPE


def test_dss_validation():
    a, c = _bm(256, 384, 3), CMatrix(256, 256, FAKE)
    b = _bm(384, 256, 3)
    assert lib.bsmm_dss_check(_r(a), 0, _r(b), 0, _r(c), 0, 0) == _lib.BSMM_ERR_MISSING_META
    b2 = _bm(384, 256, 3, meta_t=True)
    assert lib.bsmm_dss_check(_r(a), 0, _r(b2), 0, _r(c), 1, 0) == _lib.BSMM_OK
    assert lib.bsmm_dss_check(_r(a), 0, _r(_bm(256, 384, 3)), 1, _r(c), 0, 0) == _lib.BSMM_OK
    # dss_..._tn_align8.cu:76-83: both column views
    assert lib.bsmm_dss_check(_r(_bm(384, 256, 3)), 1, _r(b2), 0, _r(c), 0, 0) == \
        _lib.BSMM_ERR_MISSING_META
    assert lib.bsmm_dss_check(_r(_bm(384, 256, 3, meta_t=True)), 1, _r(b2), 0, _r(c), 0, 0) == \
        _lib.BSMM_OK
    assert lib.bsmm_dss_check(_r(a), 0, _r(b2), 0, _r(CMatrix(256, 384, FAKE)), 0, 0) == \
        _lib.BSMM_ERR_SHAPE
    # the K limit of the position table (the reference's is 32768, index_merge.h:103)
    kbig = 128 * 2049
    big_a, big_b = _bm(128, kbig, 1), _bm(128, kbig, 1)
    assert lib.bsmm_dss_check(_r(big_a), 0, _r(big_b), 1, _r(CMatrix(128, 128, FAKE)), 0, 0) == \
        _lib.BSMM_ERR_RANGE


def test_bitmask_bytes():
    n = ctypes.c_size_t()
    # BitMatrix::SizeInBytes (bit_matrix.h:16-19), view chosen as in bitmask.h:16-23
    assert lib.bsmm_bitmask_bytes(_r(_bm(256, 128 * 70, 3)), _r(n)) == 0
    assert n.value == 2 * 2 * 8
    assert lib.bsmm_bitmask_bytes(_r(_bm(256, 128 * 70, 3, meta_t=True)), _r(n)) == 0
    assert n.value == 70 * 1 * 8
    assert lib.bsmm_bitmask_bytes(_r(_bm(256, 256, 1, block=64)), _r(n)) == _lib.BSMM_ERR_BLOCK_SIZE
