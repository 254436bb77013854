"""CPU tests of the boundary: the C-ABI library loads, exports every symbol include/bsmm.h
declares, and its validation mirrors the reference's can_launch_* / ValidMatmul rules.
No compute call is made (no GPU here)."""
import ctypes
import os
import re

import pytest

import sputnik_b200 as sb
from sputnik_b200 import _lib
from sputnik_b200._lib import CBlockMatrix, CMatrix, lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FAKE = ctypes.c_void_p(0x10000)  # non-null, 16 B aligned; never dereferenced by *_check


def test_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "bsmm.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(bsmm_[a-z_0-9]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    assert declared == set(_lib.EXPORTS)
    raw = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared:
        assert hasattr(raw, name), f"libbsmm.so does not export {name}"
    assert b"sm_100a" in lib.bsmm_version()


def test_flag_values_match_the_header():
    hdr = open(os.path.join(ROOT, "include", "bsmm.h")).read()
    flags = {k: int(v) for k, v in re.findall(r"#define BSMM_FLAG_([A-Z_]+) (\d+)u", hdr)}
    assert flags.pop("NONE") == 0
    for name, value in flags.items():
        assert getattr(_lib, "FLAG_" + name) == value, name
    assert sorted(flags.values()) == [1 << i for i in range(len(flags))]  # distinct single bits


def _bm(rows, cols, nb, block=128, **kw):
    f = dict(data=FAKE, offsets=FAKE, indices=FAKE, offsets_t=None, indices_t=None,
             block_offsets=None, row_indices=None)
    f.update(kw)
    return CBlockMatrix(rows, cols, nb * 128 * 128, block, f["data"], f["offsets"], f["indices"],
                        f["offsets_t"], f["indices_t"], f["block_offsets"], f["row_indices"])


def _m(rows, cols, data=FAKE):
    return CMatrix(rows, cols, data)


def _r(x):
    return ctypes.byref(x)


def test_sdd_validation():
    a, b = _m(256, 64), _m(64, 384)
    c = _bm(256, 384, 3, row_indices=FAKE)
    assert lib.bsmm_sdd_check(_r(a), 0, _r(b), 0, _r(c), _lib.BSMM_F16) == _lib.BSMM_OK
    assert lib.bsmm_sdd_check(_r(a), 0, _r(b), 0, _r(c), _lib.BSMM_BF16) == _lib.BSMM_OK
    assert lib.bsmm_sdd_check(_r(a), 0, _r(b), 0, _r(c), 7) == _lib.BSMM_ERR_DTYPE
    # block sizes other than 128: cudaErrorNotSupported in the reference (sdd.cu:11-14)
    for bs in (16, 32, 64, 0):
        c2 = _bm(256, 384, 3, block=bs, row_indices=FAKE)
        assert lib.bsmm_sdd_check(_r(a), 0, _r(b), 0, _r(c2), 0) == _lib.BSMM_ERR_BLOCK_SIZE
    # row_indices is required (sdd_mixed_b128_128x128x32x5_nn_align8.cu:75)
    c3 = _bm(256, 384, 3)
    assert lib.bsmm_sdd_check(_r(a), 0, _r(b), 0, _r(c3), 0) == _lib.BSMM_ERR_MISSING_META
    # ValidMatmul (arguments.h:217-231)
    assert lib.bsmm_sdd_check(_r(_m(256, 72)), 0, _r(b), 0, _r(c), 0) == _lib.BSMM_ERR_SHAPE
    assert lib.bsmm_sdd_check(_r(_m(64, 256)), 1, _r(b), 0, _r(c), 0) == _lib.BSMM_OK
    assert lib.bsmm_sdd_check(_r(a), 0, _r(_m(384, 64)), 1, _r(c), 0) == _lib.BSMM_OK
    assert lib.bsmm_sdd_check(_r(a), 1, _r(b), 0, _r(c), 0) == _lib.BSMM_ERR_SHAPE
    # align8 kernels: k % 8
    assert lib.bsmm_sdd_check(_r(_m(256, 60)), 0, _r(_m(60, 384)), 0, _r(c), 0) == \
        _lib.BSMM_ERR_MISALIGNED
    # misaligned pointer
    odd = ctypes.c_void_p(0x10008)
    assert lib.bsmm_sdd_check(_r(_m(256, 64, odd)), 0, _r(b), 0, _r(c), 0) == \
        _lib.BSMM_ERR_MISALIGNED


def test_dsd_validation():
    a, b, c = _bm(256, 384, 3), _m(384, 72), _m(256, 72)
    assert lib.bsmm_dsd_check(_r(a), 0, _r(b), 0, _r(c), 0, 0) == _lib.BSMM_OK
    assert lib.bsmm_dsd_check(_r(a), 0, _r(_m(72, 384)), 1, _r(c), 0, 0) == _lib.BSMM_OK
    # transposed A without the transposed metadata: SPUTNIK_CHECK abort in the reference
    # (dsd_mixed_b128_128x256x32x3_tn_align8.cu:73-75)
    bt, ct = _m(256, 72), _m(384, 72)
    assert lib.bsmm_dsd_check(_r(a), 1, _r(bt), 0, _r(ct), 0, 0) == _lib.BSMM_ERR_MISSING_META
    a2 = _bm(256, 384, 3, offsets_t=FAKE, indices_t=FAKE, block_offsets=FAKE)
    assert lib.bsmm_dsd_check(_r(a2), 1, _r(bt), 0, _r(ct), 0, 1) == _lib.BSMM_OK
    assert lib.bsmm_dsd_check(_r(a2), 1, _r(b), 0, _r(c), 0, 0) == _lib.BSMM_ERR_SHAPE
    assert lib.bsmm_dsd_check(_r(_bm(250, 384, 3)), 0, _r(b), 0, _r(_m(250, 72)), 0, 0) == \
        _lib.BSMM_ERR_MISALIGNED
    assert lib.bsmm_dsd_check(_r(a), 0, _r(_m(384, 70)), 0, _r(_m(256, 70)), 0, 0) == \
        _lib.BSMM_ERR_MISALIGNED
    assert lib.bsmm_dsd_check(None, 0, _r(b), 0, _r(c), 0, 0) == _lib.BSMM_ERR_MISSING_META


def test_dds_validation():
    a, b, c = _m(72, 256), _bm(256, 384, 3), _m(72, 384)
    # NON-transposed B needs the transposed metadata (dds_..._nn_align8.cu:74-84)
    assert lib.bsmm_dds_check(_r(a), 0, _r(b), 0, _r(c), 0, 0) == _lib.BSMM_ERR_MISSING_META
    b2 = _bm(256, 384, 3, offsets_t=FAKE, indices_t=FAKE, block_offsets=FAKE)
    assert lib.bsmm_dds_check(_r(a), 0, _r(b2), 0, _r(c), 0, 0) == _lib.BSMM_OK
    # transposed B walks BCSR rows directly: no transposed metadata needed
    assert lib.bsmm_dds_check(_r(_m(72, 384)), 0, _r(b), 1, _r(_m(72, 256)), 0, 0) == _lib.BSMM_OK
    assert lib.bsmm_dds_check(_r(_m(256, 72)), 1, _r(b2), 0, _r(c), 1, 0) == _lib.BSMM_OK
    assert lib.bsmm_dds_check(_r(a), 0, _r(b2), 0, _r(_m(72, 256)), 0, 0) == _lib.BSMM_ERR_SHAPE


def test_transpose_buffer_bytes_and_strings():
    a = _bm(256, 384, 5)
    o, i, b = ctypes.c_size_t(), ctypes.c_size_t(), ctypes.c_size_t()
    assert lib.bsmm_transpose_buffer_bytes(_r(a), _r(o), _r(i), _r(b)) == 0
    # AllocateTransposeBuffers (arguments.h:233-245)
    assert (o.value, i.value, b.value) == ((384 // 128 + 1) * 4, 5 * 2, 5 * 4)
    for st in range(8):
        assert lib.bsmm_status_string(st)
    assert lib.bsmm_last_cuda_error() == 0
    with pytest.raises(sb.BsmmError):
        _lib.check(_lib.BSMM_ERR_SHAPE)


def test_no_oracle_on_the_product_path():
    """The shipped package must not import, link or execute anything under oracle/."""
    pkg = os.path.join(ROOT, "sputnik_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dp, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, re.M), f
                assert "liboracle" not in text and "oracle/" not in text, f
