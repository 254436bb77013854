"""ctypes front of oracle/bsmm_oracle.c — the CPU restatement of the reference's host algorithms.

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  Nothing under sputnik_b200/ may import this module.

All arrays are numpy, fp32 values (already rounded to the 16-bit device type by the caller),
int32 offsets, int16 indices.  Function names follow the reference items they restate:
  matmul            sputnik/matrix_utils.h:376-391   Matrix::operator*
  transpose         sputnik/matrix_utils.cu:635-643  Matrix::T()
  densify           sputnik/block/matrix_utils.h:81-112  ToMatrix()
  row_indices       sputnik/block/row_indices/row_indices.cu:7-19
  transpose_metadata sputnik/block/transpose/transpose.cu:69-125
  sdd / dsd / dds   the expectations of sdd_test.cu:163-195, dsd_test.cu:160-192, dds_test.cu:160-192
  ssd / sds / dss   the expectations of ssd_test.cu:215-222, sds_test.cu:220-226, dss_test.cu:226-232
  bitmask           sputnik/block/bitmask/bitmask.cu:7-45 + bit_matrix.h:12-62
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "_build", "liboracle.so")
BLOCK = 128


def build(force=False):
    src = os.path.join(_HERE, "bsmm_oracle.c")
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(src):
        os.makedirs(os.path.dirname(_LIB), exist_ok=True)
        subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-fopenmp", "-shared", "-fPIC",
                               "-o", _LIB, src])
    return _LIB


def _load():
    lib = ctypes.CDLL(build())
    lib.oracle_num_threads.restype = ctypes.c_int
    return lib


_lib = _load()
_f = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_i32 = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_i16 = np.ctypeslib.ndpointer(np.int16, flags="C_CONTIGUOUS")
_c = ctypes.c_int
_lib.oracle_matmul.argtypes = [_f, _f, _f, _c, _c, _c]
_lib.oracle_matmul_rows.argtypes = [_f, _f, _f, _c, _c, _c, _c, _c]
_lib.oracle_transpose.argtypes = [_f, _f, _c, _c]
_lib.oracle_densify.argtypes = [_f, _i32, _i16, _c, _c, _f]
_lib.oracle_sample.argtypes = [_f, _i32, _i16, _c, _c, _f]
_lib.oracle_row_indices.argtypes = [_i32, _c, _i16]
_lib.oracle_transpose_metadata.argtypes = [_i32, _i16, _c, _c, _i32, _i16, _i32]
_lib.oracle_dsd_blocked.argtypes = [_f, _i32, _i16, _c, _c, _c, _f, _c, _f, _c]
_lib.oracle_dds_blocked.argtypes = [_f, _c, _c, _f, _i32, _i16, _c, _c, _c, _f]
_lib.oracle_sdd_blocked.argtypes = [_f, _c, _f, _c, _c, _i32, _i16, _c, _c, _f]
_lib.oracle_set_threads.argtypes = [_c]
_lib.oracle_bitmask.argtypes = [_i32, _i16, _c, _c,
                                np.ctypeslib.ndpointer(np.uint64, flags="C_CONTIGUOUS")]


def num_threads():
    return int(_lib.oracle_num_threads())


def set_threads(n):
    _lib.oracle_set_threads(int(n))


def _f32(x):
    return np.ascontiguousarray(x, dtype=np.float32)


# ------------------------------------------------------------------ 16-bit rounding (host side)
def round_to(x, dtype):
    """Round fp32 -> fp16 / bf16 (round-to-nearest-even) and back to fp32."""
    x = _f32(x)
    if dtype in ("f16", "fp16", "float16"):
        return x.astype(np.float16).astype(np.float32)
    if dtype in ("f8", "e4m3", "f8e4m3"):
        return round_e4m3(x)
    u = x.view(np.uint32).astype(np.uint64)
    u = (u + 0x7FFF + ((u >> 16) & 1)) & 0xFFFF0000
    return u.astype(np.uint32).view(np.float32).reshape(x.shape)


def round_e4m3(x):
    """fp32 -> FP8 e4m3 (the `fn` flavour: 4 exponent bits, bias 7, 3 mantissa bits, no infinity,
    largest finite 448, subnormals down to 2^-9) -> fp32, round-to-nearest-even, saturating.
    The operand type of BSMM_F8E4M3 products; pinned against torch.float8_e4m3fn in
    tests/test_oracle.py."""
    x = np.asarray(x, np.float64)
    a = np.abs(x)
    with np.errstate(divide="ignore"):
        e = np.floor(np.log2(np.where(a > 0, a, 1.0)))
    e = np.maximum(e, -6.0)                    # below 2^-6: the subnormal quantum 2^-9
    q = np.exp2(e - 3.0)
    r = np.minimum(np.round(a / q) * q, 448.0)  # np.round: half to even
    return (np.sign(x) * r).astype(np.float32)


def to_bits16(x, dtype):
    """fp32 (already representable) -> uint16 bit pattern of fp16 / bf16."""
    x = _f32(x)
    if dtype in ("f16", "fp16", "float16"):
        return x.astype(np.float16).view(np.uint16)
    return (round_to(x, "bf16").view(np.uint32) >> 16).astype(np.uint16)


def from_bits16(b, dtype):
    b = np.ascontiguousarray(b, dtype=np.uint16)
    if dtype in ("f16", "fp16", "float16"):
        return b.view(np.float16).astype(np.float32)
    return (b.astype(np.uint32) << 16).view(np.float32)


# ------------------------------------------------------------------ as-written (dense) path
def matmul(a, b):
    a, b = _f32(a), _f32(b)
    m, k = a.shape
    k2, n = b.shape
    assert k == k2
    out = np.empty((m, n), np.float32)
    _lib.oracle_matmul(a, b, out, m, k, n)
    return out


def matmul_rows(a, b, row0, row1):
    a, b = _f32(a), _f32(b)
    m, k = a.shape
    n = b.shape[1]
    out = np.zeros((m, n), np.float32)
    _lib.oracle_matmul_rows(a, b, out, m, k, n, row0, row1)
    return out[row0:row1]


def transpose(a):
    a = _f32(a)
    out = np.empty((a.shape[1], a.shape[0]), np.float32)
    _lib.oracle_transpose(a, out, a.shape[0], a.shape[1])
    return out


def densify(values, offsets, indices, block_rows, block_cols):
    out = np.empty((block_rows * BLOCK, block_cols * BLOCK), np.float32)
    _lib.oracle_densify(_f32(values), offsets, indices, block_rows, block_cols, out)
    return out


def sample(dense, offsets, indices, block_rows, block_cols):
    out = np.empty((len(indices), BLOCK, BLOCK), np.float32)
    _lib.oracle_sample(_f32(dense), offsets, indices, block_rows, block_cols, out)
    return out


def row_indices(offsets, block_rows):
    out = np.empty(int(offsets[block_rows]), np.int16)
    _lib.oracle_row_indices(offsets, block_rows, out)
    return out


def transpose_metadata(offsets, indices, block_rows, block_cols):
    nb = int(offsets[block_rows])
    ot = np.empty(block_cols + 1, np.int32)
    it = np.empty(nb, np.int16)
    bo = np.empty(nb, np.int32)
    _lib.oracle_transpose_metadata(offsets, indices, block_rows, block_cols, ot, it, bo)
    return ot, it, bo


def _op(x, t):
    return transpose(x) if t else _f32(x)


def sdd_dense(a, ta, b, tb, topo):
    """sdd_test.cu:163-195: ((ta ? A.T : A) * (tb ? B.T : B)) masked at the topology."""
    full = matmul(_op(a, ta), _op(b, tb))
    return sample(full, topo.offsets, topo.indices, topo.block_rows, topo.block_cols)


def dsd_dense(values, topo, ta, b, tb):
    ad = densify(values, topo.offsets, topo.indices, topo.block_rows, topo.block_cols)
    return matmul(_op(ad, ta), _op(b, tb))


def dds_dense(a, ta, values, topo, tb):
    bd = densify(values, topo.offsets, topo.indices, topo.block_rows, topo.block_cols)
    return matmul(_op(a, ta), _op(bd, tb))


# ------------------------------------------------------------------ block-aware (same bits)
def sdd(a, ta, b, tb, topo):
    a, b = _f32(a), _f32(b)
    k = a.shape[0] if ta else a.shape[1]
    out = np.empty((topo.nnz_blocks, BLOCK, BLOCK), np.float32)
    _lib.oracle_sdd_blocked(a, int(ta), b, int(tb), k, topo.offsets, topo.indices,
                            topo.block_rows, topo.block_cols, out)
    return out


def dsd(values, topo, ta, b, tb):
    b = _f32(b)
    n = b.shape[0] if tb else b.shape[1]
    m = (topo.block_cols if ta else topo.block_rows) * BLOCK
    out = np.empty((m, n), np.float32)
    _lib.oracle_dsd_blocked(_f32(values), topo.offsets, topo.indices, topo.block_rows,
                            topo.block_cols, int(ta), b, int(tb), out, n)
    return out


def dds(a, ta, values, topo, tb):
    a = _f32(a)
    m = a.shape[1] if ta else a.shape[0]
    n = (topo.block_rows if tb else topo.block_cols) * BLOCK
    out = np.empty((m, n), np.float32)
    _lib.oracle_dds_blocked(a, int(ta), m, _f32(values), topo.offsets, topo.indices,
                            topo.block_rows, topo.block_cols, int(tb), out)
    return out


# ------------------------------------------------------------------ siblings
# Compositions of the functions above: the reference's tests densify every sparse operand and use
# the same host matmul (ssd_test.cu:215-222); skipping zero blocks does not change a bit, because
# adding an exact 0.0 product to the double accumulator is the identity.
def ssd(values_a, topo_a, ta, b, tb, topo_c):
    """((ta ? A.T : A) * op(B)) masked at C's topology; A sparse."""
    full = dsd(values_a, topo_a, ta, b, tb)
    return sample(full, topo_c.offsets, topo_c.indices, topo_c.block_rows, topo_c.block_cols)


def sds(a, ta, values_b, topo_b, tb, topo_c):
    """(op(A) * (tb ? B.T : B)) masked at C's topology; B sparse."""
    full = dds(a, ta, values_b, topo_b, tb)
    return sample(full, topo_c.offsets, topo_c.indices, topo_c.block_rows, topo_c.block_cols)


def dss(values_a, topo_a, ta, values_b, topo_b, tb):
    """op(A) * op(B), both sparse, dense result (dss_test.cu:226-232)."""
    bd = densify(values_b, topo_b.offsets, topo_b.indices, topo_b.block_rows, topo_b.block_cols)
    return dsd(values_a, topo_a, ta, bd, tb)


def bitmask(offsets, indices, walk_rows, walk_cols):
    out = np.empty(walk_rows * ((walk_cols + 63) // 64), np.uint64)
    _lib.oracle_bitmask(np.ascontiguousarray(offsets, np.int32),
                        np.ascontiguousarray(indices, np.int16), walk_rows, walk_cols, out)
    return out
