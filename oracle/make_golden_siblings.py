"""Generate tests/golden/host_matmul_siblings.npz from the REFERENCE ITSELF
(oracle/_ref/libsputnik_ref_host.so, built from /root/reference by `make -C oracle ref_host`):

    python oracle/make_golden_siblings.py

Each fixture: a seeded SibCase (tests/util.py) -> the reference's expectation exactly as its
tests form it (ssd_test.cu:215-222, sds_test.cu:220-226, dss_test.cu:226-232): every sparse
operand densified, Matrix::T() for the transpose flags, Matrix::operator*, and for the sparse
outputs Matrix::Mul with the densified all-ones output topology.  Densification is an index copy
done here in numpy; every floating-point operation goes through the reference library.
TEST INFRASTRUCTURE ONLY.
"""
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path = [p for p in sys.path if os.path.abspath(p or ".") != os.path.join(ROOT, "oracle")]
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from util import SibCase, REF_HOST_LIB  # noqa: E402

CASES = [  # (op, M, K, N, nonzeros1, nonzeros2, ta, tb, unordered, dtype, seed)
    ("ssd", 256, 384, 256, 3 * 16384, 3 * 16384, False, False, False, "f16", 1),
    ("ssd", 256, 384, 256, 4 * 16384, 2 * 16384, True, False, True, "bf16", 2),
    ("ssd", 384, 256, 256, 4 * 16384, 5 * 16384, False, True, True, "f16", 3),
    ("ssd", 256, 256, 384, 2 * 16384, 4 * 16384, True, True, False, "bf16", 4),
    ("sds", 256, 384, 256, 3 * 16384, 3 * 16384, False, False, True, "f16", 5),
    ("sds", 256, 256, 384, 4 * 16384, 4 * 16384, True, False, False, "bf16", 6),
    ("sds", 384, 256, 256, 2 * 16384, 5 * 16384, False, True, False, "f16", 7),
    ("sds", 256, 384, 256, 5 * 16384, 2 * 16384, True, True, True, "bf16", 8),
    ("dss", 256, 384, 256, 3 * 16384, 4 * 16384, False, False, False, "f16", 9),
    ("dss", 384, 256, 256, 4 * 16384, 3 * 16384, False, True, True, "bf16", 10),
    ("dss", 256, 384, 256, 4 * 16384, 4 * 16384, True, False, True, "f16", 11),
    ("dss", 256, 256, 384, 3 * 16384, 5 * 16384, True, True, False, "bf16", 12),
]


def main():
    h = ctypes.CDLL(REF_HOST_LIB)
    f = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
    h.ref_host_matmul.argtypes = [f, f, f, ctypes.c_int, ctypes.c_int, ctypes.c_int]
    h.ref_host_transpose.argtypes = [f, f, ctypes.c_int, ctypes.c_int]
    h.ref_host_mul.argtypes = [f, f, ctypes.c_int, ctypes.c_int]

    def T(x):
        out = np.empty((x.shape[1], x.shape[0]), np.float32)
        h.ref_host_transpose(np.ascontiguousarray(x), out, x.shape[0], x.shape[1])
        return out

    def mm(a, b):
        out = np.empty((a.shape[0], b.shape[1]), np.float32)
        h.ref_host_matmul(np.ascontiguousarray(a), np.ascontiguousarray(b), out, a.shape[0],
                          a.shape[1], b.shape[1])
        return out

    def dense_of(topo, values):
        d = np.zeros((topo.rows, topo.cols), np.float32)
        ri = topo.row_indices()
        for j in range(topo.nnz_blocks):
            r, c = int(ri[j]), int(topo.indices[j])
            d[r * 128:(r + 1) * 128, c * 128:(c + 1) * 128] = values[j]
        return d

    out = {}
    for i, spec in enumerate(CASES):
        c = SibCase(*spec)
        lhs = dense_of(c.topo_a, c.a) if c.topo_a is not None else c.a
        rhs = dense_of(c.topo_b, c.b) if c.topo_b is not None else c.b
        exp = mm(T(lhs) if c.ta else lhs, T(rhs) if c.tb else rhs)
        if c.topo_c is not None:
            mask = dense_of(c.topo_c, np.ones((c.topo_c.nnz_blocks, 128, 128), np.float32))
            h.ref_host_mul(exp, mask, exp.shape[0], exp.shape[1])
        out[f"spec{i}"] = np.array([str(x) for x in spec])
        out[f"expected{i}"] = exp
    path = os.path.join(ROOT, "tests", "golden", "host_matmul_siblings.npz")
    np.savez_compressed(path, n=np.array(len(CASES)), **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
