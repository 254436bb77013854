"""Shared helpers of the parity tests: seeded operands, device handles, error metrics.

The CUDA path is always reached through sputnik_b200.block, which is a ctypes veneer over the
C-ABI (include/bsmm.h); the expectation always comes from oracle/ (CPU)."""
import ctypes
import os

import numpy as np

from oracle import oracle
from sputnik_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_GPU_LIB = os.path.join(ROOT, "oracle", "_ref", "libsputnik_ref.so")
REF_HOST_LIB = os.path.join(ROOT, "oracle", "_ref", "libsputnik_ref_host.so")

REL_TOL = 1e-2   # north_star: max relative error vs the FP32 host reference
ABS_TOL = 5e-2   # the reference's own bar, fp16 (sdd_test.cu:195, dsd_test.cu:192, dds_test.cu:192)


def errors(got, ref):
    """(max relative error with the tau guard of SURVEY 8c, max absolute error)."""
    got = np.asarray(got, np.float64)
    ref = np.asarray(ref, np.float64)
    if ref.size == 0:
        return 0.0, 0.0
    tau = 1e-2 * float(np.sqrt(np.mean(ref ** 2)))
    diff = np.abs(got - ref)
    return float(np.max(diff / np.maximum(np.abs(ref), max(tau, 1e-30)))), float(np.max(diff))


def assert_close(got, ref, dtype, what=""):
    assert np.isfinite(got).all(), f"{what}: non-finite output"
    rel, ab = errors(got, ref)
    assert rel <= REL_TOL, f"{what}: max rel err {rel:.3e} > {REL_TOL}"
    if dtype == "f16":
        assert ab <= ABS_TOL, f"{what}: max abs err {ab:.3e} > {ABS_TOL}"
    return rel, ab


class Case:
    """One seeded problem in the reference's parametrisation (M, K, N, nonzeros, ta, tb,
    unordered): host fp32 operands rounded to `dtype`, plus the topology."""

    def __init__(self, op, m, k, n, nonzeros, ta, tb, unordered, dtype, seed):
        self.op, self.m, self.k, self.n = op, m, k, n
        self.ta, self.tb, self.dtype, self.seed = bool(ta), bool(tb), dtype, seed
        nb = nonzeros // (128 * 128)
        if op == "sdd":
            sr, sc = m, n
        elif op == "dsd":                      # dsd_test.cu:160-166
            sr, sc = (k, m) if ta else (m, k)
        else:                                  # dds_test.cu: sparse rhs
            sr, sc = (n, k) if tb else (k, n)
        self.topo = synth.random_topology(sr // 128, sc // 128, nb, seed, unordered)
        r = lambda shape, s: oracle.round_to(synth.uniform_values(shape, seed * 7919 + s), dtype)
        self.values = r((nb, 128, 128), 1)
        if op != "dsd":
            self.a = r((k, m) if ta else (m, k), 2)
        if op != "dds":
            self.b = r((n, k) if tb else (k, n), 3)

    def expected(self, dense=False):
        o = oracle
        if self.op == "sdd":
            return (o.sdd_dense if dense else o.sdd)(self.a, self.ta, self.b, self.tb, self.topo)
        if self.op == "dsd":
            return (o.dsd_dense if dense else o.dsd)(self.values, self.topo, self.ta, self.b, self.tb)
        return (o.dds_dense if dense else o.dds)(self.a, self.ta, self.values, self.topo, self.tb)


class SibCase:
    """One seeded SSD / SDS / DSS problem in the reference's parametrisation (ssd_test.cu:30-52,
    sds_test.cu:30-52, dss_test.cu:30-57): (M, K, N, nonzeros1, nonzeros2, ta, tb, unordered);
    nonzeros1 belongs to the first sparse operand, nonzeros2 to the sparse output (SSD, SDS) or
    the second sparse operand (DSS)."""

    def __init__(self, op, m, k, n, nnz1, nnz2, ta, tb, unordered, dtype, seed):
        self.op, self.m, self.k, self.n = op, m, k, n
        self.ta, self.tb, self.dtype, self.seed = bool(ta), bool(tb), dtype, seed
        r = lambda shape, s: oracle.round_to(synth.uniform_values(shape, seed * 7919 + s), dtype)
        topo = lambda rows, cols, nnz, s: synth.random_topology(rows // 128, cols // 128,
                                                                nnz // 16384, seed + s, unordered)
        a_shape = (k, m) if ta else (m, k)
        b_shape = (n, k) if tb else (k, n)
        self.topo_a = self.topo_b = self.topo_c = None
        if op in ("ssd", "dss"):
            self.topo_a = topo(*a_shape, nnz1, 0)
            self.a = r((self.topo_a.nnz_blocks, 128, 128), 1)
        else:
            self.a = r(a_shape, 1)
        if op == "ssd":
            self.b = r(b_shape, 2)
        else:
            self.topo_b = topo(*b_shape, nnz1 if op == "sds" else nnz2, 101)
            self.b = r((self.topo_b.nnz_blocks, 128, 128), 2)
        if op != "dss":
            self.topo_c = topo(m, n, nnz2, 202)

    def expected(self):
        if self.op == "ssd":
            return oracle.ssd(self.a, self.topo_a, self.ta, self.b, self.tb, self.topo_c)
        if self.op == "sds":
            return oracle.sds(self.a, self.ta, self.b, self.topo_b, self.tb, self.topo_c)
        return oracle.dss(self.a, self.topo_a, self.ta, self.b, self.topo_b, self.tb)


# ---------------------------------------------------------------------------- device side
def torch_dtype(dtype):
    import torch
    if dtype == "f8":
        return torch.float8_e4m3fn
    return torch.float16 if dtype == "f16" else torch.bfloat16


def out_dtype(dtype):
    """Output type of a product whose operands are `dtype` (e4m3 operands give bf16)."""
    return "bf16" if dtype == "f8" else dtype


def to_dev(x, dtype=None, dev="cuda:0"):
    import torch
    t = torch.from_numpy(np.ascontiguousarray(x)).to(dev)
    return t.to(torch_dtype(dtype)) if dtype else t


def dev_block_matrix(topo, values, dtype, dev="cuda:0", fill=None):
    import torch
    import sputnik_b200 as sb
    if fill is not None:
        v = torch.full((topo.nnz_blocks, 128, 128), fill, dtype=torch.float32,
                       device=dev).to(torch_dtype(dtype))
    else:
        v = to_dev(values, dtype, dev)
    return sb.BlockMatrix(topo.rows, topo.cols, 128, topo.nonzeros, v,
                          to_dev(topo.offsets, None, dev), to_dev(topo.indices, None, dev))


def run_ours(case, ex=False, dev="cuda:0", prefill=None, hints=False, **kw):
    """Run the CUDA path on a Case; returns the result as fp32 numpy.  `prefill`: initial
    contents of the output (fp32 numpy, representable in the 16-bit type); `hints`: pass the
    topology's twin / even hints (2-SM kernels); `kw`: accumulate= / activation= of matmul."""
    import torch
    import sputnik_b200 as sb
    mm = (lambda *a: sb.matmul_ex(*a, **kw)) if ex else (lambda *a: sb.matmul(*a, **kw))
    dt, odt = case.dtype, out_dtype(case.dtype)
    if case.op == "sdd":
        c = dev_block_matrix(case.topo, prefill, odt, dev, fill=None if prefill is not None else float("nan"))
        if hints:
            c.__dict__.update(case.topo.hints())
        sb.allocate_row_indices_buffer(c)
        sb.row_indices(c)
        mm(sb.Matrix.of(to_dev(case.a, dt, dev)), case.ta, sb.Matrix.of(to_dev(case.b, dt, dev)),
           case.tb, c)
        out = c.data
    else:
        sp = dev_block_matrix(case.topo, case.values, dt, dev)
        if hints:
            sp.__dict__.update(case.topo.hints())
        sb.allocate_transpose_buffers(sp)
        if ex:
            sb.transpose(sp)
        # random pre-fill: every output element must be overwritten (dsd_test.cu:176)
        out = (to_dev(prefill, odt, dev) if prefill is not None
               else torch.randn(case.m, case.n, device=dev).to(torch_dtype(odt)))
        if case.op == "dsd":
            mm(sp, case.ta, sb.Matrix.of(to_dev(case.b, dt, dev)), case.tb, sb.Matrix.of(out))
        else:
            mm(sb.Matrix.of(to_dev(case.a, dt, dev)), case.ta, sp, case.tb, sb.Matrix.of(out))
    torch.cuda.synchronize()
    return out.float().cpu().numpy()


def run_sibling(case, ex=False, dev="cuda:0"):
    """Run SSD / SDS / DSS of the CUDA path on a SibCase; returns the result as fp32 numpy."""
    import torch
    import sputnik_b200 as sb
    mm = sb.matmul_ex if ex else sb.matmul
    dt = case.dtype

    def sparse(topo, values):
        sp = dev_block_matrix(topo, values, dt, dev)
        sb.allocate_transpose_buffers(sp)
        if ex:
            sb.transpose(sp)
        return sp

    a = sparse(case.topo_a, case.a) if case.topo_a is not None else sb.Matrix.of(to_dev(case.a, dt, dev))
    b = sparse(case.topo_b, case.b) if case.topo_b is not None else sb.Matrix.of(to_dev(case.b, dt, dev))
    if case.topo_c is not None:
        c = dev_block_matrix(case.topo_c, None, dt, dev, fill=float("nan"))
        sb.allocate_row_indices_buffer(c)
        sb.row_indices(c)
        out = c.data
    else:
        out = torch.randn(case.m, case.n, device=dev).to(torch_dtype(dt))
        c = sb.Matrix.of(out)
    mm(a, case.ta, b, case.tb, c)
    torch.cuda.synchronize()
    return out.float().cpu().numpy()


# ---------------------------------------------------------------------------- the reference
def load_ref_gpu():
    if not os.path.exists(REF_GPU_LIB):
        return None
    return ctypes.CDLL(REF_GPU_LIB)


def _p(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def run_ref_gpu(lib, case, dev="cuda:0"):
    """Run the reference's own CUTLASS kernels (oracle/_ref, fp16 only) on a Case."""
    import torch
    assert case.dtype == "f16"
    t = case.topo
    off, idx = to_dev(t.offsets), to_dev(t.indices)
    ot = torch.empty(t.block_cols + 1, dtype=torch.int32, device=dev)
    it = torch.empty(t.nnz_blocks, dtype=torch.int16, device=dev)
    bo = torch.empty(t.nnz_blocks, dtype=torch.int32, device=dev)
    if case.op == "sdd":
        a, b = to_dev(case.a, "f16"), to_dev(case.b, "f16")
        v = torch.full((t.nnz_blocks, 128, 128), float("nan"), dtype=torch.float16, device=dev)
        ri = torch.empty(t.nnz_blocks, dtype=torch.int16, device=dev)
        st = lib.ref_row_indices(t.rows, t.cols, t.nonzeros, _p(off), _p(idx), _p(ri), None)
        assert st == 0
        st = lib.ref_sdd(a.shape[0], a.shape[1], _p(a), int(case.ta), b.shape[0], b.shape[1],
                         _p(b), int(case.tb), t.rows, t.cols, t.nonzeros, _p(v), _p(off), _p(idx),
                         _p(ri), None)
        out = v
    elif case.op == "dsd":
        v, b = to_dev(case.values, "f16"), to_dev(case.b, "f16")
        out = torch.randn(case.m, case.n, device=dev).half()
        st = lib.ref_dsd(t.rows, t.cols, t.nonzeros, _p(v), _p(off), _p(idx), _p(ot), _p(it),
                         _p(bo), int(case.ta), b.shape[0], b.shape[1], _p(b), int(case.tb),
                         case.m, case.n, _p(out), 1, None)
    else:
        a, v = to_dev(case.a, "f16"), to_dev(case.values, "f16")
        out = torch.randn(case.m, case.n, device=dev).half()
        st = lib.ref_dds(a.shape[0], a.shape[1], _p(a), int(case.ta), t.rows, t.cols, t.nonzeros,
                         _p(v), _p(off), _p(idx), _p(ot), _p(it), _p(bo), int(case.tb), case.m,
                         case.n, _p(out), 1, None)
    assert st == 0, f"reference returned cudaError {st}"
    torch.cuda.synchronize()
    return out.float().cpu().numpy()
