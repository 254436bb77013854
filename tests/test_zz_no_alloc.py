"""bsmm_dsd_static (include/bsmm.h): the DSD entry point that allocates nothing — its scratch (tile
counter of the dynamic scheduler + the transposed metadata it builds itself) is one file-scope
`static __device__` array of the library.  No counterpart in the reference; added
for callers that cannot call cudaMalloc.

Called through the C-ABI (ctypes on libbsmm.so) and compared with the existing entry point
bsmm_dsd on the same operands, two (three) calls in a row on one stream with DIFFERENT
topologies, so a call that read the previous call's scratch would be caught.  Tolerance:
tests/util.assert_close (max relative error <= 1e-2 with the tau guard), as everywhere.

This file sorts last on purpose."""
import ctypes

import numpy as np
import pytest

from sputnik_b200 import _lib
from sputnik_b200._lib import CBlockMatrix, CMatrix
from util import Case, assert_close

DEV = "cuda:0"
BF16 = _lib.BSMM_BF16


def _stage(case, with_transpose_buffers):
    """Device operands of a DSD Case as C structs (+ the tensors that own the memory)."""
    import torch
    import sputnik_b200 as sb
    from util import dev_block_matrix, to_dev
    sp = dev_block_matrix(case.topo, case.values, case.dtype, DEV)
    if with_transpose_buffers:
        sb.allocate_transpose_buffers(sp)
    b = to_dev(case.b, case.dtype, DEV)
    out = torch.full((case.m, case.n), float("nan"), dtype=b.dtype, device=DEV)
    keep = (sp, b, out)
    cm = lambda t: CMatrix(t.shape[0], t.shape[1], ctypes.c_void_p(t.data_ptr()))
    return sp.c(), cm(b), cm(out), out, keep


def _cases(ta, tb, dtype="bf16"):
    # (M, K, N, nonzeros): 32 x 32 block grid at 25 % with N = 1536 -> 192 tiles (> 148 SMs: the
    # dynamic tile counter is in use); then a smaller, differently shaped topology; then ragged N.
    shapes = [(4096, 4096, 1536, 256), (1024, 2560, 264, 37), (2048, 1024, 4096, 70)]
    return [Case("dsd", m, k, n, nb * 16384, ta, tb, False, dtype, 100 + i)
            for i, (m, k, n, nb) in enumerate(shapes)]


@pytest.mark.gpu
@pytest.mark.parametrize("ta,tb", [(0, 0), (1, 0), (0, 1), (1, 1)])
def test_static_scratch_dsd_matches_bsmm_dsd_on_calls_in_a_row(ta, tb):
    import torch
    lib = _lib.lib
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    cases = _cases(ta, tb)
    # expectation: the existing entry point, metadata built by the call into caller buffers
    want = []
    for case in cases:
        a, b, c, out, keep = _stage(case, with_transpose_buffers=True)
        _lib.check(lib.bsmm_dsd(ctypes.byref(a), ta, ctypes.byref(b), tb, ctypes.byref(c), BF16,
                                _lib.FLAG_CREATE_METADATA, stream))
        torch.cuda.synchronize()
        want.append(out.float().cpu().numpy())
        del keep
    assert_close(want[0], cases[0].expected(), "bf16", "bsmm_dsd vs oracle")
    # the new entry point: everything staged first (no transpose buffers at all), then the calls
    # back to back on one stream, the first problem once more at the end
    staged = [_stage(case, with_transpose_buffers=False) for case in cases]
    staged.append(_stage(cases[0], with_transpose_buffers=False))
    _lib.check(lib.bsmm_dsd_static(*_args(staged[0], ta, tb), stream))   # first use: set-up
    torch.cuda.synchronize()
    staged[0][3].fill_(float("nan"))
    torch.cuda.synchronize()
    # no allocation during the calls: torch's cumulative allocation / segment counters (frees by
    # the garbage collector of earlier tests' tensors do not move them) and the driver's free memory
    import gc
    gc.collect()
    count = lambda: tuple(torch.cuda.memory_stats()[k] for k in ("allocation.all.allocated",
                                                                 "segment.all.allocated"))
    allocs_before = count()
    free_before = torch.cuda.mem_get_info()[0]
    for st in staged:
        _lib.check(lib.bsmm_dsd_static(*_args(st, ta, tb), stream))
    torch.cuda.synchronize()
    assert count() == allocs_before, "torch allocated during bsmm_dsd_static"
    assert torch.cuda.mem_get_info()[0] >= free_before, "device memory was allocated by the call"
    for i, st in enumerate(staged):
        got = st[3].float().cpu().numpy()
        ref = want[i % len(cases)]
        assert_close(got, ref, "bf16", f"call {i} (ta={ta}, tb={tb})")
        # same kernel family, same K order: the bits agree as well
        assert np.array_equal(got, ref), f"call {i}: differs from bsmm_dsd"


def _args(st, ta, tb, dtype=BF16, flags=0):
    a, b, c = st[0], st[1], st[2]
    return ctypes.byref(a), ta, ctypes.byref(b), tb, ctypes.byref(c), dtype, flags


@pytest.mark.gpu
def test_static_scratch_dsd_accumulate_f16_and_empty():
    """fp16 with BSMM_FLAG_ACCUMULATE against bsmm_dsd, and an all-empty topology (zero fill)."""
    import torch
    lib = _lib.lib
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    case = Case("dsd", 1024, 1536, 512, 40 * 16384, True, False, False, "f16", 7)
    outs = []
    for fn, bufs in ((lib.bsmm_dsd, True), (lib.bsmm_dsd_static, False)):
        st = _stage(case, with_transpose_buffers=bufs)
        st[3].fill_(0.5)
        _lib.check(fn(*_args(st, 1, 0, _lib.BSMM_F16,
                             _lib.FLAG_ACCUMULATE | _lib.FLAG_CREATE_METADATA), stream))
        torch.cuda.synchronize()
        outs.append(st[3].float().cpu().numpy())
    assert_close(outs[1], outs[0], "f16", "accumulate, fp16")
    assert_close(outs[1], case.expected() + 0.5, "f16", "accumulate, fp16 vs oracle")
    empty = Case("dsd", 512, 512, 256, 0, False, False, False, "bf16", 3)
    st = _stage(empty, with_transpose_buffers=False)
    _lib.check(lib.bsmm_dsd_static(*_args(st, 0, 0), stream))
    torch.cuda.synchronize()
    assert float(st[3].float().abs().max()) == 0.0


def test_static_scratch_dsd_rejects_sizes_beyond_the_static_array():
    """Beyond BSMM_STATIC_MAX_BLOCKS / BSMM_STATIC_MAX_BLOCK_COLS: BSMM_ERR_STATIC_LIMIT, decided
    before any CUDA call (runs without a GPU; the pointers are never dereferenced)."""
    lib = _lib.lib
    fake = ctypes.c_void_p(0x10000)

    def call(rows, cols, nb, ta):
        a = CBlockMatrix(rows, cols, nb * 16384, 128, fake, fake, fake, None, None, None, None)
        m, k = (cols, rows) if ta else (rows, cols)
        b = CMatrix(k, 256, fake)
        c = CMatrix(m, 256, fake)
        return lib.bsmm_dsd_static(ctypes.byref(a), ta, ctypes.byref(b), 0, ctypes.byref(c), BF16, 0,
                                   None)

    big = 200 * 128   # 200 x 200 block grid: room for 40 000 blocks
    assert call(big, big, _lib.STATIC_MAX_BLOCKS + 1, 0) == _lib.BSMM_ERR_STATIC_LIMIT
    assert call(big, big, _lib.STATIC_MAX_BLOCKS + 1, 1) == _lib.BSMM_ERR_STATIC_LIMIT
    assert call(128, (_lib.STATIC_MAX_BLOCK_COLS + 1) * 128, 4, 1) == _lib.BSMM_ERR_STATIC_LIMIT
    msg = lib.bsmm_status_string(_lib.BSMM_ERR_STATIC_LIMIT).decode()
    assert "16384" in msg and "2048" in msg
    # the ordinary argument errors come first, as for bsmm_dsd
    a = CBlockMatrix(256, 256, 16384, 64, fake, fake, fake, None, None, None, None)
    b, c = CMatrix(256, 256, fake), CMatrix(256, 256, fake)
    assert lib.bsmm_dsd_static(ctypes.byref(a), 0, ctypes.byref(b), 0, ctypes.byref(c), BF16, 0,
                               None) == _lib.BSMM_ERR_BLOCK_SIZE
