"""Python mirror of the reference's block-sparse operator interface, over the C-ABI.

Same names and argument meaning as `namespace sputnik::block` (sputnik/sputnik.h:18-25,
sputnik/block/arguments.h:48-274): `BlockMatrix`, `Matrix`, `matmul` (the overload set
`Matmul`: SDD, DSD, DDS and the siblings SSD, SDS, DSS), `matmul_ex` (`MatmulEx`), `row_indices`
(`RowIndices`), `transpose` (`Transpose`), `bitmask` (`Bitmask`), `allocate_transpose_buffers`,
`allocate_row_indices_buffer`, `allocate_bitmask_buffers`.  torch is used only to own device
memory and streams; every computation goes through libbsmm.so.
"""
import ctypes

import torch

from . import _lib
from ._lib import BsmmError, CBlockMatrix, CMatrix, lib  # noqa: F401

BLOCK = 128


def _dtype_tag(t: torch.dtype) -> int:
    if t == torch.float16:
        return _lib.BSMM_F16
    if t == torch.bfloat16:
        return _lib.BSMM_BF16
    raise BsmmError(_lib.BSMM_ERR_DTYPE)


def _product_tag(inputs, out):
    """dtype tag of a product from its operand tensors: all 16-bit and equal (fp16 / bf16), or
    every input e4m3 with a bf16 output (BSMM_F8E4M3).  Anything else — e.g. a mix of fp16 and
    bf16, which the library would silently reinterpret — is BSMM_ERR_DTYPE."""
    f8 = getattr(torch, "float8_e4m3fn", None)
    ins = {t.dtype for t in inputs}
    if f8 is not None and ins == {f8}:
        if out.dtype != torch.bfloat16:
            raise BsmmError(_lib.BSMM_ERR_DTYPE)
        return _lib.BSMM_F8E4M3
    if len(ins) != 1 or out.dtype not in ins:
        raise BsmmError(_lib.BSMM_ERR_DTYPE)
    return _dtype_tag(out.dtype)


def _ptr(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


class Matrix:
    """Dense row-major operand {rows, cols, data} (arguments.h:155-162)."""

    def __init__(self, rows, cols, data):
        self.rows, self.cols, self.data = int(rows), int(cols), data

    @classmethod
    def of(cls, t: torch.Tensor):
        assert t.dim() == 2 and t.is_contiguous()
        return cls(t.shape[0], t.shape[1], t)

    def c(self):
        return CMatrix(self.rows, self.cols, _ptr(self.data))


class BlockMatrix:
    """BCSR handle (arguments.h:48-153). rows/cols/nonzeros in elements; `data` is
    [nnz_blocks, 128, 128]; offsets int32; indices/indices_t/row_indices int16;
    offsets_t/block_offsets int32."""

    def __init__(self, rows, cols, block_size, nonzeros, data, offsets, indices,
                 offsets_t=None, indices_t=None, block_offsets=None, row_indices=None):
        self.rows, self.cols, self.nonzeros = int(rows), int(cols), int(nonzeros)
        self.block_size = int(block_size)
        self.data, self.offsets, self.indices = data, offsets, indices
        self.offsets_t, self.indices_t, self.block_offsets = offsets_t, indices_t, block_offsets
        self.row_indices = row_indices
        self.create_metadata = True  # arguments.h:67-68
        # Topology hints (include/bsmm.h BSMM_FLAG_TWIN_*): block rows (cols) 2i, 2i+1 have
        # identical index lists.  Set by whoever built the topology; never guessed.
        self.twin_rows = False
        self.twin_cols = False
        self.even_rows = False  # every block row holds an even number of blocks
        # BSMM_FLAG_DEVICE_NNZ: sized for a capacity; the live block count is offsets[-1] on
        # the device (set by moe_topology)
        self.device_nnz = False

    @property
    def nnz_blocks(self):
        return self.nonzeros // (self.block_size * self.block_size)

    def c(self):
        return CBlockMatrix(self.rows, self.cols, self.nonzeros, self.block_size,
                            _ptr(self.data), _ptr(self.offsets), _ptr(self.indices),
                            _ptr(self.offsets_t), _ptr(self.indices_t),
                            _ptr(self.block_offsets), _ptr(self.row_indices))


def _stream_ptr(stream):
    if stream is None:
        stream = torch.cuda.current_stream()
    return ctypes.c_void_p(stream.cuda_stream)


def allocate_transpose_buffers(a: BlockMatrix):
    """AllocateTransposeBuffers (arguments.h:233-245)."""
    dev = a.offsets.device
    a.offsets_t = torch.empty(a.cols // a.block_size + 1, dtype=torch.int32, device=dev)
    a.indices_t = torch.empty(a.nnz_blocks, dtype=torch.int16, device=dev)
    a.block_offsets = torch.empty(a.nnz_blocks, dtype=torch.int32, device=dev)


def allocate_row_indices_buffer(a: BlockMatrix):
    """AllocateRowIndicesBuffer (arguments.h:259-263)."""
    a.row_indices = torch.empty(a.nnz_blocks, dtype=torch.int16, device=a.offsets.device)


def row_indices(a: BlockMatrix, out=None, stream=None):
    """RowIndices(BlockMatrix, short*, stream) (row_indices/row_indices.cu:24-36)."""
    out = a.row_indices if out is None else out
    ca = a.c()
    _lib.check(lib.bsmm_row_indices(ctypes.byref(ca), _ptr(out), _stream_ptr(stream)))
    return out


def analyze(a: BlockMatrix, stream=None):
    """bsmm_analyze_topology: compute the hint flags on the device (synchronous) and set
    a.twin_rows / a.even_rows (and a.twin_cols when the transposed metadata is present)."""
    ca = a.c()
    flags = ctypes.c_uint(0)
    _lib.check(lib.bsmm_analyze_topology(ctypes.byref(ca), ctypes.byref(flags), _stream_ptr(stream)))
    a.twin_rows = bool(flags.value & _lib.FLAG_TWIN_ROWS)
    a.twin_cols = bool(flags.value & _lib.FLAG_TWIN_COLS)
    a.even_rows = bool(flags.value & _lib.FLAG_EVEN_ROWS)
    return {"twin_rows": a.twin_rows, "twin_cols": a.twin_cols, "even_rows": a.even_rows}


def row_pointers(dest, row_index=None, base_ptr=None):
    """Device int64 table for dsd_scatter: entry i = address of row row_index[i] of `dest`
    ([rows, cols], row-major, contiguous), 0 where row_index[i] < 0.  `base_ptr` overrides
    dest.data_ptr() (a peer GPU's mapping of the same buffer)."""
    assert dest.is_contiguous() and dest.dim() == 2
    base = dest.data_ptr() if base_ptr is None else int(base_ptr)
    pitch = dest.shape[1] * dest.element_size()
    if row_index is None:
        row_index = torch.arange(dest.shape[0], device=dest.device)
    idx = row_index.to(device=dest.device, dtype=torch.int64)
    return torch.where(idx >= 0, base + idx * pitch, torch.zeros_like(idx))


def dsd_scatter(a: BlockMatrix, transpose_a, b: Matrix, transpose_b, row_ptrs, cols, dtype=None,
                stream=None):
    """bsmm_dsd_scatter: op(A) * op(B) with output row i stored at address row_ptrs[i] (device
    int64 tensor; 0 drops the row).  The addresses may be in peer-GPU memory."""
    assert row_ptrs.dtype == torch.int64 and row_ptrs.is_contiguous()
    ca, cb = a.c(), b.c()
    dt = _dtype_tag(a.data.dtype if dtype is None else dtype)
    _lib.check(lib.bsmm_dsd_scatter(ctypes.byref(ca), int(bool(transpose_a)), ctypes.byref(cb),
                                    int(bool(transpose_b)), int(row_ptrs.numel()), int(cols),
                                    _ptr(row_ptrs), dt, _hints(a), _stream_ptr(stream)))


def gather_rows(row_ptrs, dst, stream=None):
    """bsmm_gather_rows: dst[i] = the row at address row_ptrs[i] (device int64; 0 -> zero row).
    The addresses may be in peer-GPU memory."""
    assert row_ptrs.dtype == torch.int64 and row_ptrs.is_contiguous()
    assert dst.is_contiguous() and dst.dim() == 2 and dst.element_size() == 2
    assert row_ptrs.numel() == dst.shape[0]
    _lib.check(lib.bsmm_gather_rows(_ptr(row_ptrs), dst.shape[0], dst.shape[1], _ptr(dst),
                                    _stream_ptr(stream)))
    return dst


def moe_topology(tokens_per_expert, ffn_blocks, cap_block_rows, row_align_blocks=2,
                 transposed=True, stream=None):
    """bsmm_moe_topology: the dMoE block-diagonal topology for the device-resident router
    histogram `tokens_per_expert` (int32 [E]), built by one kernel with no host round trip.
    Returns (BlockMatrix sized for `cap_block_rows` block rows, data=None, hints set,
    device_nnz=True; padded_bins int32 [E + 1])."""
    t = tokens_per_expert
    assert t.dtype == torch.int32 and t.is_contiguous()
    E, dev = t.numel(), t.device
    cap_nnz = cap_block_rows * ffn_blocks
    bs = 128
    m = BlockMatrix(cap_block_rows * bs, E * ffn_blocks * bs, bs, cap_nnz * bs * bs, None,
                    torch.empty(cap_block_rows + 1, dtype=torch.int32, device=dev),
                    torch.empty(cap_nnz, dtype=torch.int16, device=dev))
    allocate_row_indices_buffer(m)
    if transposed:
        allocate_transpose_buffers(m)
    bins = torch.empty(E + 1, dtype=torch.int32, device=dev)
    cm = m.c()
    _lib.check(lib.bsmm_moe_topology(_ptr(t), E, ffn_blocks, row_align_blocks, ctypes.byref(cm),
                                     _ptr(bins), _stream_ptr(stream)))
    # hints hold by construction (empty tail rows are twins of each other)
    m.twin_rows = row_align_blocks == 2
    m.twin_cols = m.even_rows = ffn_blocks % 2 == 0
    m.device_nnz = True
    return m, bins


def moe_exchange(counts, rank, cap_block_rows, row_pitch_bytes, peer_bases, tables,
                 tokens_per_expert=None, row_align_blocks=2, stream=None):
    """bsmm_moe_exchange: the gather / scatter address tables of an expert-sharded dMoE layer from
    the device-resident routing counts (int32 [world, experts]) — one kernel, no host round trip.
    `peer_bases[t][s]` = address of rank s's mapping of the buffer table t points into;
    `tables[t]`: int64 [cap_block_rows * 128]; `tokens_per_expert`: int32 [experts / world] out."""
    assert counts.dtype == torch.int32 and counts.is_contiguous() and counts.dim() == 2
    world, experts = counts.shape
    nt = len(tables)
    assert len(peer_bases) == nt and all(len(b) == world for b in peer_bases)
    for t in tables:
        assert t.dtype == torch.int64 and t.is_contiguous() and t.numel() == cap_block_rows * BLOCK
    bases = (ctypes.c_void_p * (nt * world))(*[int(a) for b in peer_bases for a in b])
    tabs = (ctypes.c_void_p * nt)(*[t.data_ptr() for t in tables])
    _lib.check(lib.bsmm_moe_exchange(_ptr(counts), world, experts, int(rank), int(row_align_blocks),
                                     int(cap_block_rows), int(row_pitch_bytes), nt, bases, tabs,
                                     _ptr(tokens_per_expert), _stream_ptr(stream)))


def allocate_bitmask_buffers(m: BlockMatrix):
    """AllocateBitmaskBuffers (bitmask/bitmask.h:16-23): uint64 words, K-contiguous; sized for
    the transposed view when the transposed metadata buffers exist."""
    cm = m.c()
    n = ctypes.c_size_t(0)
    _lib.check(lib.bsmm_bitmask_bytes(ctypes.byref(cm), ctypes.byref(n)))
    m.bitmask = torch.empty(n.value // 8, dtype=torch.int64, device=m.offsets.device)
    return m.bitmask


def bitmask(m: BlockMatrix, out=None, stream=None):
    """Bitmask(BlockMatrix, stream) (bitmask/bitmask.cu:7-45) — on the device.  The library's DSS
    does not need it; it exists for callers that consume the reference's bitmask layout."""
    out = getattr(m, "bitmask", None) if out is None else out
    cm = m.c()
    _lib.check(lib.bsmm_bitmask(ctypes.byref(cm), _ptr(out), _stream_ptr(stream)))
    return out


def transpose(a: BlockMatrix, stream=None):
    """Transpose(BlockMatrix, stream) (transpose/transpose.cu:69-125) — on the device."""
    ca = a.c()
    _lib.check(lib.bsmm_transpose_metadata(ctypes.byref(ca), _stream_ptr(stream)))


def _hints(m):
    return ((_lib.FLAG_TWIN_ROWS if m.twin_rows else 0) | (_lib.FLAG_TWIN_COLS if m.twin_cols else 0)
            | (_lib.FLAG_EVEN_ROWS if m.even_rows else 0)
            | (_lib.FLAG_DEVICE_NNZ if getattr(m, "device_nnz", False) else 0))


_ACT_FLAGS = {None: 0, "none": 0, "relu": _lib.FLAG_RELU, "gelu": _lib.FLAG_GELU,
              "swiglu": _lib.FLAG_SWIGLU}   # swiglu: SDD only, op(B) has 2 x the columns of C


def _matmul(a, transpose_a, b, transpose_b, c, stream, create_metadata, accumulate=False,
            activation=None):
    ta, tb = int(bool(transpose_a)), int(bool(transpose_b))
    sp = _stream_ptr(stream)
    epi = (_lib.FLAG_ACCUMULATE if accumulate else 0) | _ACT_FLAGS[activation]
    if isinstance(a, Matrix) and isinstance(b, Matrix) and isinstance(c, BlockMatrix):
        ca, cb, cc = a.c(), b.c(), c.c()
        st = lib.bsmm_sdd(ctypes.byref(ca), ta, ctypes.byref(cb), tb, ctypes.byref(cc),
                          _product_tag((a.data, b.data), c.data), _hints(c) | epi, sp)
    elif isinstance(a, BlockMatrix) and isinstance(b, Matrix) and isinstance(c, Matrix):
        ca, cb, cc = a.c(), b.c(), c.c()
        flags = _lib.FLAG_CREATE_METADATA if (create_metadata and a.create_metadata) else 0
        flags |= _hints(a)
        st = lib.bsmm_dsd(ctypes.byref(ca), ta, ctypes.byref(cb), tb, ctypes.byref(cc),
                          _product_tag((a.data, b.data), c.data), flags | epi, sp)
    elif isinstance(a, Matrix) and isinstance(b, BlockMatrix) and isinstance(c, Matrix):
        ca, cb, cc = a.c(), b.c(), c.c()
        flags = _lib.FLAG_CREATE_METADATA if (create_metadata and b.create_metadata) else 0
        flags |= _hints(b)
        st = lib.bsmm_dds(ctypes.byref(ca), ta, ctypes.byref(cb), tb, ctypes.byref(cc),
                          _product_tag((a.data, b.data), c.data), flags | epi, sp)
    elif isinstance(a, BlockMatrix) and isinstance(b, Matrix) and isinstance(c, BlockMatrix):
        ca, cb, cc = a.c(), b.c(), c.c()
        flags = _lib.FLAG_CREATE_METADATA if (create_metadata and a.create_metadata) else 0
        st = lib.bsmm_ssd(ctypes.byref(ca), ta, ctypes.byref(cb), tb, ctypes.byref(cc),
                          _dtype_tag(c.data.dtype), flags | epi, sp)
    elif isinstance(a, Matrix) and isinstance(b, BlockMatrix) and isinstance(c, BlockMatrix):
        ca, cb, cc = a.c(), b.c(), c.c()
        flags = _lib.FLAG_CREATE_METADATA if (create_metadata and b.create_metadata) else 0
        st = lib.bsmm_sds(ctypes.byref(ca), ta, ctypes.byref(cb), tb, ctypes.byref(cc),
                          _dtype_tag(c.data.dtype), flags | epi, sp)
    elif isinstance(a, BlockMatrix) and isinstance(b, BlockMatrix) and isinstance(c, Matrix):
        # create_metadata is per matrix in the reference (dss_..._tn_align8.cu:85-97): build the
        # column views that are asked for here, then run the product on valid metadata.
        if create_metadata and ta and a.create_metadata:
            transpose(a, stream)
        if create_metadata and not tb and b.create_metadata:
            transpose(b, stream)
        ca, cb, cc = a.c(), b.c(), c.c()
        st = lib.bsmm_dss(ctypes.byref(ca), ta, ctypes.byref(cb), tb, ctypes.byref(cc),
                          _dtype_tag(c.data.dtype), epi, sp)
    else:
        raise TypeError("unsupported operand combination (SDD, DSD, DDS, SSD, SDS, DSS exist; "
                        "sparse x sparse -> sparse does not in the reference either)")
    _lib.check(st)
    return c


def matmul(a, transpose_a, b, transpose_b, c, stream=None, accumulate=False, activation=None):
    """sputnik::block::Matmul overload set (sputnik.h:18-25): SDD / DSD / DDS / SSD / SDS / DSS
    chosen by which operands are BlockMatrix. Builds transposed metadata first when it is needed.
    Extensions (include/bsmm.h BSMM_FLAG_ACCUMULATE / RELU / GELU): accumulate=True computes
    C += op(A) op(B); activation in {None, "relu", "gelu"} is fused into the epilogue."""
    return _matmul(a, transpose_a, b, transpose_b, c, stream, True, accumulate, activation)


def matmul_ex(a, transpose_a, b, transpose_b, c, stream=None, accumulate=False, activation=None):
    """sputnik::block::MatmulEx (dsd.cu:20-27, dds.cu:17-24): transposed metadata is taken as
    already valid."""
    return _matmul(a, transpose_a, b, transpose_b, c, stream, False, accumulate, activation)
