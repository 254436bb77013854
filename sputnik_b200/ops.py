"""PyTorch operator binding of the block-sparse products: `sdd`, `dsd`, `dds`
(and the siblings `ssd`, `sds`, `dss`, end of file)
as `torch.library` custom operators (`torch.ops.sputnik_b200.*`, with fake-tensor shape functions
and autograd formulas for every transpose pair) over a device-resident `Topology`, so the path
can be used as a layer.  torch owns memory, streams and the autograd graph; every product —
forward and backward — is one call of the C-ABI (libbsmm.so), in the pattern MegaBlocks' dMoE uses:

    h = sdd(x, w1, topo)          dx  = dsd(dh, topo, w1^T)     dw1 = dds(x^T, dh, topo)
    y = dsd(h, topo, w2)          dh  = sdd(dy, w2^T, topo)     dw2 = dsd(h^T, topo, dy)

Sparse values travel as a plain tensor `[nnz_blocks, 128, 128]`; the topology (offsets, indices,
row_indices and the transposed metadata, built once on the device) is shared by every product of
a layer, exactly like the reference's `BlockMatrix` metadata (sputnik/block/arguments.h:48-68).
"""
import torch

from . import block as sb

BLOCK = 128


class Topology:
    """Device-resident BCSR metadata + topology hints.  Built from host index arrays (a
    `synth.Topology`) or from device tensors; row_indices and the transposed metadata are
    produced by the library's device kernels (bsmm_row_indices, bsmm_transpose_metadata)."""

    def __init__(self, rows, cols, offsets, indices, device=None, hints=None):
        dev = torch.device(device) if device is not None else offsets.device
        self.rows, self.cols = int(rows), int(cols)
        self.offsets = torch.as_tensor(offsets, dtype=torch.int32).to(dev).contiguous()
        idx = torch.as_tensor(indices)
        if self.rows // BLOCK > 32767 or self.cols // BLOCK > 32767 or \
                (idx.dtype != torch.int16 and idx.numel() and int(idx.max()) > 32767):
            raise ValueError("block indices are int16: at most 32767 block rows / columns")
        self.indices = idx.to(torch.int16).to(dev).contiguous()
        self.nnz_blocks = int(self.indices.numel())
        self.hints = dict(hints or {})
        m = self.matrix(None)
        sb.allocate_transpose_buffers(m)
        sb.allocate_row_indices_buffer(m)
        if self.nnz_blocks:
            sb.transpose(m)
            sb.row_indices(m)
        self.offsets_t, self.indices_t = m.offsets_t, m.indices_t
        self.block_offsets, self.row_indices = m.block_offsets, m.row_indices
        if hints is None and self.nnz_blocks:
            self.hints = sb.analyze(m)  # device-side analysis (one synchronous call per topology)

    @classmethod
    def from_host(cls, topo, device="cuda"):
        """`topo`: a sputnik_b200.synth.Topology (hints are derived exactly on the host)."""
        return cls(topo.rows, topo.cols, topo.offsets, topo.indices, device, topo.hints())

    @classmethod
    def for_moe(cls, tokens_per_expert, ffn, capacity_rows, row_align=256):
        """The dMoE topology of a device-resident router histogram (int32 [E]) — one kernel, no
        host synchronisation (bsmm_moe_topology).  Sized for `capacity_rows`
        padded token rows (>= sum(tokens) + E * (row_align - 1), rounded up to row_align); the
        live block count stays on the device and every product honours it (DEVICE_NNZ).
        `self.padded_bins` (int32 [E + 1]): first padded token row of each expert."""
        assert ffn % BLOCK == 0 and row_align in (128, 256) and capacity_rows % row_align == 0
        m, bins = sb.moe_topology(tokens_per_expert, ffn // BLOCK, capacity_rows // BLOCK,
                                  row_align // BLOCK)
        self = cls.__new__(cls)
        self.rows, self.cols = m.rows, m.cols
        self.offsets, self.indices = m.offsets, m.indices
        self.nnz_blocks = m.nnz_blocks
        self.offsets_t, self.indices_t = m.offsets_t, m.indices_t
        self.block_offsets, self.row_indices = m.block_offsets, m.row_indices
        self.hints = {"twin_rows": m.twin_rows, "twin_cols": m.twin_cols,
                      "even_rows": m.even_rows, "device_nnz": True}
        self.padded_bins = bins
        return self

    def matrix(self, values):
        """A BlockMatrix handle (arguments.h:48-68) over `values` with this topology."""
        m = sb.BlockMatrix(self.rows, self.cols, BLOCK, self.nnz_blocks * BLOCK * BLOCK, values,
                           self.offsets, self.indices)
        for name in ("offsets_t", "indices_t", "block_offsets", "row_indices"):
            setattr(m, name, getattr(self, name, None))
        m.__dict__.update(self.hints)
        return m

    def new_values(self, dtype, device=None):
        return torch.empty((self.nnz_blocks, BLOCK, BLOCK), dtype=dtype,
                           device=device or self.offsets.device)


def _dense(t):
    return sb.Matrix.of(t.contiguous())


# --------------------------------------------------------------------------- torch.library ops
#
# The three products as registered custom operators (`torch.ops.sputnik_b200.{sdd,dsd,dds}`):
# plain tensors and flags in, one tensor out, so they carry fake-tensor shape functions and
# autograd formulas and the surrounding model can be traced / exported without graph breaks.
# `meta` = the six topology tensors (offsets, indices, row_indices, offsets_t, indices_t,
# block_offsets) and `hints` the BSMM_FLAG_* word; `rows` / `cols` are the sparse matrix's extents.
# Every backward product is again one of the three operators (any transpose pair):
#
#   SDD  C = op(A) op(B) | topo     dA = ta ? dds(B, tb, dC, T)  : dsd(dC, N, B, !tb)
#                                   dB = tb ? dsd(dC, T, A, ta)  : dds(A, !ta, dC, N)
#   DSD  C = op(S) op(B)            dS = ta ? sdd(B, tb, dC, T)  : sdd(dC, N, B, !tb)
#                                   dB = tb ? dds(dC, T, S, ta)  : dsd(S, !ta, dC, N)
#   DDS  C = op(A) op(S)            dA = ta ? dsd(S, tb, dC, T)  : dds(dC, N, S, !tb)
#                                   dS = tb ? sdd(dC, T, A, ta)  : sdd(A, !ta, dC, N)

_META_DTYPES = (torch.int32, torch.int16, torch.int16, torch.int32, torch.int16, torch.int32)


def _check_operands(tensors, meta):
    """One device for everything, contiguous metadata of the BCSR types (include/bsmm.h:52-61);
    operand dtype agreement is checked by block._product_tag (BSMM_ERR_DTYPE)."""
    dev = tensors[0].device
    if dev.type != "cuda":
        raise ValueError("sputnik_b200 ops need CUDA tensors (there is no CPU fallback)")
    for t in tensors:
        if t.device != dev:
            raise ValueError(f"operands on different devices: {t.device} vs {dev}")
    for t, dt, name in zip(meta, _META_DTYPES, ("offsets", "indices", "row_indices", "offsets_t",
                                                "indices_t", "block_offsets")):
        if t.device != dev or t.dtype != dt or not t.is_contiguous():
            raise ValueError(f"topology tensor `{name}` must be a contiguous {dt} tensor on {dev}")
    return torch.cuda.device(dev)


def _matrix_of(rows, cols, values, meta, hints):
    offsets, indices, row_indices, offsets_t, indices_t, block_offsets = meta
    m = sb.BlockMatrix(rows, cols, BLOCK, indices.numel() * BLOCK * BLOCK, values, offsets, indices,
                       offsets_t, indices_t, block_offsets, row_indices)
    m.twin_rows = bool(hints & 2)
    m.twin_cols = bool(hints & 4)
    m.even_rows = bool(hints & 8)
    m.device_nnz = bool(hints & 16)
    return m


@torch.library.custom_op("sputnik_b200::sdd", mutates_args=())
def sdd_op(a: torch.Tensor, ta: bool, b: torch.Tensor, tb: bool, rows: int, cols: int,
           offsets: torch.Tensor, indices: torch.Tensor, row_indices: torch.Tensor,
           offsets_t: torch.Tensor, indices_t: torch.Tensor, block_offsets: torch.Tensor,
           hints: int) -> torch.Tensor:
    # capacity-sized topologies (DEVICE_NNZ): blocks past the live count are never written —
    # zero them so that reductions over the values (or their gradient) stay finite
    alloc = torch.zeros if hints & 16 else torch.empty
    out = alloc((indices.numel(), BLOCK, BLOCK), dtype=a.dtype, device=a.device)
    meta = (offsets, indices, row_indices, offsets_t, indices_t, block_offsets)
    with _check_operands((a, b), meta):
        sb.matmul_ex(_dense(a), ta, _dense(b), tb, _matrix_of(rows, cols, out, meta, hints))
    return out


@torch.library.custom_op("sputnik_b200::dsd", mutates_args=())
def dsd_op(values: torch.Tensor, ta: bool, b: torch.Tensor, tb: bool, rows: int, cols: int,
           offsets: torch.Tensor, indices: torch.Tensor, row_indices: torch.Tensor,
           offsets_t: torch.Tensor, indices_t: torch.Tensor, block_offsets: torch.Tensor,
           hints: int) -> torch.Tensor:
    m = cols if ta else rows
    n = b.shape[0] if tb else b.shape[1]
    out = torch.empty((m, n), dtype=b.dtype, device=b.device)
    meta = (offsets, indices, row_indices, offsets_t, indices_t, block_offsets)
    with _check_operands((values, b), meta):
        sb.matmul_ex(_matrix_of(rows, cols, values.contiguous(), meta, hints), ta, _dense(b), tb,
                     sb.Matrix.of(out))
    return out


@torch.library.custom_op("sputnik_b200::dds", mutates_args=())
def dds_op(a: torch.Tensor, ta: bool, values: torch.Tensor, tb: bool, rows: int, cols: int,
           offsets: torch.Tensor, indices: torch.Tensor, row_indices: torch.Tensor,
           offsets_t: torch.Tensor, indices_t: torch.Tensor, block_offsets: torch.Tensor,
           hints: int) -> torch.Tensor:
    m = a.shape[1] if ta else a.shape[0]
    n = rows if tb else cols
    out = torch.empty((m, n), dtype=a.dtype, device=a.device)
    meta = (offsets, indices, row_indices, offsets_t, indices_t, block_offsets)
    with _check_operands((a, values), meta):
        sb.matmul_ex(_dense(a), ta, _matrix_of(rows, cols, values.contiguous(), meta, hints), tb,
                     sb.Matrix.of(out))
    return out


@sdd_op.register_fake
def _(a, ta, b, tb, rows, cols, offsets, indices, row_indices, offsets_t, indices_t,
      block_offsets, hints):
    return a.new_empty((indices.shape[0], BLOCK, BLOCK))


@dsd_op.register_fake
def _(values, ta, b, tb, rows, cols, offsets, indices, row_indices, offsets_t, indices_t,
      block_offsets, hints):
    return b.new_empty((cols if ta else rows, b.shape[0] if tb else b.shape[1]))


@dds_op.register_fake
def _(a, ta, values, tb, rows, cols, offsets, indices, row_indices, offsets_t, indices_t,
      block_offsets, hints):
    return a.new_empty((a.shape[1] if ta else a.shape[0], rows if tb else cols))


def _ctx(ctx, inputs, output):
    x, tx, y, ty, rows, cols = inputs[:6]
    ctx.save_for_backward(x, y, *inputs[6:12])
    ctx.flags = (tx, ty, rows, cols, inputs[12])


def _unpack(ctx):
    x, y, *meta = ctx.saved_tensors
    tx, ty, rows, cols, hints = ctx.flags
    return x, tx, y, ty, (rows, cols, *meta, hints)


def _sdd_backward(ctx, dc):
    a, ta, b, tb, topo = _unpack(ctx)
    dc = dc.contiguous()
    da = db = None
    if ctx.needs_input_grad[0]:
        da = dds_op(b, tb, dc, True, *topo) if ta else dsd_op(dc, False, b, not tb, *topo)
    if ctx.needs_input_grad[2]:
        db = dsd_op(dc, True, a, ta, *topo) if tb else dds_op(a, not ta, dc, False, *topo)
    return (da, None, db) + (None,) * 10


def _dsd_backward(ctx, dc):
    s, ta, b, tb, topo = _unpack(ctx)
    dc = dc.contiguous()
    ds = db = None
    if ctx.needs_input_grad[0]:
        ds = sdd_op(b, tb, dc, True, *topo) if ta else sdd_op(dc, False, b, not tb, *topo)
    if ctx.needs_input_grad[2]:
        db = dds_op(dc, True, s, ta, *topo) if tb else dsd_op(s, not ta, dc, False, *topo)
    return (ds, None, db) + (None,) * 10


def _dds_backward(ctx, dc):
    a, ta, s, tb, topo = _unpack(ctx)
    dc = dc.contiguous()
    da = ds = None
    if ctx.needs_input_grad[0]:
        da = dsd_op(s, tb, dc, True, *topo) if ta else dds_op(dc, False, s, not tb, *topo)
    if ctx.needs_input_grad[2]:
        ds = sdd_op(dc, True, a, ta, *topo) if tb else sdd_op(a, not ta, dc, False, *topo)
    return (da, None, ds) + (None,) * 10


sdd_op.register_autograd(_sdd_backward, setup_context=_ctx)
dsd_op.register_autograd(_dsd_backward, setup_context=_ctx)
dds_op.register_autograd(_dds_backward, setup_context=_ctx)


def _topo_args(topo):
    h = topo.hints
    word = (2 if h.get("twin_rows") else 0) | (4 if h.get("twin_cols") else 0) | \
        (8 if h.get("even_rows") else 0) | (16 if h.get("device_nnz") else 0)
    return (topo.rows, topo.cols, topo.offsets, topo.indices, topo.row_indices, topo.offsets_t,
            topo.indices_t, topo.block_offsets, word)


def sdd(a, b, topo, transpose_a=False, transpose_b=False):
    """values[nnz_blocks,128,128] = (op(a) @ op(b)) sampled at `topo` (Matmul(Matrix, bool, Matrix,
    bool, BlockMatrix), sputnik/block/sdd/sdd.h:10-15).  op(a): [topo.rows, k], op(b): [k, topo.cols]."""
    return sdd_op(a, transpose_a, b, transpose_b, *_topo_args(topo))


def dsd(values, topo, b, transpose_a=False, transpose_b=False):
    """op(S) @ op(b) with S = (values, topo) (Matmul(BlockMatrix, bool, Matrix, bool, Matrix),
    sputnik/block/dsd/dsd.h:10-22)."""
    return dsd_op(values, transpose_a, b, transpose_b, *_topo_args(topo))


def dds(a, values, topo, transpose_a=False, transpose_b=False):
    """op(a) @ op(S) (Matmul(Matrix, bool, BlockMatrix, bool, Matrix), sputnik/block/dds/dds.h:10-22)."""
    return dds_op(a, transpose_a, values, transpose_b, *_topo_args(topo))


# Plain (non-differentiable) calls used by tests and by callers that manage their own buffers.
def _sdd(a, ta, b, tb, topo):
    return sdd_op(a, ta, b, tb, *_topo_args(topo))


def _dsd(values, topo, ta, b, tb):
    return dsd_op(values, ta, b, tb, *_topo_args(topo))


def _dds(a, ta, values, topo, tb):
    return dds_op(a, ta, values, tb, *_topo_args(topo))


# ------------------------------------------------------------------- siblings: SSD / SDS / DSS
#
# The products with a sparse operand AND a sparse output or a second sparse operand
# (sputnik/sputnik.h:21-23) as operators with autograd.  Each takes TWO
# topologies (9 schema arguments each: rows, cols, the six metadata tensors, the hint word), so the
# schemas are written out instead of inferred.  The gradients close over the same three products:
#
#   SSD  C|tc = op(A|ta_) op(B)     dA|ta_ = tA ? sds(B, tB, dC, T) : ssd(dC, N, B, !tB)
#                                   dB     = tB ? dss(dC, T, A, tA) : dss(A, !tA, dC, N)
#   SDS  C|tc = op(A) op(B|tb_)     dA     = tA ? dss(B, tB, dC, T) : dss(dC, N, B, !tB)
#                                   dB|tb_ = tB ? ssd(dC, T, A, tA) : sds(A, !tA, dC, N)
#   DSS  C    = op(A|ta_) op(B|tb_) dA|ta_ = tA ? ssd(B, tB, dC, T) : sds(dC, N, B, !tB)
#                                   dB|tb_ = tB ? sds(dC, T, A, tA) : ssd(A, !tA, dC, N)
#
# (X|t: X sampled at / stored with topology t.)  No topology hints are consumed by these kernels.

_TOPO_SCHEMA = ("int {p}rows, int {p}cols, Tensor {p}offsets, Tensor {p}indices, Tensor {p}row_indices, "
                "Tensor {p}offsets_t, Tensor {p}indices_t, Tensor {p}block_offsets, int {p}hints")


def _sib_schema(p1, p2):
    return (f"(Tensor x, bool tx, Tensor y, bool ty, {_TOPO_SCHEMA.format(p=p1)}, "
            f"{_TOPO_SCHEMA.format(p=p2)}) -> Tensor")


def _split(t):
    """18 flat topology arguments -> ((rows, cols, meta6, hints), (rows, cols, meta6, hints))."""
    return (t[0], t[1], tuple(t[2:8]), t[8]), (t[9], t[10], tuple(t[11:17]), t[17])


def _sparse(topo, values):
    rows, cols, meta, hints = topo
    return _matrix_of(rows, cols, values, meta, hints)


def _blocks_of(topo, like):
    return torch.empty((topo[2][1].numel(), BLOCK, BLOCK), dtype=like.dtype, device=like.device)


@torch.library.custom_op("sputnik_b200::ssd", mutates_args=(), schema=_sib_schema("a_", "c_"))
def ssd_op(x, tx, y, ty, *t):
    ta_, tc = _split(t)
    out = _blocks_of(tc, y)
    with _check_operands((x, y), ta_[2]), _check_operands((x, y), tc[2]):
        sb.matmul_ex(_sparse(ta_, x.contiguous()), tx, _dense(y), ty, _sparse(tc, out))
    return out


@torch.library.custom_op("sputnik_b200::sds", mutates_args=(), schema=_sib_schema("b_", "c_"))
def sds_op(x, tx, y, ty, *t):
    tb_, tc = _split(t)
    out = _blocks_of(tc, x)
    with _check_operands((x, y), tb_[2]), _check_operands((x, y), tc[2]):
        sb.matmul_ex(_dense(x), tx, _sparse(tb_, y.contiguous()), ty, _sparse(tc, out))
    return out


@torch.library.custom_op("sputnik_b200::dss", mutates_args=(), schema=_sib_schema("a_", "b_"))
def dss_op(x, tx, y, ty, *t):
    ta_, tb_ = _split(t)
    m = ta_[1] if tx else ta_[0]
    n = tb_[0] if ty else tb_[1]
    out = torch.empty((m, n), dtype=x.dtype, device=x.device)
    with _check_operands((x, y), ta_[2]), _check_operands((x, y), tb_[2]):
        sb.matmul_ex(_sparse(ta_, x.contiguous()), tx, _sparse(tb_, y.contiguous()), ty,
                     sb.Matrix.of(out))
    return out


@ssd_op.register_fake
def _(x, tx, y, ty, *t):
    return y.new_empty((t[12].shape[0], BLOCK, BLOCK))


@sds_op.register_fake
def _(x, tx, y, ty, *t):
    return x.new_empty((t[12].shape[0], BLOCK, BLOCK))


@dss_op.register_fake
def _(x, tx, y, ty, *t):
    return x.new_empty((t[1] if tx else t[0], t[9] if ty else t[10]))


def _sib_ctx(ctx, inputs, output):
    x, tx, y, ty = inputs[:4]
    t = inputs[4:]
    ctx.save_for_backward(x, y, *t[2:8], *t[11:17])
    ctx.flags = (tx, ty, t[0], t[1], t[8], t[9], t[10], t[17])


def _sib_unpack(ctx):
    x, y, *meta = ctx.saved_tensors
    tx, ty, r1, c1, h1, r2, c2, h2 = ctx.flags
    return x, tx, y, ty, (r1, c1, *meta[:6], h1), (r2, c2, *meta[6:], h2)


_NONE18 = (None,) * 18


def _ssd_backward(ctx, dc):
    a, ta, b, tb, t_a, t_c = _sib_unpack(ctx)
    dc = dc.contiguous()
    da = db = None
    if ctx.needs_input_grad[0]:
        da = sds_op(b, tb, dc, True, *t_c, *t_a) if ta else ssd_op(dc, False, b, not tb, *t_c, *t_a)
    if ctx.needs_input_grad[2]:
        db = dss_op(dc, True, a, ta, *t_c, *t_a) if tb else dss_op(a, not ta, dc, False, *t_a, *t_c)
    return (da, None, db, None) + _NONE18


def _sds_backward(ctx, dc):
    a, ta, b, tb, t_b, t_c = _sib_unpack(ctx)
    dc = dc.contiguous()
    da = db = None
    if ctx.needs_input_grad[0]:
        da = dss_op(b, tb, dc, True, *t_b, *t_c) if ta else dss_op(dc, False, b, not tb, *t_c, *t_b)
    if ctx.needs_input_grad[2]:
        db = ssd_op(dc, True, a, ta, *t_c, *t_b) if tb else sds_op(a, not ta, dc, False, *t_c, *t_b)
    return (da, None, db, None) + _NONE18


def _dss_backward(ctx, dc):
    a, ta, b, tb, t_a, t_b = _sib_unpack(ctx)
    dc = dc.contiguous()
    da = db = None
    if ctx.needs_input_grad[0]:
        da = ssd_op(b, tb, dc, True, *t_b, *t_a) if ta else sds_op(dc, False, b, not tb, *t_b, *t_a)
    if ctx.needs_input_grad[2]:
        db = sds_op(dc, True, a, ta, *t_a, *t_b) if tb else ssd_op(a, not ta, dc, False, *t_a, *t_b)
    return (da, None, db, None) + _NONE18


ssd_op.register_autograd(_ssd_backward, setup_context=_sib_ctx)
sds_op.register_autograd(_sds_backward, setup_context=_sib_ctx)
dss_op.register_autograd(_dss_backward, setup_context=_sib_ctx)


def ssd(values_a, topo_a, b, topo_c, transpose_a=False, transpose_b=False):
    """values_c = (op(A) @ op(b)) sampled at `topo_c`, A = (values_a, topo_a)
    (Matmul(BlockMatrix, bool, Matrix, bool, BlockMatrix), sputnik/block/ssd/ssd.h:10-22)."""
    return ssd_op(values_a, transpose_a, b, transpose_b, *_topo_args(topo_a), *_topo_args(topo_c))


def sds(a, values_b, topo_b, topo_c, transpose_a=False, transpose_b=False):
    """values_c = (op(a) @ op(B)) sampled at `topo_c`, B = (values_b, topo_b)
    (Matmul(Matrix, bool, BlockMatrix, bool, BlockMatrix), sputnik/block/sds/sds.h:10-22)."""
    return sds_op(a, transpose_a, values_b, transpose_b, *_topo_args(topo_b), *_topo_args(topo_c))


def dss(values_a, topo_a, values_b, topo_b, transpose_a=False, transpose_b=False):
    """op(A) @ op(B), both sparse, dense result
    (Matmul(BlockMatrix, bool, BlockMatrix, bool, Matrix), sputnik/block/dss/dss.h:10-22)."""
    return dss_op(values_a, transpose_a, values_b, transpose_b, *_topo_args(topo_a),
                  *_topo_args(topo_b))
