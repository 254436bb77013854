"""PyTorch operator binding of the block-sparse products: `sdd`, `dsd`, `dds`
as differentiable functions over a device-resident `Topology`, so the path can be used as a
layer.  torch owns memory, streams and the autograd graph; every product — forward and backward —
is one call of the C-ABI (libbsmm.so), in the pattern MegaBlocks' dMoE uses:

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
        self.indices = torch.as_tensor(indices, dtype=torch.int16).to(dev).contiguous()
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


def _sdd(a, ta, b, tb, topo):
    out = topo.new_values(a.dtype)
    sb.matmul_ex(_dense(a), ta, _dense(b), tb, topo.matrix(out))
    return out


def _dsd(values, topo, ta, b, tb):
    m = topo.cols if ta else topo.rows
    n = b.shape[0] if tb else b.shape[1]
    out = torch.empty((m, n), dtype=b.dtype, device=b.device)
    sb.matmul_ex(topo.matrix(values.contiguous()), ta, _dense(b), tb, sb.Matrix.of(out))
    return out


def _dds(a, ta, values, topo, tb):
    m = a.shape[1] if ta else a.shape[0]
    n = topo.rows if tb else topo.cols
    out = torch.empty((m, n), dtype=a.dtype, device=a.device)
    sb.matmul_ex(_dense(a), ta, topo.matrix(values.contiguous()), tb, sb.Matrix.of(out))
    return out


class _Sdd(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, b, topo):
        ctx.save_for_backward(a, b)
        ctx.topo = topo
        return _sdd(a, False, b, False, topo)

    @staticmethod
    def backward(ctx, dc):
        a, b = ctx.saved_tensors
        da = _dsd(dc, ctx.topo, False, b, True) if ctx.needs_input_grad[0] else None   # dC B^T
        db = _dds(a, True, dc, ctx.topo, False) if ctx.needs_input_grad[1] else None   # A^T dC
        return da, db, None


class _Dsd(torch.autograd.Function):
    @staticmethod
    def forward(ctx, values, topo, b):
        ctx.save_for_backward(values, b)
        ctx.topo = topo
        return _dsd(values, topo, False, b, False)

    @staticmethod
    def backward(ctx, dc):
        values, b = ctx.saved_tensors
        dv = _sdd(dc, False, b, True, ctx.topo) if ctx.needs_input_grad[0] else None       # dC B^T | topo
        db = _dsd(values, ctx.topo, True, dc, False) if ctx.needs_input_grad[2] else None  # S^T dC
        return dv, None, db


class _Dds(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, values, topo):
        ctx.save_for_backward(a, values)
        ctx.topo = topo
        return _dds(a, False, values, topo, False)

    @staticmethod
    def backward(ctx, dc):
        a, values = ctx.saved_tensors
        da = _dds(dc, False, values, ctx.topo, True) if ctx.needs_input_grad[0] else None  # dC S^T
        dv = _sdd(a, True, dc, False, ctx.topo) if ctx.needs_input_grad[1] else None       # A^T dC | topo
        return da, dv, None


def sdd(a, b, topo):
    """values[nnz_blocks,128,128] = (a @ b) sampled at `topo` (Matmul(Matrix, Matrix, BlockMatrix),
    sputnik/block/sdd/sdd.h:10-15).  a: [topo.rows, k], b: [k, topo.cols]."""
    return _Sdd.apply(a, b, topo)


def dsd(values, topo, b):
    """[topo.rows, n] = S @ b with S = (values, topo) (Matmul(BlockMatrix, Matrix, Matrix),
    sputnik/block/dsd/dsd.h:10-22).  b: [topo.cols, n]."""
    return _Dsd.apply(values, topo, b)


def dds(a, values, topo):
    """[m, topo.cols] = a @ S (Matmul(Matrix, BlockMatrix, Matrix), sputnik/block/dds/dds.h:10-22).
    a: [m, topo.rows]."""
    return _Dds.apply(a, values, topo)
