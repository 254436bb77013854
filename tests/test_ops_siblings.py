"""The sibling operators of sputnik_b200/ops.py — ssd / sds / dss (sparse operand with a sparse
output, or two sparse operands) — forward and backward for every transpose pair against dense
fp32 autograd on the same 16-bit operands.  Every gradient is again one of the three products
(the table in ops.py), so this also exercises each kernel with each transpose flag."""
import numpy as np
import pytest
import torch

from sputnik_b200 import ops, synth
from test_ops import _rand, _rel, _to_blocks, _to_dense

gpu = pytest.mark.gpu
DEV = "cuda:0"
PAIRS = [(False, False), (True, False), (False, True), (True, True)]


def _mask(topo):
    m = torch.zeros((topo.block_rows, topo.block_cols), dtype=torch.bool)
    m[torch.from_numpy(topo.row_indices().astype(np.int64)),
      torch.from_numpy(topo.indices.astype(np.int64))] = True
    return m.to(DEV).repeat_interleave(128, 0).repeat_interleave(128, 1).float()


def _op(x, t):
    return x.t() if t else x


def _topos(m, k, n, ta, tb, seed):
    """Host topologies for op(A) [m, k] and op(B) [k, n] as stored, and for C [m, n] (blocks)."""
    sa = (k, m) if ta else (m, k)
    sb_ = (n, k) if tb else (k, n)
    mk = lambda shape, s: synth.random_topology(shape[0], shape[1], max(2, (shape[0] * shape[1] * 2) // 3),
                                                seed=seed + s, unordered=True)
    return mk(sa, 1), mk(sb_, 2), mk((m, n), 3)


@gpu
@pytest.mark.parametrize("ta,tb", PAIRS)
def test_ssd_forward_backward(ta, tb):
    m, k, n = 3, 4, 2                                    # in blocks
    ha, _, hc = _topos(m, k, n, ta, tb, 10)
    t_a, t_c = ops.Topology.from_host(ha, DEV), ops.Topology.from_host(hc, DEV)
    va = _rand((ha.nnz_blocks, 128, 128), 1).requires_grad_()
    b = _rand((n * 128, k * 128) if tb else (k * 128, n * 128), 2).requires_grad_()
    dc = _rand((hc.nnz_blocks, 128, 128), 3)
    out = ops.ssd(va, t_a, b, t_c, ta, tb)
    out.backward(dc)
    a32 = _to_dense(va.detach().float(), ha).requires_grad_()
    b32 = b.detach().float().requires_grad_()
    ref = (_op(a32, ta) @ _op(b32, tb)) * _mask(hc)
    ref.backward(_to_dense(dc.float(), hc))
    assert _rel(out.float(), _to_blocks(ref.detach(), hc)) <= 1e-2
    assert _rel(va.grad.float(), _to_blocks(a32.grad, ha)) <= 1e-2
    assert _rel(b.grad.float(), b32.grad) <= 1e-2


@gpu
@pytest.mark.parametrize("ta,tb", PAIRS)
def test_sds_forward_backward(ta, tb):
    m, k, n = 2, 4, 3
    _, hb, hc = _topos(m, k, n, ta, tb, 20)
    t_b, t_c = ops.Topology.from_host(hb, DEV), ops.Topology.from_host(hc, DEV)
    a = _rand((k * 128, m * 128) if ta else (m * 128, k * 128), 4).requires_grad_()
    vb = _rand((hb.nnz_blocks, 128, 128), 5).requires_grad_()
    dc = _rand((hc.nnz_blocks, 128, 128), 6)
    out = ops.sds(a, vb, t_b, t_c, ta, tb)
    out.backward(dc)
    a32 = a.detach().float().requires_grad_()
    b32 = _to_dense(vb.detach().float(), hb).requires_grad_()
    ref = (_op(a32, ta) @ _op(b32, tb)) * _mask(hc)
    ref.backward(_to_dense(dc.float(), hc))
    assert _rel(out.float(), _to_blocks(ref.detach(), hc)) <= 1e-2
    assert _rel(a.grad.float(), a32.grad) <= 1e-2
    assert _rel(vb.grad.float(), _to_blocks(b32.grad, hb)) <= 1e-2


@gpu
@pytest.mark.parametrize("ta,tb", PAIRS)
def test_dss_forward_backward(ta, tb):
    m, k, n = 3, 4, 3
    ha, hb, _ = _topos(m, k, n, ta, tb, 30)
    t_a, t_b = ops.Topology.from_host(ha, DEV), ops.Topology.from_host(hb, DEV)
    va = _rand((ha.nnz_blocks, 128, 128), 7).requires_grad_()
    vb = _rand((hb.nnz_blocks, 128, 128), 8).requires_grad_()
    dc = _rand((m * 128, n * 128), 9)
    out = ops.dss(va, t_a, vb, t_b, ta, tb)
    out.backward(dc)
    a32 = _to_dense(va.detach().float(), ha).requires_grad_()
    b32 = _to_dense(vb.detach().float(), hb).requires_grad_()
    ref = _op(a32, ta) @ _op(b32, tb)
    ref.backward(dc.float())
    assert _rel(out.float(), ref.detach()) <= 1e-2
    assert _rel(va.grad.float(), _to_blocks(a32.grad, ha)) <= 1e-2
    assert _rel(vb.grad.float(), _to_blocks(b32.grad, hb)) <= 1e-2


def test_sibling_ops_are_registered_with_shape_functions():
    """CPU: the operators exist with fake-tensor kernels (shapes only; no compute without a GPU)."""
    from torch._subclasses.fake_tensor import FakeTensorMode
    with FakeTensorMode():
        meta = lambda nb, br, bc: (br * 128, bc * 128, torch.empty(br + 1, dtype=torch.int32),
                                   torch.empty(nb, dtype=torch.int16), torch.empty(nb, dtype=torch.int16),
                                   torch.empty(bc + 1, dtype=torch.int32), torch.empty(nb, dtype=torch.int16),
                                   torch.empty(nb, dtype=torch.int32), 0)
        va = torch.empty((5, 128, 128), dtype=torch.bfloat16)
        b = torch.empty((512, 256), dtype=torch.bfloat16)
        c = torch.ops.sputnik_b200.ssd(va, False, b, False, *meta(5, 3, 4), *meta(4, 3, 2))
        assert tuple(c.shape) == (4, 128, 128)
        d = torch.ops.sputnik_b200.dss(va, True, va, False, *meta(5, 3, 4), *meta(5, 3, 4))
        assert tuple(d.shape) == (512, 512)
