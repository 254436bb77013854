"""The PyTorch operator binding (sputnik_b200/ops.py): sdd / dsd / dds forward and backward
against dense fp32 autograd on the same 16-bit operands, on a random topology (1-SM kernels) and
on a dMoE topology (2-SM kernels)."""
import numpy as np
import pytest
import torch

from sputnik_b200 import ops, synth

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _rel(got, ref):
    got, ref = got.detach(), ref.detach()
    ref = ref.double()
    tau = 1e-2 * ref.pow(2).mean().sqrt()
    return float(((got.double() - ref).abs() / torch.maximum(ref.abs(), tau)).max())


def _mask(topo):
    m = torch.zeros((topo.block_rows, topo.block_cols), dtype=torch.bool)
    rows = topo.row_indices().astype(np.int64)
    m[torch.from_numpy(rows), torch.from_numpy(topo.indices.astype(np.int64))] = True
    return m.to(DEV).repeat_interleave(128, 0).repeat_interleave(128, 1)


def _to_blocks(dense, topo):
    rows = topo.row_indices().astype(np.int64)
    cols = topo.indices.astype(np.int64)
    d = dense.reshape(topo.block_rows, 128, topo.block_cols, 128).permute(0, 2, 1, 3)
    return d[torch.from_numpy(rows).to(DEV), torch.from_numpy(cols).to(DEV)]


def _to_dense(values, topo):
    out = torch.zeros((topo.block_rows, topo.block_cols, 128, 128), device=DEV, dtype=values.dtype)
    rows = topo.row_indices().astype(np.int64)
    cols = topo.indices.astype(np.int64)
    out[torch.from_numpy(rows).to(DEV), torch.from_numpy(cols).to(DEV)] = values
    return out.permute(0, 2, 1, 3).reshape(topo.rows, topo.cols)


def _rand(shape, seed):
    g = torch.Generator(device=DEV).manual_seed(seed)
    return torch.empty(shape, device=DEV).uniform_(-1, 1, generator=g).bfloat16()


TOPOS = {"random": lambda: synth.random_topology(5, 6, 13, seed=3, unordered=True),
         "moe": lambda: synth.moe_topology(3, 4, 2)}


@pytest.mark.parametrize("kind", list(TOPOS))
def test_sdd_forward_backward(kind):
    host = TOPOS[kind]()
    topo = ops.Topology.from_host(host, DEV)
    a, b = _rand((host.rows, 264), 1).requires_grad_(), _rand((264, host.cols), 2).requires_grad_()
    dc = _rand((host.nnz_blocks, 128, 128), 3)
    out = ops.sdd(a, b, topo)
    out.backward(dc)
    a32, b32 = a.detach().float().requires_grad_(), b.detach().float().requires_grad_()
    ref = (a32 @ b32) * _mask(host)
    ref.backward(_to_dense(dc.float(), host))
    assert _rel(out.float(), _to_blocks(ref.detach(), host)) <= 1e-2
    assert _rel(a.grad.float(), a32.grad) <= 1e-2
    assert _rel(b.grad.float(), b32.grad) <= 1e-2


@pytest.mark.parametrize("kind", list(TOPOS))
def test_dsd_forward_backward(kind):
    host = TOPOS[kind]()
    topo = ops.Topology.from_host(host, DEV)
    v = _rand((host.nnz_blocks, 128, 128), 4).requires_grad_()
    b = _rand((host.cols, 264), 5).requires_grad_()
    dc = _rand((host.rows, 264), 6)
    out = ops.dsd(v, topo, b)
    out.backward(dc)
    s32 = _to_dense(v.detach().float(), host).requires_grad_()
    b32 = b.detach().float().requires_grad_()
    ref = s32 @ b32
    ref.backward(dc.float())
    assert _rel(out.float(), ref.detach()) <= 1e-2
    assert _rel(v.grad.float(), _to_blocks(s32.grad, host)) <= 1e-2
    assert _rel(b.grad.float(), b32.grad) <= 1e-2


@pytest.mark.parametrize("kind", list(TOPOS))
def test_dds_forward_backward(kind):
    host = TOPOS[kind]()
    topo = ops.Topology.from_host(host, DEV)
    a = _rand((264, host.rows), 7).requires_grad_()
    v = _rand((host.nnz_blocks, 128, 128), 8).requires_grad_()
    dc = _rand((264, host.cols), 9)
    out = ops.dds(a, v, topo)
    out.backward(dc)
    a32 = a.detach().float().requires_grad_()
    s32 = _to_dense(v.detach().float(), host).requires_grad_()
    ref = a32 @ s32
    ref.backward(dc.float())
    assert _rel(out.float(), ref.detach()) <= 1e-2
    assert _rel(a.grad.float(), a32.grad) <= 1e-2
    assert _rel(v.grad.float(), _to_blocks(s32.grad, host)) <= 1e-2


def test_dmoe_ffn_as_autograd_graph():
    """h = sdd(x, w1); y = dsd(h, w2) as a differentiable layer: gradients of all three leaves
    against dense fp32 autograd of the same block-diagonal computation."""
    host = synth.moe_topology(2, 2, 2)
    topo = ops.Topology.from_host(host, DEV)
    x = _rand((host.rows, 136), 10).requires_grad_()
    w1 = (_rand((136, host.cols), 11) * 0.1).bfloat16().requires_grad_()
    w2 = (_rand((host.cols, 136), 12) * 0.1).bfloat16().requires_grad_()
    dy = _rand((host.rows, 136), 13)
    y = ops.dsd(ops.sdd(x, w1, topo), topo, w2)
    y.backward(dy)
    x32, w132, w232 = (t.detach().float().requires_grad_() for t in (x, w1, w2))
    h = ((x32 @ w132) * _mask(host))
    yref = h @ w232
    yref.backward(dy.float())
    # h (and dh) are rounded to bf16 between two products on the device while the fp32 graph
    # keeps them exact, so the chain is compared in the rms norm: max |diff| <= 2 % of rms(ref)
    # (each product on its own is held to the 1e-2 element-wise bar by the tests above).
    def chain_err(got, ref):
        ref = ref.detach().double()
        return float((got.detach().double() - ref).abs().max() / ref.pow(2).mean().sqrt())
    assert chain_err(y, yref) <= 2e-2
    assert chain_err(x.grad, x32.grad) <= 2e-2
    assert chain_err(w1.grad, w132.grad) <= 2e-2
    assert chain_err(w2.grad, w232.grad) <= 2e-2


def _near_twin(seed):
    """A dMoE topology with one index of one row changed: twin_rows must drop, even_rows stay."""
    t = synth.moe_topology(4, 2, 4)
    idx = t.indices.copy()
    idx[5] = (idx[5] + 4) % t.block_cols
    # keep the row sorted-free (unordered rows are legal BCSR here) but still unique per row
    return synth.Topology(t.block_rows, t.block_cols, t.offsets, idx)


ANALYZE = {"moe_even": lambda: synth.moe_topology(3, 4, 2),
           "moe_odd_cols": lambda: synth.moe_topology(3, 2, 3),
           "moe_odd_rows": lambda: synth.moe_topology(3, 3, 2),
           "moe_one_expert": lambda: synth.moe_topology(1, 2, 2),
           "random": lambda: synth.random_topology(5, 6, 13, seed=3, unordered=True),
           "random_big": lambda: synth.random_topology(64, 48, 700, seed=5),
           "near_twin": lambda: _near_twin(0),
           "dense": lambda: synth.random_topology(4, 4, 16, seed=1)}


@pytest.mark.parametrize("kind", list(ANALYZE))
def test_device_analysis_matches_host_hints(kind):
    """bsmm_analyze_topology (device kernels) == synth.Topology.hints() (numpy), flag for flag."""
    host = ANALYZE[kind]()
    topo = ops.Topology(host.rows, host.cols, torch.from_numpy(host.offsets),
                        torch.from_numpy(host.indices), DEV)
    assert topo.hints == host.hints(), (topo.hints, host.hints())


def test_ops_with_device_derived_hints_use_pair_kernels():
    """A Topology built from device tensors (no host hints) reaches the 2-SM kernels and gives the
    same bits as the host-hinted one."""
    host = synth.moe_topology(4, 4, 4)
    t_dev = ops.Topology(host.rows, host.cols, torch.from_numpy(host.offsets).to(DEV),
                         torch.from_numpy(host.indices).to(DEV))
    t_host = ops.Topology.from_host(host, DEV)
    assert t_dev.hints == t_host.hints and t_dev.hints["twin_rows"]
    a, b = _rand((host.rows, 256), 21), _rand((256, host.cols), 22)
    h1, h2 = ops.sdd(a, b, t_dev), ops.sdd(a, b, t_host)
    assert torch.equal(h1, h2)
    w = _rand((host.cols, 256), 23)
    assert torch.equal(ops.dsd(h1, t_dev, w), ops.dsd(h2, t_host, w))


@pytest.mark.parametrize("ta,tb", [(0, 0), (0, 1), (1, 0), (1, 1)])
@pytest.mark.parametrize("op", ["sdd", "dsd", "dds"])
@pytest.mark.parametrize("kind", ["random", "moe"])
def test_custom_ops_every_transpose_pair_forward_backward(kind, op, ta, tb):
    """torch.ops.sputnik_b200.{sdd,dsd,dds} with transpose flags: forward and both gradients
    against dense fp32 autograd of the same stored operands."""
    host = TOPOS[kind]()
    topo = ops.Topology.from_host(host, DEV)
    R, C, D = host.rows, host.cols, 264
    T = lambda x, t: x.t() if t else x
    mask = _mask(host)
    if op == "sdd":      # op(a): [R, D], op(b): [D, C]
        a = _rand((D, R) if ta else (R, D), 31).requires_grad_()
        b = _rand((C, D) if tb else (D, C), 32).requires_grad_()
        out = ops.sdd(a, b, topo, ta, tb)
        dc = _rand(tuple(out.shape), 33)
        out.backward(dc)
        a32, b32 = a.detach().float().requires_grad_(), b.detach().float().requires_grad_()
        ref = (T(a32, ta) @ T(b32, tb)) * mask
        ref.backward(_to_dense(dc.float(), host))
        assert _rel(out.float(), _to_blocks(ref.detach(), host)) <= 1e-2
        assert _rel(a.grad.float(), a32.grad) <= 1e-2 and _rel(b.grad.float(), b32.grad) <= 1e-2
    elif op == "dsd":    # op(S): [R, C] or [C, R]; op(b): [inner, D]
        inner = R if ta else C
        v = _rand((host.nnz_blocks, 128, 128), 34).requires_grad_()
        b = _rand((D, inner) if tb else (inner, D), 35).requires_grad_()
        out = ops.dsd(v, topo, b, ta, tb)
        dc = _rand(tuple(out.shape), 36)
        out.backward(dc)
        s32 = _to_dense(v.detach().float(), host).requires_grad_()
        b32 = b.detach().float().requires_grad_()
        ref = T(s32, ta) @ T(b32, tb)
        ref.backward(dc.float())
        assert _rel(out.float(), ref.detach()) <= 1e-2
        assert _rel(v.grad.float(), _to_blocks(s32.grad, host)) <= 1e-2
        assert _rel(b.grad.float(), b32.grad) <= 1e-2
    else:                # op(a): [D, inner]; op(S): [R, C] or [C, R]
        inner = C if tb else R
        a = _rand((inner, D) if ta else (D, inner), 37).requires_grad_()
        v = _rand((host.nnz_blocks, 128, 128), 38).requires_grad_()
        out = ops.dds(a, v, topo, ta, tb)
        dc = _rand(tuple(out.shape), 39)
        out.backward(dc)
        a32 = a.detach().float().requires_grad_()
        s32 = _to_dense(v.detach().float(), host).requires_grad_()
        ref = T(a32, ta) @ T(s32, tb)
        ref.backward(dc.float())
        assert _rel(out.float(), ref.detach()) <= 1e-2
        assert _rel(a.grad.float(), a32.grad) <= 1e-2
        assert _rel(v.grad.float(), _to_blocks(s32.grad, host)) <= 1e-2
