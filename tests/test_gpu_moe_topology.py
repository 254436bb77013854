"""The on-device dMoE topology builder (bsmm_moe_topology) and the
device-sized topologies it produces (BSMM_FLAG_DEVICE_NNZ):

* metadata bit-exact against a numpy construction whose transposed view / row_indices come from
  the oracle (itself pinned to the reference's Transpose / RowIndices);
* SDD / DSD / DDS on the capacity-sized topology give the same bits as on the exact topology
  (live region), zeros for padded rows, and leave the dead tail of the value array untouched;
* a whole differentiable FFN over ops.Topology.for_moe against the host-built topology.
"""
import numpy as np
import pytest
import torch

import sputnik_b200 as sb
from oracle import oracle
from sputnik_b200 import ops, synth

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def host_moe(tokens, ffn_blocks, align, cap_rows):
    """numpy statement of the builder's contract; returns (live Topology, padded bins)."""
    n = [(max(t, 0) + 128 * align - 1) // (128 * align) * align for t in tokens]
    start = np.minimum(np.concatenate([[0], np.cumsum(n)]), cap_rows)
    bins = np.concatenate([[0], np.cumsum(n)]) * 128
    R = int(start[-1])
    offsets = np.arange(R + 1, dtype=np.int32) * ffn_blocks
    indices = np.zeros(R * ffn_blocks, dtype=np.int16)
    for e in range(len(tokens)):
        for r in range(start[e], start[e + 1]):
            indices[r * ffn_blocks:(r + 1) * ffn_blocks] = e * ffn_blocks + np.arange(ffn_blocks)
    return synth.Topology(R, len(tokens) * ffn_blocks, offsets, indices), bins.astype(np.int32)


CASES = {
    "uniform": ([256, 256, 256, 256], 2, 2, 8),
    "ragged": ([100, 0, 700, 1, 128, 129], 3, 1, 16),
    "ragged_align2": ([100, 0, 700, 1, 128, 129], 4, 2, 24),
    "leading_empty": ([0, 0, 300], 2, 2, 6),
    "all_empty": ([0, 0, 0, 0], 2, 2, 4),
    "exact_fit": ([128, 256, 384], 1, 1, 6),
    "spare_capacity": ([513, 5], 2, 2, 32),
    "truncated": ([600, 600, 600], 2, 1, 8),       # 15 block rows wanted, 8 available
    "many_experts": (list(np.random.default_rng(0).integers(0, 400, 70)), 2, 2, 400),
    "one_expert": ([1000], 5, 2, 8),
}


@pytest.mark.parametrize("name", list(CASES))
def test_builder_bit_exact(name):
    tokens, F, align, cap = CASES[name]
    t = torch.tensor(tokens, dtype=torch.int32, device=DEV)
    m, bins = sb.moe_topology(t, F, cap, align)
    torch.cuda.synchronize()
    host, hbins = host_moe(tokens, F, align, cap)
    R, nnz = host.block_rows, host.nnz_blocks
    off = m.offsets.cpu().numpy()
    assert np.array_equal(off[:R + 1], host.offsets)
    assert np.all(off[R:] == nnz)                                   # empty tail rows
    assert np.array_equal(m.indices.cpu().numpy()[:nnz], host.indices)
    assert np.array_equal(m.row_indices.cpu().numpy()[:nnz], oracle.row_indices(host.offsets, R))
    ot, it, bo = oracle.transpose_metadata(host.offsets, host.indices, R, host.block_cols)
    assert np.array_equal(m.offsets_t.cpu().numpy(), ot)
    assert np.array_equal(m.indices_t.cpu().numpy()[:nnz], it)
    assert np.array_equal(m.block_offsets.cpu().numpy()[:nnz], bo)
    for arr in (m.indices, m.row_indices, m.indices_t, m.block_offsets):
        assert not arr.cpu().numpy()[nnz:].any()                    # zero-filled tails
    assert np.array_equal(bins.cpu().numpy(), hbins)
    if name == "truncated":
        assert int(bins[-1]) > cap * 128                            # how a caller detects it
    # the library's own device kernels on the live topology agree too
    live = sb.BlockMatrix(R * 128, host.cols, 128, nnz * 128 * 128, None,
                          m.offsets[:R + 1].contiguous(), m.indices[:nnz].contiguous())
    sb.allocate_transpose_buffers(live)
    sb.allocate_row_indices_buffer(live)
    if nnz:
        sb.transpose(live)
        sb.row_indices(live)
        assert torch.equal(live.offsets_t, m.offsets_t)
        assert torch.equal(live.indices_t, m.indices_t[:nnz])
        assert torch.equal(live.block_offsets, m.block_offsets[:nnz])
        assert torch.equal(live.row_indices, m.row_indices[:nnz])
        # hints claimed by construction == device analysis of the live topology
        got = sb.analyze(live)
        if m.twin_rows:
            assert got["twin_rows"]
        if m.even_rows:
            assert got["even_rows"] and got["twin_cols"]


def test_builder_rejects_bad_arguments():
    t = torch.zeros(4, dtype=torch.int32, device=DEV)
    with pytest.raises(sb.BsmmError):
        sb.moe_topology(t, 2, 7, 2)          # capacity not a multiple of the row alignment
    with pytest.raises(sb.BsmmError):
        sb.moe_topology(t, 0, 8, 2)
    with pytest.raises(sb.BsmmError):
        sb.moe_topology(t, 2, 8, 3)
    with pytest.raises(sb.BsmmError):
        sb.moe_topology(t, 9000, 8, 2)       # > 32767 block columns


def _rand(shape, seed, dtype=torch.bfloat16):
    g = torch.Generator(device=DEV).manual_seed(seed)
    return torch.empty(shape, device=DEV).uniform_(-1, 1, generator=g).to(dtype)


@pytest.mark.parametrize("align", [1, 2])
@pytest.mark.parametrize("F", [2, 3])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_products_on_device_sized_topology(align, F, dtype):
    """Capacity-sized topology + DEVICE_NNZ == exact topology, product by product.  F = 2 with
    align = 2 takes the 2-SM kernels, the other combinations the 1-SM ones."""
    tokens = [300, 0, 129, 640, 5]
    cap = 24
    t = torch.tensor(tokens, dtype=torch.int32, device=DEV)
    host, _ = host_moe(tokens, F, align, cap)
    R, nnz = host.block_rows, host.nnz_blocks
    big = ops.Topology.for_moe(t, F * 128, cap * 128, 128 * align)
    exact = ops.Topology.from_host(host, DEV)
    assert big.nnz_blocks == cap * F and exact.nnz_blocks == nnz
    hid = 264
    x_live = _rand((R * 128, hid), 1, dtype)
    x = torch.cat([x_live, _rand(((cap - R) * 128, hid), 2, dtype)])      # padding rows: junk
    w1, w2 = _rand((hid, host.cols), 3, dtype), _rand((host.cols, hid), 4, dtype)

    # SDD: live blocks equal, dead tail untouched
    sentinel = 77.0
    h_big = torch.full((big.nnz_blocks, 128, 128), sentinel, device=DEV, dtype=dtype)
    sb.matmul_ex(sb.Matrix.of(x), False, sb.Matrix.of(w1), False, big.matrix(h_big))
    h = ops.sdd(x_live, w1, exact)
    assert torch.equal(h_big[:nnz], h)
    assert bool((h_big[nnz:] == sentinel).all())
    # SDD nt / tn variants (the backward products dh = dy w2^T and the weight-gradient shape)
    h_nt = torch.full_like(h_big, sentinel)
    sb.matmul_ex(sb.Matrix.of(x), False, sb.Matrix.of(w2), True, big.matrix(h_nt))
    ref_nt = exact.new_values(dtype)
    sb.matmul_ex(sb.Matrix.of(x_live), False, sb.Matrix.of(w2), True, exact.matrix(ref_nt))
    assert torch.equal(h_nt[:nnz], ref_nt) and bool((h_nt[nnz:] == sentinel).all())

    # DSD nn: live rows equal, padded rows zero; tn: S^T dy over the capacity K extent
    y_big = ops.dsd(h_big, big, w2)
    y = ops.dsd(h, exact, w2)
    assert torch.equal(y_big[:R * 128], y)
    assert not y_big[R * 128:].any()
    dy = _rand((cap * 128, hid), 5, dtype)
    g_big = ops._dsd(h_big, big, True, dy, False)
    g = ops._dsd(h, exact, True, dy[:R * 128].contiguous(), False)
    assert torch.equal(g_big, g)

    # DDS nn (x^T-shaped lhs times S) and nt (dy S^T: output has capacity columns, dead ones zero)
    lhs = _rand((hid, cap * 128), 6, dtype)
    d_big = ops.dds(lhs, h_big, big)
    d = ops.dds(lhs[:, :R * 128].contiguous(), h, exact)
    assert torch.equal(d_big, d)
    lhs2 = _rand((hid, host.cols), 7, dtype)
    e_big = ops._dds(lhs2, False, h_big, big, True)
    e = ops._dds(lhs2, False, h, exact, True)
    assert torch.equal(e_big[:, :R * 128], e)
    assert not e_big[:, R * 128:].any()


def test_sdd_device_nnz_many_blocks():
    """More live blocks than SMs (dynamic scheduler + clamp) and an odd live block count."""
    tokens = [128 * 37, 128 * 20]
    F, cap = 3, 128   # 384 capacity blocks = 192 tiles > 148 SMs: the dynamic scheduler runs
    t = torch.tensor(tokens, dtype=torch.int32, device=DEV)
    host, _ = host_moe(tokens, F, 1, cap)
    big = ops.Topology.for_moe(t, F * 128, cap * 128, 128)
    exact = ops.Topology.from_host(host, DEV)
    assert host.nnz_blocks == 171
    x = _rand((cap * 128, 128), 1)
    w = _rand((128, host.cols), 2)
    out = torch.full((big.nnz_blocks, 128, 128), 5.0, device=DEV, dtype=torch.bfloat16)
    for _ in range(3):   # the self-resetting tile counter must survive the clamped tile count
        sb.matmul_ex(sb.Matrix.of(x), False, sb.Matrix.of(w), False, big.matrix(out))
    ref = ops.sdd(x[:host.rows].contiguous(), w, exact)
    assert torch.equal(out[:171], ref)
    assert bool((out[171:] == 5.0).all())


def test_ffn_on_device_built_topology():
    """h = sdd(x, w1); y = dsd(h, w2) with gradients, topology from the device histogram: same
    bits as the host-built exact topology for y and dx, dw1, dw2."""
    tokens = [256, 512, 0, 256]
    F, cap = 2, 12
    t = torch.tensor(tokens, dtype=torch.int32, device=DEV)
    host, _ = host_moe(tokens, F, 2, cap)
    R = host.block_rows
    big = ops.Topology.for_moe(t, F * 128, cap * 128, 256)
    exact = ops.Topology.from_host(host, DEV)
    hid = 256
    x_live = _rand((R * 128, hid), 1)
    pad = torch.zeros(((cap - R) * 128, hid), device=DEV, dtype=torch.bfloat16)
    w1 = (_rand((hid, host.cols), 2) * 0.1).bfloat16()
    w2 = (_rand((host.cols, hid), 3) * 0.1).bfloat16()
    dy_live = _rand((R * 128, hid), 4)

    def run(topo, x, dy):
        x, a, b = x.clone().requires_grad_(), w1.clone().requires_grad_(), w2.clone().requires_grad_()
        h = ops.sdd(x, a, topo)
        if topo is big:   # dead blocks are uninitialised by contract: give them a value
            h = torch.cat([h[:exact.nnz_blocks], torch.zeros_like(h[exact.nnz_blocks:])])
        y = ops.dsd(h, topo, b)
        y.backward(dy)
        return y, x.grad, a.grad, b.grad

    y1, dx1, da1, db1 = run(exact, x_live, dy_live)
    y2, dx2, da2, db2 = run(big, torch.cat([x_live, pad]), torch.cat([dy_live, torch.zeros_like(pad)]))
    assert torch.equal(y2[:R * 128], y1) and not y2[R * 128:].any()
    assert torch.equal(dx2[:R * 128], dx1) and not dx2[R * 128:].any()
    assert torch.equal(da2, da1)
    assert torch.equal(db2, db1)


def test_builder_stays_inside_its_buffers():
    """Guard zones around every output array of bsmm_moe_topology (compute-sanitizer is not
    available on the GPU pool): nothing outside [0, n) is written, for a ragged and a truncated
    histogram."""
    import ctypes
    from sputnik_b200 import _lib
    from sputnik_b200.block import _ptr, _stream_ptr
    for tokens, F, align, cap in (([100, 0, 700, 1, 128, 129], 3, 1, 16), ([600, 600, 600], 2, 2, 8)):
        E = len(tokens)
        t = torch.tensor(tokens, dtype=torch.int32, device=DEV)
        nnz, G = cap * F, 64

        def guarded(n, dtype):
            base = torch.full((n + 2 * G,), 77, dtype=dtype, device=DEV)
            return base, base[G:G + n]
        bufs = {"offsets": guarded(cap + 1, torch.int32), "indices": guarded(nnz, torch.int16),
                "row_indices": guarded(nnz, torch.int16), "offsets_t": guarded(E * F + 1, torch.int32),
                "indices_t": guarded(nnz, torch.int16), "block_offsets": guarded(nnz, torch.int32),
                "bins": guarded(E + 1, torch.int32)}
        m = sb.BlockMatrix(cap * 128, E * F * 128, 128, nnz * 128 * 128, None,
                           bufs["offsets"][1], bufs["indices"][1], bufs["offsets_t"][1],
                           bufs["indices_t"][1], bufs["block_offsets"][1], bufs["row_indices"][1])
        cm = m.c()
        _lib.check(_lib.lib.bsmm_moe_topology(_ptr(t), E, F, align, ctypes.byref(cm),
                                              _ptr(bufs["bins"][1]), _stream_ptr(None)))
        torch.cuda.synchronize()
        for name, (base, view) in bufs.items():
            n = view.numel()
            assert bool((base[:G] == 77).all()) and bool((base[G + n:] == 77).all()), name
        host, hbins = host_moe(tokens, F, align, cap)
        assert np.array_equal(bufs["offsets"][1].cpu().numpy()[:host.block_rows + 1], host.offsets)
        assert np.array_equal(bufs["bins"][1].cpu().numpy(), hbins)
