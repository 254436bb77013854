"""GPU tests of the fused epilogues (include/bsmm.h BSMM_FLAG_ACCUMULATE /
RELU / GELU) on every kernel family: 1-SM (random topologies), 2-SM (dMoE topologies with the
twin / even hints) and the siblings.

 * accumulate: C += op(A) op(B) must equal, BIT FOR BIT, the unfused sequence on the same GPU
   (product into a temporary, then an elementwise add in fp32 rounded to the 16-bit type), and
   agree with the CPU oracle's `c0 + product` within the path's tolerance (1e-2 relative);
 * relu: equal to relu of the unfused product (rounding commutes with max(x, 0));
 * gelu (tanh form): within tolerance of the oracle product passed through the same formula in
   float64 (applied to the fp32 accumulator before the one rounding)."""
import numpy as np
import pytest

from oracle import oracle
from sputnik_b200 import synth
from util import Case, SibCase, assert_close, run_ours, run_sibling, torch_dtype

pytestmark = pytest.mark.gpu

TRANSPOSES = [(False, False), (False, True), (True, False), (True, True)]


def random_case(op, ta, tb, dtype, seed=1):
    # 1-SM kernels: unstructured topology, ragged dense extents for DSD / DDS
    if op == "sdd":
        return Case(op, 512, 200, 768, 11 * 16384, ta, tb, True, dtype, seed)
    if op == "dsd":
        return Case(op, 512, 768, 328, 13 * 16384, ta, tb, True, dtype, seed)
    return Case(op, 328, 768, 512, 13 * 16384, ta, tb, True, dtype, seed)


def moe_case(op, ta, tb, dtype, seed=2):
    """dMoE block-diagonal topology (3 experts x 4 block rows x 2 block columns): twin rows, twin
    columns, even rows -> the 2-SM kernels when the hints are passed."""
    topo = synth.moe_topology(3, 4, 2)
    sr, sc = topo.rows, topo.cols
    if op == "sdd":
        m, n, k = sr, sc, 256
    elif op == "dsd":
        (k, m), n = ((sr, sc) if ta else (sc, sr)), 512
    else:
        (n, k), m = ((sr, sc) if tb else (sc, sr)), 512
    c = Case.__new__(Case)
    c.op, c.m, c.k, c.n, c.ta, c.tb, c.dtype, c.seed, c.topo = op, m, k, n, ta, tb, dtype, seed, topo
    r = lambda shape, s: oracle.round_to(synth.uniform_values(shape, seed * 7919 + s), dtype)
    c.values = r((topo.nnz_blocks, 128, 128), 1)
    if op != "dsd":
        c.a = r((k, m) if ta else (m, k), 2)
    if op != "dds":
        c.b = r((n, k) if tb else (k, n), 3)
    return c


def out_shape(c):
    return (c.topo.nnz_blocks, 128, 128) if c.op == "sdd" else (c.m, c.n)


def add16(c0, t, dtype):
    """fp32 add of two 16-bit-representable arrays, rounded to the 16-bit type (on the host)."""
    return oracle.round_to(np.asarray(c0, np.float32) + np.asarray(t, np.float32), dtype)


def assert_accumulated(got, c0, product, dtype, what):
    """Against the CPU oracle: two roundings to the 16-bit type (the product at its own magnitude,
    then the sum), so the bound is half an ulp of each — not a relative error of the sum, which
    cancellation between c0 and the product can make arbitrarily small."""
    eps = 2.0 ** -8 if dtype == "bf16" else 2.0 ** -11     # half an ulp, relative
    ref = np.asarray(c0, np.float64) + np.asarray(product, np.float64)
    bound = 1.01 * eps * (np.abs(product) + np.abs(ref)) + 1e-4 * np.sqrt(np.mean(product ** 2))
    err = np.abs(np.asarray(got, np.float64) - ref)
    assert np.isfinite(got).all() and (err <= bound).all(), \
        f"{what}: worst excess {float((err - bound).max()):.3e}"


@pytest.mark.parametrize("dtype", ["bf16", "f16"])
@pytest.mark.parametrize("ta,tb", TRANSPOSES)
@pytest.mark.parametrize("op", ["sdd", "dsd", "dds"])
@pytest.mark.parametrize("family", ["1sm", "2sm"])
def test_accumulate(family, op, ta, tb, dtype):
    c = (random_case if family == "1sm" else moe_case)(op, ta, tb, dtype)
    hints = family == "2sm"
    c0 = oracle.round_to(8.0 * synth.uniform_values(out_shape(c), 77), dtype)
    plain = run_ours(c, ex=False, hints=hints)
    fused = run_ours(c, ex=False, hints=hints, prefill=c0, accumulate=True)
    assert np.array_equal(fused, add16(c0, plain, dtype)), "fused accumulate != unfused sequence"
    assert_accumulated(fused, c0, c.expected(), dtype, f"accumulate {family} {op} {ta} {tb}")
    # twice: C = c0 + 2 P (the second launch reads what the first one wrote)
    # (checked against the unfused chain, again bit for bit)
    if op == "dsd" and not ta and not tb:
        import torch
        import sputnik_b200 as sb
        from util import dev_block_matrix, to_dev
        sp = dev_block_matrix(c.topo, c.values, dtype)
        if hints:
            sp.__dict__.update(c.topo.hints())
        out = to_dev(c0, dtype)
        b = sb.Matrix.of(to_dev(c.b, dtype))
        for _ in range(2):
            sb.matmul(sp, False, b, False, sb.Matrix.of(out), accumulate=True)
        torch.cuda.synchronize()
        want = add16(add16(c0, plain, dtype), plain, dtype)
        assert np.array_equal(out.float().cpu().numpy(), want)


@pytest.mark.parametrize("dtype", ["bf16", "f16"])
@pytest.mark.parametrize("op", ["sdd", "dsd", "dds"])
@pytest.mark.parametrize("family", ["1sm", "2sm"])
def test_relu(family, op, dtype):
    for ta, tb in TRANSPOSES:
        c = (random_case if family == "1sm" else moe_case)(op, ta, tb, dtype)
        hints = family == "2sm"
        plain = run_ours(c, hints=hints)
        fused = run_ours(c, hints=hints, activation="relu")
        assert np.array_equal(fused, np.maximum(plain, 0.0)), f"relu {family} {op} {ta} {tb}"


def gelu64(x):
    x = np.asarray(x, np.float64)
    return 0.5 * x * (1.0 + np.tanh(0.7978845608028654 * (x + 0.044715 * x ** 3)))


@pytest.mark.parametrize("dtype", ["bf16", "f16"])
@pytest.mark.parametrize("op", ["sdd", "dsd", "dds"])
@pytest.mark.parametrize("family", ["1sm", "2sm"])
def test_gelu(family, op, dtype):
    for ta, tb in TRANSPOSES[:2] if op != "sdd" else TRANSPOSES:
        c = (random_case if family == "1sm" else moe_case)(op, ta, tb, dtype)
        # small operands so that the products land where GELU is not yet the identity
        c.values = oracle.round_to(c.values * 0.125, dtype)
        if op != "dsd":
            c.a = oracle.round_to(c.a * 0.125, dtype)
        if op != "dds":
            c.b = oracle.round_to(c.b * 0.125, dtype)
        fused = run_ours(c, hints=family == "2sm", activation="gelu")
        ref = gelu64(c.expected()).astype(np.float32)
        assert_close(fused, ref, dtype, f"gelu {family} {op} {ta} {tb}")
        assert (fused < 0).any() and (fused > 0).any()


def test_accumulate_leaves_c_alone_without_contributions():
    """Empty sparse operand / empty walks: C += 0 (the plain product would zero-fill)."""
    c = Case("dsd", 256, 384, 136, 0, False, False, False, "bf16", 0)
    c0 = oracle.round_to(synth.uniform_values((256, 136), 5), "bf16")
    assert np.array_equal(run_ours(c, prefill=c0, accumulate=True), c0)
    assert not run_ours(c, prefill=c0).any()
    c = Case("dsd", 512, 384, 136, 2 * 16384, False, False, False, "bf16", 3)  # some rows empty
    c0 = oracle.round_to(synth.uniform_values((512, 136), 6), "bf16")
    got = run_ours(c, prefill=c0, accumulate=True)
    assert np.array_equal(got, add16(c0, run_ours(c), "bf16"))


@pytest.mark.parametrize("op", ["ssd", "sds", "dss"])
def test_siblings_take_the_flags(op):
    import torch
    import sputnik_b200 as sb
    from util import dev_block_matrix, to_dev
    c = SibCase(op, 512, 512, 512, 9 * 16384, 7 * 16384, False, True, True, "bf16", 4)
    plain = run_sibling(c)

    def run(prefill, **kw):
        def sparse(topo, values):
            sp = dev_block_matrix(topo, values, "bf16")
            sb.allocate_transpose_buffers(sp)
            return sp
        a = sparse(c.topo_a, c.a) if c.topo_a is not None else sb.Matrix.of(to_dev(c.a, "bf16"))
        b = sparse(c.topo_b, c.b) if c.topo_b is not None else sb.Matrix.of(to_dev(c.b, "bf16"))
        if c.topo_c is not None:
            out_m = dev_block_matrix(c.topo_c, prefill, "bf16")
            sb.allocate_row_indices_buffer(out_m)
            sb.row_indices(out_m)
            out = out_m.data
        else:
            out = to_dev(prefill, "bf16")
            out_m = sb.Matrix.of(out)
        sb.matmul(a, c.ta, b, c.tb, out_m, **kw)
        torch.cuda.synchronize()
        return out.float().cpu().numpy()

    c0 = oracle.round_to(4.0 * synth.uniform_values(plain.shape, 9), "bf16")
    assert np.array_equal(run(c0, accumulate=True), add16(c0, plain, "bf16"))
    assert np.array_equal(run(c0, activation="relu"), np.maximum(plain, 0.0))


# ------------------------------------------------------------------ SwiGLU (BSMM_FLAG_SWIGLU, SDD)
def _swiglu_case(topo, k, ta, tb, dtype, seed):
    """A [m, k], B [k, 2n] (gate panel of output block column j at block column 2j, up at 2j + 1),
    both rounded to `dtype`, stored as the transpose flags say; + the fp64 expectation per block."""
    m, n = topo.rows, topo.cols
    r = lambda shape, s: oracle.round_to(synth.uniform_values(shape, seed * 7919 + s), dtype)
    a, b = r((m, k), 1), r((k, 2 * n), 2)
    bb = b.reshape(k, n // 128, 2, 128)
    gate, up = bb[:, :, 0].reshape(k, n), bb[:, :, 1].reshape(k, n)
    g = oracle.sdd(a, False, np.ascontiguousarray(gate), False, topo).astype(np.float64)
    u = oracle.sdd(a, False, np.ascontiguousarray(up), False, topo).astype(np.float64)
    ref = g / (1.0 + np.exp(-g)) * u
    return (np.ascontiguousarray(a.T) if ta else a), (np.ascontiguousarray(b.T) if tb else b), ref


@pytest.mark.parametrize("dtype", ["bf16", "f16"])
@pytest.mark.parametrize("ta,tb", TRANSPOSES)
@pytest.mark.parametrize("kind", ["random", "moe", "many", "moe_many"])
def test_sdd_swiglu(kind, ta, tb, dtype):
    """silu(op(A) gate) * (op(A) up) sampled at the topology, one fused SDD: against the oracle's
    two SDD products combined in float64.  `many`: more tiles than SMs (dynamic scheduler);
    `moe*`: twin-row hints -> 2-SM kernel, `moe_many` with several tiles per cluster and the group
    raster (4 row pairs per expert)."""
    import torch
    import sputnik_b200 as sb
    from util import dev_block_matrix, to_dev
    topo = {"random": lambda: synth.random_topology(4, 6, 11, seed=5, unordered=True),
            "moe": lambda: synth.moe_topology(3, 4, 2),
            "many": lambda: synth.random_topology(16, 24, 200, seed=6),
            "moe_many": lambda: synth.moe_topology(8, 8, 12)}[kind]()   # 384 tiles on 74 clusters
    k = 200 if kind in ("random", "moe") else 64       # ragged K (multiple of 8)
    a, b, ref = _swiglu_case(topo, k, ta, tb, dtype, 3)
    c = dev_block_matrix(topo, None, dtype, fill=float("nan"))
    sb.allocate_row_indices_buffer(c)
    sb.row_indices(c)
    if kind.startswith("moe"):
        c.__dict__.update(topo.hints())     # twin rows: the 2-SM kernel (gate / up panel per CTA)
    sb.matmul(sb.Matrix.of(to_dev(a, dtype)), ta, sb.Matrix.of(to_dev(b, dtype)), tb, c,
              activation="swiglu")
    torch.cuda.synchronize()
    # relative bound only: the outputs are products of two accumulators (magnitudes up to ~100),
    # so the reference's absolute fp16 bar for plain products (5e-2) does not apply
    got = c.data.float().cpu().numpy()
    from util import REL_TOL, errors
    rel, _ = errors(got, ref)
    assert np.isfinite(got).all() and rel <= REL_TOL, f"swiglu {kind} ta={ta} tb={tb}: rel {rel:.3e}"


def test_sdd_swiglu_argument_errors():
    import torch
    import sputnik_b200 as sb
    from sputnik_b200 import _lib
    from util import dev_block_matrix
    topo = synth.random_topology(2, 2, 3, seed=1)
    c = dev_block_matrix(topo, None, "bf16", fill=0.0)
    sb.allocate_row_indices_buffer(c)
    sb.row_indices(c)
    z = lambda *s: sb.Matrix.of(torch.zeros(s, dtype=torch.bfloat16, device="cuda:0"))
    with pytest.raises(sb.BsmmError) as e:      # op(B) with the columns of C, not twice as many
        sb.matmul(z(256, 64), False, z(64, 256), False, c, activation="swiglu")
    assert e.value.status in (_lib.BSMM_ERR_SHAPE, _lib.BSMM_ERR_MISALIGNED)
    sp = dev_block_matrix(topo, np.zeros((3, 128, 128), np.float32), "bf16")
    with pytest.raises(sb.BsmmError) as e:      # an epilogue of SDD only
        sb.matmul(sp, False, z(256, 64), False, z(256, 64), activation="swiglu")
    assert e.value.status == _lib.BSMM_ERR_SHAPE
    with pytest.raises(sb.BsmmError) as e:      # not combinable with accumulate
        sb.matmul(z(256, 64), False, z(64, 512), False, c, activation="swiglu", accumulate=True)
    assert e.value.status == _lib.BSMM_ERR_SHAPE
