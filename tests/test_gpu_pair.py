"""GPU parity of the 2-SM (CTA-pair, tcgen05.mma.cta_group::2) DSD / DDS kernel against the oracle
on topologies with twin block rows / columns (dMoE block-diagonal and hand-made), every transpose
pair, both dtypes, ragged dense extents, empty twin pairs, unordered indices — and equality with
the 1-SM kernel's result (same inputs, BSMM hint off)."""
import numpy as np
import pytest

from oracle import oracle
from sputnik_b200 import synth
from util import assert_close, dev_block_matrix, to_dev, torch_dtype

pytestmark = pytest.mark.gpu


def twin_topology(pairs, block_cols, seed, unordered=False, empty_pairs=()):
    """Block rows 2i, 2i+1 share a random column list; pairs in `empty_pairs` are empty."""
    rng = np.random.default_rng(seed)
    offsets, indices = [0], []
    for pr in range(pairs):
        n = 0 if pr in empty_pairs else int(rng.integers(1, block_cols + 1))
        cols = np.sort(rng.permutation(block_cols)[:n])
        if unordered:
            rng.shuffle(cols)
        for _ in range(2):
            indices.extend(cols.tolist())
            offsets.append(len(indices))
    return synth.Topology(2 * pairs, block_cols, np.array(offsets), np.array(indices, dtype=np.int16))


def _run(op, topo, ta, tb, dense_ext, dtype, seed, hint=True):
    import torch
    import sputnik_b200 as sb
    r = lambda shape, s: oracle.round_to(synth.uniform_values(shape, seed * 131 + s), dtype)
    vals = r((topo.nnz_blocks, 128, 128), 1)
    sp = dev_block_matrix(topo, vals, dtype)
    sb.allocate_transpose_buffers(sp)
    sb.transpose(sp)
    if hint:
        sp.twin_rows, sp.twin_cols = topo.twin_rows(), topo.twin_cols()
    if op == "dsd":
        kdim, m = (topo.rows, topo.cols) if ta else (topo.cols, topo.rows)
        b = r((dense_ext, kdim) if tb else (kdim, dense_ext), 2)
        out = torch.full((m, dense_ext), float("nan"), device="cuda:0", dtype=torch_dtype(dtype))
        sb.matmul_ex(sp, ta, sb.Matrix.of(to_dev(b, dtype)), tb, sb.Matrix.of(out))
        ref = oracle.dsd(vals, topo, ta, b, tb)
    else:
        kdim, n = (topo.cols, topo.rows) if tb else (topo.rows, topo.cols)
        a = r((kdim, dense_ext) if ta else (dense_ext, kdim), 2)
        out = torch.full((dense_ext, n), float("nan"), device="cuda:0", dtype=torch_dtype(dtype))
        sb.matmul_ex(sb.Matrix.of(to_dev(a, dtype)), ta, sp, tb, sb.Matrix.of(out))
        ref = oracle.dds(a, ta, vals, topo, tb)
    torch.cuda.synchronize()
    return out.float().cpu().numpy(), ref


@pytest.mark.parametrize("dtype", ["bf16", "f16"])
@pytest.mark.parametrize("ta,tb", [(0, 0), (0, 1), (1, 0), (1, 1)])
@pytest.mark.parametrize("op", ["dsd", "dds"])
def test_moe_topology_all_transposes(op, ta, tb, dtype):
    # 3 experts x 4 block rows x 2 block cols: twin rows AND twin columns
    topo = synth.moe_topology(3, 4, 2)
    assert topo.twin_rows() and topo.twin_cols()
    got, ref = _run(op, topo, ta, tb, 264, dtype, seed=3)   # 264: one full + one 8-wide tile
    assert_close(got, ref, dtype, f"{op} {ta}{tb}")


@pytest.mark.parametrize("dense_ext", [136, 256, 520])
@pytest.mark.parametrize("unordered", [False, True])
def test_twin_rows_random_lists(dense_ext, unordered):
    topo = twin_topology(5, 6, seed=11, unordered=unordered, empty_pairs=(1,))
    assert topo.twin_rows() and not topo.twin_cols()
    got, ref = _run("dsd", topo, 0, 0, dense_ext, "bf16", seed=5)        # row walk: pair kernel
    assert_close(got, ref, "bf16", "dsd nn twins")
    got, ref = _run("dds", topo, 0, 1, dense_ext, "bf16", seed=6)        # row walk, transposed out
    assert_close(got, ref, "bf16", "dds nt twins")
    got, ref = _run("dsd", topo, 1, 0, dense_ext, "bf16", seed=7)        # column walk: 1-SM kernel
    assert_close(got, ref, "bf16", "dsd tn (no twin cols)")


def test_pair_kernel_equals_single_sm_kernel():
    import sputnik_b200 as sb
    topo = synth.moe_topology(2, 6, 4)
    before = sb._lib.lib.bsmm_launch_count()
    a, ref = _run("dsd", topo, 0, 0, 512, "bf16", seed=9, hint=True)
    b, _ = _run("dsd", topo, 0, 0, 512, "bf16", seed=9, hint=False)
    assert sb._lib.lib.bsmm_launch_count() - before == 4   # 2 x (transpose + matmul)
    assert_close(a, ref, "bf16", "pair")
    # same products, same fp32 accumulation order along K: identical 16-bit results
    assert np.array_equal(a, b)


def test_c2_sized_property_linearity():
    """BASELINE C2 at full size: dsd(h, w2 + w2') == dsd(h, w2) + dsd(h, w2') within rounding."""
    import torch
    import sputnik_b200 as sb
    from sputnik_b200 import harness, workloads
    wl = workloads.get("c2")
    prob = harness.DeviceProblem(wl, "cuda:0", sets=1, seed=3)
    s = prob.sets[0]
    assert s.sparse.twin_rows
    b1 = s.b.data.clone()
    b2 = torch.randn_like(b1)
    outs = []
    for bb in (b1, b2, (b1.float() + b2.float()).to(b1.dtype)):
        s.b.data.copy_(bb)
        prob.call(s)
        outs.append(s.c.data.float().clone())
    torch.cuda.synchronize()
    ref = outs[0] + outs[1]
    rms = ref.pow(2).mean().sqrt()
    # the summed operand is itself rounded to bf16 (2^-9 relative per element of w2)
    assert float((outs[2] - ref).abs().max() / rms) < 5e-2
    assert float((outs[2] - ref).pow(2).mean().sqrt() / rms) < 1e-2


def _run_sdd(topo, ta, tb, k, dtype, seed, hint=True):
    import torch
    import sputnik_b200 as sb
    r = lambda shape, s: oracle.round_to(synth.uniform_values(shape, seed * 977 + s), dtype)
    m, n = topo.rows, topo.cols
    a = r((k, m) if ta else (m, k), 1)
    b = r((n, k) if tb else (k, n), 2)
    c = dev_block_matrix(topo, None, dtype, fill=float("nan"))
    sb.allocate_row_indices_buffer(c)
    sb.row_indices(c)
    if hint:
        c.__dict__.update(topo.hints())
    sb.matmul(sb.Matrix.of(to_dev(a, dtype)), ta, sb.Matrix.of(to_dev(b, dtype)), tb, c)
    torch.cuda.synchronize()
    return c.data.float().cpu().numpy(), oracle.sdd(a, ta, b, tb, topo)


@pytest.mark.parametrize("dtype", ["bf16", "f16"])
@pytest.mark.parametrize("ta,tb", [(0, 0), (0, 1), (1, 0), (1, 1)])
def test_sdd_cliques_all_transposes(ta, tb, dtype):
    topo = synth.moe_topology(3, 4, 2)
    assert topo.twin_rows() and topo.even_rows()
    got, ref = _run_sdd(topo, ta, tb, 200, dtype, seed=4)   # k = 3 full steps + 8
    assert_close(got, ref, dtype, f"sdd {ta}{tb}")


@pytest.mark.parametrize("unordered", [False, True])
def test_sdd_cliques_ragged_even_rows(unordered):
    # twin rows with different (even) lengths per pair, one empty pair, unordered columns
    rng = np.random.default_rng(21)
    offsets, indices = [0], []
    for n in (4, 0, 2, 6, 2):
        cols = np.sort(rng.permutation(7)[:n])
        if unordered:
            rng.shuffle(cols)
        for _ in range(2):
            indices.extend(cols.tolist())
            offsets.append(len(indices))
    topo = synth.Topology(10, 7, np.array(offsets), np.array(indices, dtype=np.int16))
    assert topo.twin_rows() and topo.even_rows()
    got, ref = _run_sdd(topo, 0, 0, 264, "bf16", seed=8)
    assert_close(got, ref, "bf16", "sdd ragged cliques")
    single, _ = _run_sdd(topo, 0, 0, 264, "bf16", seed=8, hint=False)
    assert np.array_equal(got, single)
