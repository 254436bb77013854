"""Split-K of the remainder wave in the 2-SM DSD kernel (bsmm_pair_sk_kernel.cuh): parity against
the oracle with the schedule forced on small twin-row topologies (every tile cut in two across
clusters, empty pairs, one-block rows, ragged dense extents, both walk directions, both dtypes),
agreement with the classic schedule, determinism, back-to-back launches (workspace regions /
epochs), and the full-size C2 case.  The schedule is an opt-in experiment (BSMM_SPLITK=1: on
B200 the hand-off costs more than the balance saves), so every test forces it."""
import os

import numpy as np
import pytest

from sputnik_b200 import synth
from test_gpu_pair import _run, twin_topology
from util import assert_close

pytestmark = pytest.mark.gpu


@pytest.fixture
def streamk():
    """Force the split-K schedule (the library reads BSMM_SPLITK on every call)."""
    old = os.environ.get("BSMM_SPLITK")
    os.environ["BSMM_SPLITK"] = "1"
    yield
    if old is None:
        del os.environ["BSMM_SPLITK"]
    else:
        os.environ["BSMM_SPLITK"] = old


def _classic(*args, **kw):
    old = os.environ.get("BSMM_SPLITK")
    os.environ["BSMM_SPLITK"] = "0"
    try:
        return _run(*args, **kw)
    finally:
        if old is None:
            del os.environ["BSMM_SPLITK"]
        else:
            os.environ["BSMM_SPLITK"] = old


def _close_to_classic(got, classic, ref):
    """Same products, different fp32 association: within a few 16-bit ulps of each other."""
    rms = float(np.sqrt(np.mean(ref.astype(np.float64) ** 2))) or 1.0
    ulp2 = 2.0 ** -6 * np.maximum(np.abs(classic), rms)      # two bf16 ulps of the value
    assert bool((np.abs(got - classic) <= ulp2).all())


@pytest.mark.parametrize("dtype", ["bf16", "f16"])
@pytest.mark.parametrize("ta,tb", [(0, 0), (0, 1), (1, 0), (1, 1)])
def test_forced_moe_all_transposes(streamk, ta, tb, dtype):
    # 3 experts x 4 block rows x 2 block cols: 12 tiles of 2 blocks on 74 clusters, each tile
    # has one finisher (block 0) and one writer (block 1)
    topo = synth.moe_topology(3, 4, 2)
    got, ref = _run("dsd", topo, ta, tb, 264, dtype, seed=3)
    assert_close(got, ref, dtype, f"split-K dsd {ta}{tb}")
    _close_to_classic(got, _classic("dsd", topo, ta, tb, 264, dtype, seed=3)[0], ref)


@pytest.mark.parametrize("dense_ext", [136, 256, 520, 1032])
@pytest.mark.parametrize("unordered", [False, True])
def test_forced_ragged_lists_and_empty_pairs(streamk, dense_ext, unordered):
    topo = twin_topology(9, 12, seed=11, unordered=unordered, empty_pairs=(0, 4, 8))
    got, ref = _run("dsd", topo, 0, 0, dense_ext, "bf16", seed=5)
    assert_close(got, ref, "bf16", "split-K dsd ragged")
    _close_to_classic(got, _classic("dsd", topo, 0, 0, dense_ext, "bf16", seed=5)[0], ref)


def test_forced_long_rows_many_way_split(streamk):
    # 2 row pairs x 60 blocks, one 256-wide tile each: heads and tails of 30 blocks
    topo = synth.Topology(4, 60, np.arange(5, dtype=np.int32) * 60,
                          np.tile(np.arange(60, dtype=np.int16), 4))
    got, ref = _run("dsd", topo, 0, 0, 256, "bf16", seed=7)
    assert_close(got, ref, "bf16", "split-K long rows")


def test_forced_deterministic_and_repeatable(streamk):
    """Same launch 6 times in a row (more than the workspace regions): identical bits."""
    topo = twin_topology(16, 40, seed=4)
    first, ref = _run("dsd", topo, 0, 0, 1024, "bf16", seed=9)
    assert_close(first, ref, "bf16", "split-K")
    for _ in range(5):
        again, _ = _run("dsd", topo, 0, 0, 1024, "bf16", seed=9)
        assert np.array_equal(first, again)


def test_c2_full_size():
    """BASELINE C2 at full size (128 tiles on 74 CTA pairs, 54 of them cut 24 + 8): split-K
    agrees with the classic schedule and with an fp32 dense product, every output element."""
    import torch
    from sputnik_b200 import harness, workloads
    wl = workloads.get("c2")
    prob = harness.DeviceProblem(wl, "cuda:0", sets=1, seed=3)
    s = prob.sets[0]
    assert s.sparse.twin_rows
    outs = {}
    for mode in ("0", "1"):
        old = os.environ.pop("BSMM_SPLITK", None)
        if mode is not None:
            os.environ["BSMM_SPLITK"] = mode
        try:
            s.c.data.fill_(float("nan"))
            prob.call(s)
            torch.cuda.synchronize()
            outs[mode] = s.c.data.float().clone()
        finally:
            os.environ.pop("BSMM_SPLITK", None)
            if old is not None:
                os.environ["BSMM_SPLITK"] = old
    assert torch.isfinite(outs["1"]).all()
    rms = outs["0"].pow(2).mean().sqrt()
    diff = (outs["1"] - outs["0"]).abs()
    assert bool((diff <= 2.0 ** -6 * torch.maximum(outs["0"].abs(), rms)).all())
    assert float((diff > 0).float().mean()) > 0.0        # it really is another schedule
    # every output element against an fp32 dense product of the same 16-bit operands
    t = wl.topo
    rows = t.row_indices().astype(np.int64)
    dense = torch.zeros((t.block_rows, t.block_cols, 128, 128), device="cuda:0")
    dense[torch.from_numpy(rows).cuda(), torch.from_numpy(t.indices.astype(np.int64)).cuda()] = \
        s.sparse.data.float()
    dense = dense.permute(0, 2, 1, 3).reshape(t.rows, t.cols)
    ref = (dense.double() @ s.b.data.double())
    tau = 1e-2 * ref.pow(2).mean().sqrt()
    rel = ((outs["1"].double() - ref).abs() / torch.maximum(ref.abs(), tau)).max()
    assert float(rel) <= 1e-2
