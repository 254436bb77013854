"""Device-side kernel selection for hint-less callers (bsmm_api.cu "auto pair"): with no topology
hint, a gate kernel classifies the topology on the device and exactly one of the 2-SM / 1-SM
launches runs.  Forced here for small problems with BSMM_AUTO_PAIR=1 (the default threshold only
applies it to products >= 1.5e11 flops).  Checked: results against the oracle on twin, non-twin and
almost-twin topologies (a wrong gate decision gives wrong numbers), equality with the hinted
call, and the launch count (gate + two gated kernels)."""
import os

import numpy as np
import pytest

import sputnik_b200 as sb
from sputnik_b200 import synth
from test_gpu_pair import _run, _run_sdd, twin_topology
from util import assert_close

pytestmark = pytest.mark.gpu


@pytest.fixture
def auto_pair():
    old = os.environ.get("BSMM_AUTO_PAIR")
    os.environ["BSMM_AUTO_PAIR"] = "1"
    yield
    if old is None:
        del os.environ["BSMM_AUTO_PAIR"]
    else:
        os.environ["BSMM_AUTO_PAIR"] = old


def _almost_twin():
    """dMoE topology with one column index of one block row changed."""
    t = synth.moe_topology(4, 2, 4)
    idx = t.indices.copy()
    idx[5] = (idx[5] + 4) % t.block_cols
    return synth.Topology(t.block_rows, t.block_cols, t.offsets, idx)


TOPOS = {"twin": lambda: synth.moe_topology(3, 4, 2),
         "twin_ragged": lambda: twin_topology(5, 6, seed=11, empty_pairs=(1,)),
         "random": lambda: synth.random_topology(6, 6, 17, seed=3, unordered=True),
         "almost_twin": _almost_twin}


@pytest.mark.parametrize("ta,tb", [(0, 0), (0, 1), (1, 0), (1, 1)])
@pytest.mark.parametrize("op", ["dsd", "dds"])
@pytest.mark.parametrize("kind", list(TOPOS))
def test_hintless_dsd_dds(auto_pair, kind, op, ta, tb):
    topo = TOPOS[kind]()
    before = sb._lib.lib.bsmm_launch_count()
    got, ref = _run(op, topo, ta, tb, 264, "bf16", seed=5, hint=False)
    launches = sb._lib.lib.bsmm_launch_count() - before
    assert_close(got, ref, "bf16", f"auto {op} {ta}{tb} {kind}")
    hinted, _ = _run(op, topo, ta, tb, 264, "bf16", seed=5, hint=True)
    assert np.array_equal(got, hinted)
    walk_even = ((topo.block_cols if (ta if op == "dsd" else not tb) else topo.block_rows) % 2) == 0
    assert launches == (4 if walk_even else 2), launches   # transpose + [gate + 2 gated] GEMM(s)


@pytest.mark.parametrize("ta,tb", [(0, 0), (0, 1), (1, 0), (1, 1)])
@pytest.mark.parametrize("kind", ["twin", "random", "almost_twin"])
def test_hintless_sdd(auto_pair, kind, ta, tb):
    topo = TOPOS[kind]()
    before = sb._lib.lib.bsmm_launch_count()
    got, ref = _run_sdd(topo, ta, tb, 200, "bf16", seed=4, hint=False)
    launches = sb._lib.lib.bsmm_launch_count() - before
    assert_close(got, ref, "bf16", f"auto sdd {ta}{tb} {kind}")
    hinted, _ = _run_sdd(topo, ta, tb, 200, "bf16", seed=4, hint=True)
    assert np.array_equal(got, hinted)
    clique_shape = topo.nnz_blocks % 4 == 0 and topo.block_rows % 2 == 0
    assert launches == (4 if clique_shape else 2), launches  # row_indices + [gate + 2] GEMM(s)


def test_default_threshold_leaves_small_products_alone():
    os.environ.pop("BSMM_AUTO_PAIR", None)
    topo = synth.moe_topology(3, 4, 2)
    before = sb._lib.lib.bsmm_launch_count()
    _run("dsd", topo, 0, 0, 264, "bf16", seed=5, hint=False)
    assert sb._lib.lib.bsmm_launch_count() - before == 2        # transpose + one GEMM
