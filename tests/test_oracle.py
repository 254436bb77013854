"""CPU tests of the oracle: against the reference-generated golden fixtures, against the
reference's own host code when oracle/_ref is present, and internal consistency
(block-aware variant == as-written dense variant, bit for bit)."""
import ctypes
import os

import numpy as np
import pytest

from oracle import oracle
from sputnik_b200 import synth
from util import Case, REF_HOST_LIB

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "host_matmul.npz")


def _bits(x):
    return np.ascontiguousarray(x, np.float32).view(np.uint32)


def _golden_cases():
    z = np.load(GOLDEN)
    for i in range(int(z["n"])):
        s = list(z[f"spec{i}"])
        spec = (s[0], int(s[1]), int(s[2]), int(s[3]), int(s[4]), s[5] == "True", s[6] == "True",
                s[7] == "True", s[8], int(s[9]))
        yield spec, z[f"expected{i}"]


@pytest.mark.parametrize("spec,expected", list(_golden_cases()),
                         ids=lambda v: "-".join(map(str, v)) if isinstance(v, tuple) else "exp")
def test_oracle_matches_reference_golden(spec, expected):
    """Fixtures were produced by the reference's Matrix::operator* / T / Mul
    (oracle/make_golden.py); the restatement must reproduce them bit for bit."""
    c = Case(*spec)
    dense = c.expected(dense=True)
    blocked = c.expected(dense=False)
    if c.op == "sdd":  # golden is the masked dense matrix
        exp = oracle.sample(expected, c.topo.offsets, c.topo.indices, c.topo.block_rows,
                            c.topo.block_cols)
        full = oracle.densify(dense, c.topo.offsets, c.topo.indices, c.topo.block_rows,
                              c.topo.block_cols)
        assert np.array_equal(full, expected)
    else:
        exp = expected
    assert np.array_equal(_bits(dense), _bits(exp))
    assert np.array_equal(blocked, exp)  # numeric equality (-0.0 == 0.0 allowed)


@pytest.mark.skipif(not os.path.exists(REF_HOST_LIB), reason="oracle/_ref not built here")
def test_oracle_matmul_bit_identical_to_reference_host():
    h = ctypes.CDLL(REF_HOST_LIB)
    f = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
    h.ref_host_matmul.argtypes = [f, f, f, ctypes.c_int, ctypes.c_int, ctypes.c_int]
    h.ref_host_transpose.argtypes = [f, f, ctypes.c_int, ctypes.c_int]
    rng = np.random.default_rng(5)
    for (m, k, n) in [(8, 128, 128), (136, 264, 72), (1, 1, 1), (257, 33, 5)]:
        a = rng.uniform(-1, 1, (m, k)).astype(np.float32)   # un-rounded fp32, as the reference feeds
        b = rng.uniform(-1, 1, (k, n)).astype(np.float32)
        out = np.empty((m, n), np.float32)
        h.ref_host_matmul(a, b, out, m, k, n)
        assert np.array_equal(_bits(out), _bits(oracle.matmul(a, b)))
        t = np.empty((k, m), np.float32)
        h.ref_host_transpose(a, t, m, k)
        assert np.array_equal(t, oracle.transpose(a))


@pytest.mark.parametrize("op", ["sdd", "dsd", "dds"])
@pytest.mark.parametrize("ta", [False, True])
@pytest.mark.parametrize("tb", [False, True])
@pytest.mark.parametrize("unordered", [False, True])
def test_blocked_equals_dense(op, ta, tb, unordered):
    c = Case(op, 384, 256, 136 if op == "dsd" else 256, 3 * 16384, ta, tb, unordered, "bf16", 21)
    if op == "dds":
        c = Case(op, 72, 256, 384, 3 * 16384, ta, tb, unordered, "f16", 22)
    assert np.array_equal(c.expected(dense=False), c.expected(dense=True))


def test_empty_rows_give_zero_tiles():
    topo = synth.Topology(3, 2, np.array([0, 2, 2, 3], np.int32), np.array([1, 0, 1], np.int16))
    vals = synth.uniform_values((3, 128, 128), 1)
    b = synth.uniform_values((256, 8), 2)
    out = oracle.dsd(vals, topo, False, b, False)
    assert np.all(out[128:256] == 0) and np.any(out[:128] != 0)
    outt = oracle.dsd(vals, topo, True, synth.uniform_values((384, 8), 3), False)
    assert outt.shape == (256, 8)


@pytest.mark.parametrize("br,bc,nb,unordered,seed", [
    (4, 4, 8, False, 0), (8, 8, 32, True, 1), (64, 256, 2048, False, 2), (40, 3, 100, True, 3),
    (3, 5, 1, False, 4), (5, 7, 0, False, 5), (1, 1, 1, False, 6), (16, 16, 256, True, 7)])
def test_transpose_metadata_is_stable_sort(br, bc, nb, unordered, seed):
    """transpose.cu:87-104 restated; compared with numpy's stable argsort."""
    t = synth.random_topology(br, bc, nb, seed, unordered)
    ot, it, bo = oracle.transpose_metadata(t.offsets, t.indices, br, bc)
    order = np.argsort(t.indices, kind="stable")
    assert np.array_equal(bo, order.astype(np.int32))
    assert np.array_equal(it, t.row_indices()[order])
    exp_ot = np.zeros(bc + 1, np.int32)
    np.cumsum(np.bincount(t.indices, minlength=bc), out=exp_ot[1:])
    assert np.array_equal(ot, exp_ot)
    assert np.array_equal(oracle.row_indices(t.offsets, br), t.row_indices())


def test_rounding_helpers():
    x = synth.uniform_values((1000,), 3)
    import torch
    tb = torch.from_numpy(x).to(torch.bfloat16)
    assert np.array_equal(oracle.round_to(x, "bf16"), tb.float().numpy())
    assert np.array_equal(oracle.to_bits16(x, "bf16"), tb.view(torch.int16).numpy().view(np.uint16))
    assert np.array_equal(oracle.round_to(x, "f16"), torch.from_numpy(x).half().float().numpy())
    assert np.array_equal(oracle.from_bits16(oracle.to_bits16(x, "bf16"), "bf16"),
                          oracle.round_to(x, "bf16"))


def test_moe_topology():
    t = synth.moe_topology(8, 8, 32)
    assert (t.block_rows, t.block_cols, t.nnz_blocks) == (64, 256, 2048)
    assert t.offsets[-1] == 2048 and t.indices[32 * 8] == 32 and t.indices[-1] == 255


def test_round_e4m3_matches_torch_float8():
    """oracle.round_e4m3 == a round trip through torch.float8_e4m3fn (round-to-nearest-even,
    subnormals, the largest finite value 448) on every value the tests feed it."""
    import torch
    rng = np.random.default_rng(0)
    x = np.concatenate([rng.uniform(-1, 1, 200000), rng.uniform(-0.02, 0.02, 100000),
                        rng.uniform(-448, 448, 10000),
                        [0.0, 448.0, -448.0, 2.0 ** -9, 2.0 ** -10, 3 * 2.0 ** -10, 0.0175, 1.0625,
                         1.1875]]).astype(np.float32)
    ref = torch.from_numpy(x).to(torch.float8_e4m3fn).float().numpy()
    assert np.array_equal(oracle.round_e4m3(x), ref)
    assert np.array_equal(oracle.round_to(x, "f8"), ref)
