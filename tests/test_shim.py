"""The header-only C++ shim (include/sputnik/sputnik.h): a caller written against the
reference's `sputnik::block` interface compiles, links with libbsmm.so and behaves like the
reference (tests/cxx/shim_test.cu, modelled on the reference's dsd_test.cu / sdd_test.cu)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "tests", "cxx", "_build", "shim_test")


@pytest.fixture(scope="module")
def shim_bin():
    if not os.path.exists(BIN):
        import __graft_entry__ as g
        g.build()
    return BIN


def test_shim_compiles_and_maps_errors(shim_bin):
    out = subprocess.run([shim_bin, "abi"], capture_output=True, text=True, timeout=60)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "shim abi ok" in out.stdout


@pytest.mark.gpu
def test_shim_runs_all_three_ops_on_device(shim_bin):
    out = subprocess.run([shim_bin, "gpu"], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "shim gpu ok" in out.stdout
