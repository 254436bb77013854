"""CPU: the custom operators are registered and their fake-tensor shape functions agree with the
products' shape rules (sputnik/block/arguments.h MatmulShape) for every transpose pair."""
import pytest
import torch
from torch._subclasses.fake_tensor import FakeTensorMode

from sputnik_b200 import ops


def test_ops_are_registered():
    for name in ("sdd", "dsd", "dds"):
        assert hasattr(torch.ops.sputnik_b200, name)


@pytest.mark.parametrize("ta,tb", [(False, False), (False, True), (True, False), (True, True)])
def test_fake_shapes(ta, tb):
    R, C, D, nb = 512, 768, 136, 9
    bf = torch.bfloat16
    with FakeTensorMode():
        meta = (torch.empty(R // 128 + 1, dtype=torch.int32), torch.empty(nb, dtype=torch.int16),
                torch.empty(nb, dtype=torch.int16), torch.empty(C // 128 + 1, dtype=torch.int32),
                torch.empty(nb, dtype=torch.int16), torch.empty(nb, dtype=torch.int32), 0)
        a = torch.empty((D, R) if ta else (R, D), dtype=bf)
        b = torch.empty((C, D) if tb else (D, C), dtype=bf)
        v = ops.sdd_op(a, ta, b, tb, R, C, *meta)
        assert v.shape == (nb, 128, 128) and v.dtype == bf
        inner, outer = (R, C) if ta else (C, R)
        b = torch.empty((D, inner) if tb else (inner, D), dtype=bf)
        assert ops.dsd_op(v, ta, b, tb, R, C, *meta).shape == (outer, D)
        inner, outer = (C, R) if tb else (R, C)
        a = torch.empty((inner, D) if ta else (D, inner), dtype=bf)
        assert ops.dds_op(a, ta, v, tb, R, C, *meta).shape == (D, outer)
