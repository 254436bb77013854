"""Device-side staging of a Workload for benchmarks and tools: operand sets resident in HBM,
one `step(i)` = one call of the public operator (sputnik_b200.matmul_ex -> C-ABI -> CUDA)."""
import torch

from . import block as sb

BLOCK = 128


def _dt(dtype):
    if dtype == "f8":
        return torch.float8_e4m3fn
    return torch.float16 if dtype == "f16" else torch.bfloat16


def _odt(dtype):
    """Output type: e4m3 operands produce bf16."""
    return torch.bfloat16 if dtype == "f8" else _dt(dtype)


def _uniform(shape, dtype, dev, gen):
    t = torch.empty(shape, dtype=torch.float32, device=dev)
    t.uniform_(-1.0, 1.0, generator=gen)
    return t.to(_dt(dtype))


class OperandSet:
    """One set of operands + output for a workload, on `dev`."""

    def __init__(self, wl, dev, gen, host=None):
        t, dt = wl.topo, wl.dtype
        self.offsets = torch.from_numpy(t.offsets).to(dev)
        self.indices = torch.from_numpy(t.indices).to(dev)
        shapes = wl.dense_shapes()
        mk = lambda key, shape: (host[key].to(dev, non_blocking=True) if host is not None
                                 else _uniform(shape, dt, dev, gen))
        vals_shape = (t.nnz_blocks, BLOCK, BLOCK)
        if wl.op == "sdd":
            vals = torch.empty(vals_shape, dtype=_odt(dt), device=dev)
        else:
            vals = mk("values", vals_shape)
        self.sparse = sb.BlockMatrix(t.rows, t.cols, BLOCK, t.nonzeros, vals, self.offsets,
                                     self.indices)
        self.sparse.__dict__.update(t.hints())
        self.a = sb.Matrix.of(mk("a", shapes["a"])) if shapes["a"] else None
        self.b = sb.Matrix.of(mk("b", shapes["b"])) if shapes["b"] else None
        self.c = (sb.Matrix.of(torch.empty(shapes["c"], dtype=_odt(dt), device=dev))
                  if shapes["c"] else None)
        if wl.op == "sdd":
            sb.allocate_row_indices_buffer(self.sparse)
            sb.row_indices(self.sparse)
        else:
            sb.allocate_transpose_buffers(self.sparse)
            sb.transpose(self.sparse)  # prebuilt: steps use MatmulEx semantics

    def output(self):
        return self.sparse.data if self.c is None else self.c.data


class DeviceProblem:
    """`sets` rotating operand sets so consecutive steps never hit the same lines in L2."""

    def __init__(self, wl, dev="cuda:0", sets=1, seed=0):
        self.wl, self.dev = wl, dev
        gen = torch.Generator(device=dev)
        gen.manual_seed(seed)
        self.sets = [OperandSet(wl, dev, gen) for _ in range(sets)]

    def call(self, s, with_metadata=False, stream=None):
        wl = self.wl
        mm = sb.matmul if with_metadata else sb.matmul_ex
        if wl.op == "sdd":
            mm(s.a, wl.ta, s.b, wl.tb, s.sparse, stream)
        elif wl.op == "dsd":
            mm(s.sparse, wl.ta, s.b, wl.tb, s.c, stream)
        else:
            mm(s.a, wl.ta, s.sparse, wl.tb, s.c, stream)

    def step(self, i, with_metadata=False):
        self.call(self.sets[i % len(self.sets)], with_metadata)


def host_operands(wl, seed=0, pin=True):
    """Pinned host tensors of the step's INPUTS (for the end-to-end leg) + a pinned result."""
    g = torch.Generator()
    g.manual_seed(seed)
    t = wl.topo
    shapes = wl.dense_shapes()

    def mk(shape):
        x = torch.empty(shape, dtype=torch.float32).uniform_(-1.0, 1.0, generator=g).to(_dt(wl.dtype))
        return x.pin_memory() if pin else x
    host = {}
    if wl.op != "sdd":
        host["values"] = mk((t.nnz_blocks, BLOCK, BLOCK))
    if shapes["a"]:
        host["a"] = mk(shapes["a"])
    if shapes["b"]:
        host["b"] = mk(shapes["b"])
    out_shape = shapes["c"] if shapes["c"] else (t.nnz_blocks, BLOCK, BLOCK)
    result = torch.empty(out_shape, dtype=_odt(wl.dtype))
    return host, (result.pin_memory() if pin else result)
