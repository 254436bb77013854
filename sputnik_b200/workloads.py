"""Named synthetic workloads = the configs of BASELINE.json, host side only.

A workload is one block-sparse matmul: op, transpose flags, dtype, dense extents and a seeded
BCSR topology.  FLOPs count nonzero blocks only: 2 * nb * 128^2 * D with D = K (SDD), N (DSD),
M (DDS) — the definition of the reference's benchmarks (dsd_benchmark.cu:113-114).
Algorithmic bytes count every operand once."""
from . import synth

BLOCK = 128


class Workload:
    def __init__(self, name, op, ta, tb, dtype, m, k, n, topo, note=""):
        self.name, self.op, self.ta, self.tb, self.dtype = name, op, bool(ta), bool(tb), dtype
        self.m, self.k, self.n, self.topo, self.note = m, k, n, topo, note

    @property
    def nb(self):
        return self.topo.nnz_blocks

    @property
    def flops(self):
        d = {"sdd": self.k, "dsd": self.n, "dds": self.m}[self.op]
        return 2.0 * self.nb * BLOCK * BLOCK * d

    @property
    def algorithmic_bytes(self):
        es = 1 if self.dtype == "f8" else 2          # operand bytes per element; outputs are 16-bit
        t = self.topo
        meta = 4 * (t.block_rows + 1) + 2 * self.nb
        if self.op == "sdd":
            return es * (self.m * self.k + self.k * self.n) + 2 * self.nb * BLOCK * BLOCK + meta + 2 * self.nb
        sp = es * self.nb * BLOCK * BLOCK
        transposed_walk = (self.op == "dsd" and self.ta) or (self.op == "dds" and not self.tb)
        if transposed_walk:
            meta = 4 * (t.block_cols + 1) + 2 * self.nb + 4 * self.nb
        if self.op == "dsd":
            return sp + es * self.k * self.n + 2 * self.m * self.n + meta
        return es * self.m * self.k + sp + 2 * self.m * self.n + meta

    def dense_shapes(self):
        """Stored shapes of the dense operands / output: dict a, b, c (None where sparse)."""
        a = (self.k, self.m) if self.ta else (self.m, self.k)
        b = (self.n, self.k) if self.tb else (self.k, self.n)
        c = (self.m, self.n)
        return {"a": None if self.op == "dsd" else a, "b": None if self.op == "dds" else b,
                "c": None if self.op == "sdd" else c}

    def describe(self):
        return (f"{self.name}: {self.op.upper()} {'t' if self.ta else 'n'}{'t' if self.tb else 'n'} "
                f"{self.dtype} m={self.m} k={self.k} n={self.n} nnz_blocks={self.nb} "
                f"({100.0 * self.nb / (self.topo.block_rows * self.topo.block_cols):.1f}% dense)")


def _moe(experts, tokens, hidden, ffn):
    """dMoE topology: tokens routed uniformly, expert e owns tokens/experts rows and ffn cols."""
    return synth.moe_topology(experts, tokens // experts // BLOCK, ffn // BLOCK)


def get(name, seed=0):
    """`<name>@f8`: the same workload with e4m3 operands and bf16 output (BSMM_F8E4M3).
    c1 | c2 | c3a | c3b | c4_<op>_<density%>[_<nn|nt|tn|tt>] | c5_<sdd|dsd|sdd_nt|dsd_tn|dsd_nt|dds_tn>[_e<experts>]"""
    if name.endswith("@f8"):
        wl = get(name[:-3], seed)
        wl.name, wl.dtype = name, "f8"
        return wl
    if name == "c1":  # SDD 1024^3, 50% random, fp16 (sdd_test.cu:95)
        return Workload(name, "sdd", 0, 0, "f16", 1024, 1024, 1024,
                        synth.random_topology(8, 8, 32, seed))
    if name in ("c2", "c3a", "c3b"):
        topo = _moe(8, 8192, 1024, 4096)  # sparse 8192 x 32768, nb = 2048
        if name == "c2":   # y = dsd(h, w2): C 8192 x 1024
            return Workload(name, "dsd", 0, 0, "bf16", 8192, 32768, 1024, topo, "dMoE FFN fwd")
        if name == "c3a":  # dw2 = dsd(h^T, dout): C 32768 x 1024
            return Workload(name, "dsd", 1, 0, "bf16", 32768, 8192, 1024, topo, "dMoE bwd dw2")
        return Workload(name, "dds", 1, 0, "bf16", 1024, 8192, 32768, topo, "dMoE bwd dw1")
    if name.startswith("c4_"):
        parts = name.split("_")
        op, dens = parts[1], int(parts[2])
        tr = parts[3] if len(parts) > 3 else "nn"
        nb = int(round(1024 * dens / 100.0))
        return Workload(name, op, tr[0] == "t", tr[1] == "t", "bf16", 4096, 4096, 4096,
                        synth.random_topology(32, 32, nb, seed))
    if name.startswith("c5_"):
        parts = name[3:].split("_")
        experts = 64
        if parts[-1].startswith("e"):
            experts = int(parts[-1][1:])
            parts = parts[:-1]
        kind = "_".join(parts)
        tokens, hidden, ffn = 1024 * experts, 4096, 14336
        topo = _moe(experts, tokens, hidden, ffn)  # tokens x (experts*ffn)
        F = experts * ffn
        table = {
            "sdd": ("sdd", 0, 0, tokens, hidden, F),       # h = sdd(x, w1)
            "dsd": ("dsd", 0, 0, tokens, F, hidden),       # y = dsd(h, w2)
            "sdd_nt": ("sdd", 0, 1, tokens, hidden, F),    # dh = sdd(dy, w2^T)
            "dsd_tn": ("dsd", 1, 0, F, tokens, hidden),    # dw2 = dsd(h^T, dy)
            "dsd_nt": ("dsd", 0, 1, tokens, F, hidden),    # dx = dsd(dh, w1^T)
            "dds_tn": ("dds", 1, 0, hidden, tokens, F),    # dw1 = dds(x^T, dh)
        }
        op, ta, tb, m, k, n = table[kind]
        return Workload(name, op, ta, tb, "bf16", m, k, n, topo, "Mixtral-like dMoE")
    if name.startswith("m_"):  # m_<dsd|dsdt|sdd>_<experts>_<rows/128 per expert>_<ffn/128>_<hidden>
        parts = name.split("_")
        kind, e, r, f, hidden = parts[1], int(parts[2]), int(parts[3]), int(parts[4]), int(parts[5])
        topo = synth.moe_topology(e, r, f)
        tokens, F = e * r * BLOCK, e * f * BLOCK
        table = {"dsd": ("dsd", 0, 0, tokens, F, hidden), "dsdt": ("dsd", 0, 1, tokens, F, hidden),
                 "sdd": ("sdd", 0, 0, tokens, hidden, F)}
        op, ta, tb, m, k, n = table[kind]
        return Workload(name, op, ta, tb, "bf16", m, k, n, topo, "dMoE-shaped probe")
    if name.startswith("x_"):  # x_<op>_<m>_<k>_<n>_<nb>[_<nn|nt|tn|tt>]: random topology, bf16
        parts = name.split("_")
        op, m, k, n, nb = parts[1], int(parts[2]), int(parts[3]), int(parts[4]), int(parts[5])
        tr = parts[6] if len(parts) > 6 else "nn"
        ta, tb = tr[0] == "t", tr[1] == "t"
        if op == "sdd":
            sr, sc = m, n
        elif op == "dsd":
            sr, sc = (k, m) if ta else (m, k)
        else:
            sr, sc = (n, k) if tb else (k, n)
        return Workload(name, op, ta, tb, "bf16", m, k, n,
                        synth.random_topology(sr // BLOCK, sc // BLOCK, nb, seed))
    raise KeyError(name)
