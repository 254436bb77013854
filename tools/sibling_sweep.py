"""Throughput of the SSD / SDS / DSS siblings on one B200.

    python tools/sibling_sweep.py [--size 4096] [--dens 25 50 75] [--reps 5] [--iters 20]

Random topologies (seeded), bf16, metadata prebuilt (MatmulEx semantics), CUDA events around
`iters` back-to-back launches after warm-up, median of `reps`.  FLOPs count only the block
products the operator must do: 2 * 128^3 per (output block, contributing K block) pair — for DSS
that is the size of the index intersections, computed here on the host from the topologies.
Beside each line: the same product through the path's own dense-output operator (DSD for SSD,
DDS for SDS; what a caller without the siblings would run before sampling)."""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sputnik_b200 as sb  # noqa: E402
from sputnik_b200 import synth  # noqa: E402

B = 128


def dev_sparse(topo, dev, gen, fill=True):
    v = torch.empty((topo.nnz_blocks, B, B), dtype=torch.float32, device=dev)
    if fill:
        v.uniform_(-1, 1, generator=gen)
    m = sb.BlockMatrix(topo.rows, topo.cols, B, topo.nonzeros, v.bfloat16(),
                       torch.from_numpy(topo.offsets).to(dev), torch.from_numpy(topo.indices).to(dev))
    sb.allocate_transpose_buffers(m)
    sb.transpose(m)
    sb.allocate_row_indices_buffer(m)
    sb.row_indices(m)
    return m


def dense(rows, cols, dev, gen):
    t = torch.empty((rows, cols), dtype=torch.float32, device=dev).uniform_(-1, 1, generator=gen)
    return sb.Matrix.of(t.bfloat16())


def block_sets(topo, by_cols):
    rows = [set() for _ in range(topo.block_cols if by_cols else topo.block_rows)]
    ri = topo.row_indices()
    for j in range(topo.nnz_blocks):
        r, c = int(ri[j]), int(topo.indices[j])
        if by_cols:
            rows[c].add(r)
        else:
            rows[r].add(c)
    return rows


def time_call(fn, iters, reps):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    out = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        out.append(e0.elapsed_time(e1) * 1e3 / iters)
    return float(np.median(out))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=4096)
    ap.add_argument("--dens", type=int, nargs="+", default=[25, 50, 75])
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--probe", action="store_true",
                    help="one cold launch of ssd_nn, sds_nn, dss_nn per density (for ncu captures)")
    ap.add_argument("--peak", type=float, default=None, help="bf16 dense TFLOP/s for the % column")
    a = ap.parse_args()
    peak = a.peak
    if peak is None:
        try:
            peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"])
        except Exception:
            peak = 1640.2  # MEASURED_PEAKS.json of this pool (burst) when the file did not travel
    dev = "cuda:0"
    gen = torch.Generator(device=dev)
    gen.manual_seed(0)
    n = a.size
    nbk = n // B
    print(f"# siblings, {n}^3, bf16, 128x128 blocks, peak {peak:.1f} TFLOP/s; us = median of {a.reps} x {a.iters} launches")
    print(f"{'op':<10}{'dens':>5}{'pairs':>8}{'us':>9}{'TFLOP/s':>9}{'%peak':>7} | {'dense-out op us':>16}{'x':>6}")
    for d in a.dens:
        nb = int(round(nbk * nbk * d / 100.0))
        t1 = synth.random_topology(nbk, nbk, nb, 1)
        t2 = synth.random_topology(nbk, nbk, nb, 2)
        s1, s2 = dev_sparse(t1, dev, gen), dev_sparse(t2, dev, gen)
        cs = dev_sparse(t2, dev, gen, fill=False)
        x = dense(n, n, dev, gen)
        y = dense(n, n, dev, gen)
        ri2 = t2.row_indices()
        if a.probe:
            sb.matmul_ex(s1, False, x, False, cs)
            sb.matmul_ex(x, False, s1, False, cs)
            sb.matmul_ex(s1, False, s2, False, y)
            torch.cuda.synchronize()
            print(f"probe: density {d}: ssd_nn, sds_nn, dss_nn launched once each")
            continue
        for tr in ("nn", "tt"):
            ta, tb = tr[0] == "t", tr[1] == "t"
            # op(A) row sets / op(B) column sets of the sparse inputs
            a_rows = block_sets(t1, ta)
            b_cols = block_sets(t1, not tb)
            # SSD: A = s1, C sampled at t2
            pairs = sum(len(a_rows[int(ri2[j])]) for j in range(t2.nnz_blocks))
            us = time_call(lambda: sb.matmul_ex(s1, ta, x, tb, cs), a.iters, a.reps)
            us_d = time_call(lambda: sb.matmul_ex(s1, ta, x, tb, y), a.iters, a.reps)
            line("ssd_" + tr, d, pairs, us, peak, us_d)
            # SDS: B = s1, C sampled at t2
            pairs = sum(len(b_cols[int(t2.indices[j])]) for j in range(t2.nnz_blocks))
            us = time_call(lambda: sb.matmul_ex(x, ta, s1, tb, cs), a.iters, a.reps)
            us_d = time_call(lambda: sb.matmul_ex(x, ta, s1, tb, y), a.iters, a.reps)
            line("sds_" + tr, d, pairs, us, peak, us_d)
            # DSS: A = s1, B = s2
            b2_cols = block_sets(t2, not tb)
            pairs = sum(len(ra & cb) for ra in a_rows for cb in b2_cols)
            us = time_call(lambda: sb.matmul_ex(s1, ta, s2, tb, y), a.iters, a.reps)
            line("dss_" + tr, d, pairs, us, peak, None)


def line(name, d, pairs, us, peak, us_dense):
    tf = 2.0 * pairs * B ** 3 / us * 1e-6
    extra = f"{us_dense:16.1f}{us_dense / us:6.2f}" if us_dense else f"{'-':>16}{'-':>6}"
    print(f"{name:<10}{d:>5}{pairs:>8}{us:9.1f}{tf:9.1f}{100 * tf / peak:7.1f} | {extra}", flush=True)


if __name__ == "__main__":
    main()
