"""First-light check of the CUDA path on a B200 (developer tool, not a test):
runs every op x transpose x dtype on small shapes against a torch fp32 matmul of the same
16-bit-rounded inputs and prints a per-case verdict with some failure diagnostics.

    python tools/gpu_selftest.py --op dsd
"""
import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sputnik_b200 as sb  # noqa: E402
from sputnik_b200 import synth  # noqa: E402

DEV = "cuda:0"


def dense_from_blocks(topo, vals):
    """vals [nb,128,128] torch -> dense [rows, cols] fp32 (block/matrix_utils.h:92-104)."""
    out = torch.zeros(topo.rows, topo.cols, dtype=torch.float32, device=vals.device)
    ri = topo.row_indices()
    for j in range(topo.nnz_blocks):
        r, c = int(ri[j]), int(topo.indices[j])
        out[r * 128:(r + 1) * 128, c * 128:(c + 1) * 128] = vals[j].float()
    return out


def make_sparse(topo, dtype, seed, with_values=True):
    nb = topo.nnz_blocks
    v = torch.from_numpy(synth.uniform_values((nb, 128, 128), seed)).to(DEV).to(dtype)
    if not with_values:
        v = torch.full_like(v, float("nan"))
    bm = sb.BlockMatrix(topo.rows, topo.cols, 128, topo.nonzeros, v,
                        torch.from_numpy(topo.offsets).to(DEV),
                        torch.from_numpy(topo.indices).to(DEV))
    return bm


def report(name, got, ref, extra=""):
    err = (got.float() - ref).abs()
    mx = float(err.max()) if err.numel() else 0.0
    bad = bool(torch.isnan(got.float()).any())
    ok = (mx <= 5e-2 * max(1.0, float(ref.abs().max()) / 16)) and not bad
    print(f"{'PASS' if ok else 'FAIL'} {name}: max_abs_err={mx:.4g} ref_max={float(ref.abs().max()):.3g} nan={bad} {extra}")
    if not ok and err.numel():
        e = err.cpu().numpy()
        R, C = e.shape
        # which 32-row x 64-col cells are wrong
        cells = []
        for r in range(0, min(R, 256), 32):
            row = ""
            for c in range(0, min(C, 512), 64):
                row += "X" if np.nanmax(e[r:r + 32, c:c + 64]) > 0.05 or np.isnan(e[r:r + 32, c:c + 64]).any() else "."
            cells.append(row)
        print("   bad-cell map (32 rows x 64 cols):")
        for row in cells:
            print("   ", row)
    return ok


def run_dsd(dtype, ta, tb, m_blocks, k_blocks, n, density, seed, unordered):
    # A sparse stored (ta ? k x m : m x k)
    sr, sc = (k_blocks, m_blocks) if ta else (m_blocks, k_blocks)
    nnz = max(1, int(round(sr * sc * density)))
    topo = synth.random_topology(sr, sc, nnz, seed, unordered)
    a = make_sparse(topo, dtype, seed + 1)
    sb.allocate_transpose_buffers(a)
    m, k = m_blocks * 128, k_blocks * 128
    bshape = (n, k) if tb else (k, n)
    bt = torch.from_numpy(synth.uniform_values(bshape, seed + 2)).to(DEV).to(dtype)
    ct = torch.randn(m, n, device=DEV).to(dtype)  # must be fully overwritten
    sb.matmul(a, ta, sb.Matrix.of(bt), tb, sb.Matrix.of(ct))
    torch.cuda.synchronize()
    ad = dense_from_blocks(topo, a.data)
    ref = (ad.t() if ta else ad) @ (bt.float().t() if tb else bt.float())
    return report(f"dsd {str(dtype)[6:]} ta={int(ta)} tb={int(tb)} m={m} k={k} n={n} nnz={nnz} unord={int(unordered)}", ct, ref)


def run_dds(dtype, ta, tb, m, k_blocks, n_blocks, density, seed, unordered):
    sr, sc = (n_blocks, k_blocks) if tb else (k_blocks, n_blocks)
    nnz = max(1, int(round(sr * sc * density)))
    topo = synth.random_topology(sr, sc, nnz, seed, unordered)
    b = make_sparse(topo, dtype, seed + 1)
    sb.allocate_transpose_buffers(b)
    k, n = k_blocks * 128, n_blocks * 128
    ashape = (k, m) if ta else (m, k)
    at = torch.from_numpy(synth.uniform_values(ashape, seed + 2)).to(DEV).to(dtype)
    ct = torch.randn(m, n, device=DEV).to(dtype)
    sb.matmul(sb.Matrix.of(at), ta, b, tb, sb.Matrix.of(ct))
    torch.cuda.synchronize()
    bd = dense_from_blocks(topo, b.data)
    ref = (at.float().t() if ta else at.float()) @ (bd.t() if tb else bd)
    return report(f"dds {str(dtype)[6:]} ta={int(ta)} tb={int(tb)} m={m} k={k} n={n} nnz={nnz} unord={int(unordered)}", ct, ref)


def run_sdd(dtype, ta, tb, m_blocks, k, n_blocks, density, seed, unordered):
    nnz = max(1, int(round(m_blocks * n_blocks * density)))
    topo = synth.random_topology(m_blocks, n_blocks, nnz, seed, unordered)
    c = make_sparse(topo, dtype, seed + 1, with_values=False)
    sb.allocate_row_indices_buffer(c)
    sb.row_indices(c)
    m, n = m_blocks * 128, n_blocks * 128
    at = torch.from_numpy(synth.uniform_values((k, m) if ta else (m, k), seed + 2)).to(DEV).to(dtype)
    bt = torch.from_numpy(synth.uniform_values((n, k) if tb else (k, n), seed + 3)).to(DEV).to(dtype)
    sb.matmul(sb.Matrix.of(at), ta, sb.Matrix.of(bt), tb, c)
    torch.cuda.synchronize()
    full = (at.float().t() if ta else at.float()) @ (bt.float().t() if tb else bt.float())
    ri = topo.row_indices()
    ok_ri = bool((c.row_indices.cpu().numpy() == ri).all())
    ref = torch.stack([full[int(ri[j]) * 128:(int(ri[j]) + 1) * 128,
                            int(topo.indices[j]) * 128:(int(topo.indices[j]) + 1) * 128]
                       for j in range(nnz)])
    return report(f"sdd {str(dtype)[6:]} ta={int(ta)} tb={int(tb)} m={m} k={k} n={n} nnz={nnz} unord={int(unordered)}",
                  c.data.reshape(nnz * 128, 128), ref.reshape(nnz * 128, 128), f"row_indices_ok={ok_ri}")


def run_transpose(seed, br, bc, nnz, unordered):
    topo = synth.random_topology(br, bc, nnz, seed, unordered)
    a = make_sparse(topo, torch.float16, seed)
    sb.allocate_transpose_buffers(a)
    sb.transpose(a)
    torch.cuda.synchronize()
    order = np.argsort(topo.indices, kind="stable")
    ri = topo.row_indices()
    exp_it, exp_bo = ri[order], order.astype(np.int32)
    exp_ot = np.zeros(bc + 1, dtype=np.int32)
    np.cumsum(np.bincount(topo.indices, minlength=bc), out=exp_ot[1:])
    ok = (np.array_equal(a.offsets_t.cpu().numpy(), exp_ot)
          and np.array_equal(a.indices_t.cpu().numpy(), exp_it)
          and np.array_equal(a.block_offsets.cpu().numpy(), exp_bo))
    print(f"{'PASS' if ok else 'FAIL'} transpose br={br} bc={bc} nnz={nnz} unord={int(unordered)}")
    return ok


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--op", default="dsd")
    ap.add_argument("--quick", action="store_true")
    args = ap.parse_args()
    torch.cuda.init()
    print("device:", torch.cuda.get_device_name(0), "lib:", sb._lib.lib.bsmm_version().decode())
    dtypes = [torch.bfloat16, torch.float16]
    allok = True
    if args.op == "transpose":
        for (br, bc, nnz, un) in [(4, 4, 8, False), (8, 8, 32, True), (64, 256, 2048, False),
                                  (40, 3, 100, True), (512, 896, 57344, False), (3, 5, 0 + 1, False)]:
            allok &= run_transpose(7, br, bc, nnz, un)
    elif args.op == "dsd":
        first = True
        for dt in dtypes:
            for ta in (False, True):
                for tb in (False, True):
                    allok &= run_dsd(dt, ta, tb, 2, 3, 256, 0.5, 11, False)
                    if first and args.quick:
                        return
                    first = False
                    allok &= run_dsd(dt, ta, tb, 4, 4, 512, 0.4, 12, True)
                    allok &= run_dsd(dt, ta, tb, 2, 2, 8, 0.5, 13, False)
                    allok &= run_dsd(dt, ta, tb, 3, 2, 136, 0.5, 14, True)
        allok &= run_dsd(torch.bfloat16, False, False, 16, 16, 1024, 0.25, 15, False)
    elif args.op == "dds":
        for dt in dtypes:
            for ta in (False, True):
                for tb in (False, True):
                    allok &= run_dds(dt, ta, tb, 256, 3, 2, 0.5, 21, False)
                    allok &= run_dds(dt, ta, tb, 512, 4, 4, 0.4, 22, True)
                    allok &= run_dds(dt, ta, tb, 8, 2, 2, 0.5, 23, False)
                    allok &= run_dds(dt, ta, tb, 136, 2, 3, 0.5, 24, True)
        allok &= run_dds(torch.bfloat16, False, False, 1024, 16, 16, 0.25, 25, False)
    elif args.op == "sdd":
        for dt in dtypes:
            for ta in (False, True):
                for tb in (False, True):
                    allok &= run_sdd(dt, ta, tb, 2, 384, 2, 0.5, 31, False)
                    allok &= run_sdd(dt, ta, tb, 4, 512, 4, 0.4, 32, True)
                    allok &= run_sdd(dt, ta, tb, 2, 8, 2, 0.5, 33, False)
                    allok &= run_sdd(dt, ta, tb, 3, 136, 2, 0.5, 34, True)
        allok &= run_sdd(torch.float16, False, False, 8, 1024, 8, 0.5, 35, False)
    print("ALL PASS" if allok else "SOME FAILED")
    sys.exit(0 if allok else 1)


if __name__ == "__main__":
    main()
