"""Developer tool: time every named workload (BASELINE.json configs) on one B200 and print a
table — ours, the reference's own CUTLASS kernels (oracle/_ref, fp16, legacy mma.sync path) and
the dense torch.matmul ceiling at the same shape.  Writes JSON lines to gpurun_out/sweep.jsonl.

    python tools/sweep.py [--workloads c1,c2,...] [--iters 50] [--ref] [--dense]
"""
import argparse
import ctypes
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sputnik_b200 as sb  # noqa: E402
from sputnik_b200 import harness, workloads  # noqa: E402

DEFAULT = ["c1", "c2", "c3a", "c3b"] + \
    [f"c4_{op}_{d}" for op in ("sdd", "dsd", "dds") for d in (10, 25, 50, 75, 90)] + \
    ["c4_dsd_50_tn", "c4_dds_50_nt", "c4_sdd_50_nt", "c4_sdd_50_tn", "c4_sdd_50_tt",
     "c5_sdd_e8", "c5_dsd_e8", "c5_sdd_nt_e8", "c5_dsd_tn_e8", "c5_dsd_nt_e8", "c5_dds_tn_e8"]


GRAPH = False


def time_fn(fn, iters, warmup=5):
    """ms per call.  --graph: the `iters` calls are captured into one CUDA graph and the graph is
    replayed, which takes the host (Python + ctypes, ~10 us per call) out of short kernels."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if GRAPH:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=side):
                for _ in range(iters):
                    fn()
            g.replay()
            torch.cuda.synchronize()
            e0.record()
            for _ in range(3):
                g.replay()
            e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / (3 * iters)
    e0.record()
    for i in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def ref_call(lib, wl, s):
    p = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else None
    cur = lambda: ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    t = wl.topo
    sp = s.sparse
    if wl.op == "sdd":
        return lambda: lib.ref_sdd(s.a.rows, s.a.cols, p(s.a.data), int(wl.ta), s.b.rows, s.b.cols,
                                   p(s.b.data), int(wl.tb), t.rows, t.cols, t.nonzeros, p(sp.data),
                                   p(sp.offsets), p(sp.indices), p(sp.row_indices), cur())
    if wl.op == "dsd":
        return lambda: lib.ref_dsd(t.rows, t.cols, t.nonzeros, p(sp.data), p(sp.offsets),
                                   p(sp.indices), p(sp.offsets_t), p(sp.indices_t),
                                   p(sp.block_offsets), int(wl.ta), s.b.rows, s.b.cols, p(s.b.data),
                                   int(wl.tb), wl.m, wl.n, p(s.c.data), 0, cur())
    return lambda: lib.ref_dds(s.a.rows, s.a.cols, p(s.a.data), int(wl.ta), t.rows, t.cols,
                               t.nonzeros, p(sp.data), p(sp.offsets), p(sp.indices),
                               p(sp.offsets_t), p(sp.indices_t), p(sp.block_offsets), int(wl.tb),
                               wl.m, wl.n, p(s.c.data), 0, cur())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workloads", default=",".join(DEFAULT))
    ap.add_argument("--iters", type=int, default=30)
    ap.add_argument("--ref", action="store_true")
    ap.add_argument("--dense", action="store_true")
    ap.add_argument("--graph", action="store_true", help="time CUDA-graph replays of the calls")
    ap.add_argument("--no-hints", action="store_true",
                    help="call like an unmodified reference caller: no topology hints")
    ap.add_argument("--sets", type=int, default=0, help="operand sets to rotate (0 = enough to defeat L2)")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "sweep.jsonl"))
    args = ap.parse_args()
    global GRAPH
    GRAPH = args.graph
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) \
        if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"bf16_tflops": 1590.0, "hbm_gbs": 6650.0}
    lib = None
    if args.ref:
        path = os.path.join(ROOT, "oracle", "_ref", "libsputnik_ref.so")
        lib = ctypes.CDLL(path) if os.path.exists(path) else None
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    fout = open(args.out, "a")
    print(f"{'workload':18s} {'us':>9s} {'TFLOP/s':>9s} {'%peak':>6s} {'GB/s':>7s} | "
          f"{'ref us':>9s} {'ref TF':>8s} | {'dense TF':>8s}")
    for name in args.workloads.split(","):
        wl = workloads.get(name)
        set_bytes = wl.algorithmic_bytes
        nsets = max(1, min(4, -(-2 * 126 * 2**20 // set_bytes)))
        if set_bytes * nsets > 60e9:
            nsets = 1
        if args.sets:
            nsets = args.sets
        prob = harness.DeviceProblem(wl, "cuda:0", sets=nsets)
        if args.no_hints:
            for st in prob.sets:
                st.sparse.twin_rows = st.sparse.twin_cols = st.sparse.even_rows = False
        cnt = [0]

        def ours():
            prob.step(cnt[0])
            cnt[0] += 1
        ms = time_fn(ours, args.iters)
        tf = wl.flops / ms / 1e9
        row = {"workload": name, "desc": wl.describe(), "us": ms * 1e3, "tflops": tf,
               "frac_burst": tf / peaks["bf16_tflops"],
               "gbs": wl.algorithmic_bytes / ms / 1e6, "sets": nsets}
        ref_s = "        -        -"
        if lib is not None:
            wl16 = workloads.get(name)
            wl16.dtype = "f16"
            p16 = harness.DeviceProblem(wl16, "cuda:0", sets=1)
            f = ref_call(lib, wl16, p16.sets[0])
            rms = time_fn(f, max(3, args.iters // 3), warmup=2)
            row["ref_us"], row["ref_tflops"] = rms * 1e3, wl.flops / rms / 1e9
            ref_s = f"{rms * 1e3:9.1f} {row['ref_tflops']:8.1f}"
            del p16
        dense_s = "       -"
        if args.dense and wl.m * wl.k + wl.k * wl.n + wl.m * wl.n < 8e9:
            a = torch.randn(wl.m, wl.k, device="cuda:0", dtype=torch.bfloat16)
            b = torch.randn(wl.k, wl.n, device="cuda:0", dtype=torch.bfloat16)
            dms = time_fn(lambda: torch.matmul(a, b), max(3, args.iters // 3), warmup=2)
            row["dense_tflops"] = 2.0 * wl.m * wl.k * wl.n / dms / 1e9
            dense_s = f"{row['dense_tflops']:8.1f}"
            del a, b
        print(f"{name:18s} {ms * 1e3:9.1f} {tf:9.1f} {100 * row['frac_burst']:6.1f} "
              f"{row['gbs']:7.0f} | {ref_s} | {dense_s}", flush=True)
        fout.write(json.dumps(row) + "\n")
        fout.flush()
        del prob
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
