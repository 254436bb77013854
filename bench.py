"""bench.py — the driver's benchmark contract for the block-sparse matmul path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

Default workload (every N, including 1) = BASELINE.json configs[4], the largest configuration that
fits one B200: the Mixtral-like dMoE FFN layer — 64 experts, hidden 4096, ffn 14336, 65 536 routed
token rows, SDD -> DSD forward + backward (6 block-sparse matmuls per step) — sharded BY EXPERT
over the N ranks (strong scaling: the same 64 experts / 65 536 rows at every N).  One "step" = one
forward + backward pass of the layer through the public operators (sputnik_b200.moe /
sputnik_b200.block -> C-ABI -> CUDA).

The one JSON line of rank 0 carries, besides the contract's keys:
  roofline      the dominant kernel of the step, timed with CUDA events inside the timed region;
  alt           the same layer with the NCCL all-to-all exchange instead of the fused peer-memory one;
  parity        (every N) ragged-routing check before the timed region: exchange bit-exact, 6 products
                per expert <= 1e-2 vs fp64, fused path bit-equal to the NCCL path — the run FAILS otherwise;
  e2e           token rows in / y, dx out through pinned host memory, every step;
  configs       (N = 1) one record per other BASELINE config: C1, C2, C3a / C3b (MatmulEx and
                including the device-side metadata build), C4 {10..90} % x {SDD, DSD, DDS};
                (N > 1) the C4 products sharded by nnz-balanced block-row slabs;
  legacy_gpu    (N = 1) the reference's own CUTLASS kernels (oracle/_ref/libsputnik_ref.so, fp16)
                with the reference's benchmark method (10 warm-up + 100 timed, dsd_benchmark.cu:82-107);
  ragged        (every N) the layer under a Zipf-like routing whose counts live only on the device;
  cpu_baseline  (N = 1) the oracle port on the host cores on a bounded sample.

`--workload c2` (or any name of sputnik_b200/workloads.py) times one product instead.
`--impl reference` times the reference's own host implementation (oracle/_ref, sputnik::Matrix
operator* / T() / Mul(), compiled from the reference's sources) on all host threads on a bounded
sample of the same workload.  See DESIGN.md "Measurement".
"""
import argparse
import ctypes
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "block-sparse matmul effective TFLOP/s (nnz blocks) and % of BF16 tensor peak"
UNIT = "TFLOP/s"
L2_BYTES = 126 * 1024 * 1024
PROFILE_ROUND = "r02"


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"bf16_burst": d["bf16_tflops"], "bf16_sustained": d["bf16_tflops_sustained"],
                "hbm_gbs": d["hbm_gbs"], "source": "measured"}
    return {"bf16_burst": 1590.0, "bf16_sustained": 1400.0, "hbm_gbs": 6650.0,
            "source": "fallback"}


def ncu_traffic(kernel_key):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, parsed at
    run time from the committed `ncu --set full` summary profiles/<round>_ncu_traffic.json
    ({"<key>": {"bytes": ..., "source": "profiles/..."}}); (None, None) when not captured."""
    p = os.path.join(ROOT, "profiles", f"{PROFILE_ROUND}_ncu_traffic.json")
    try:
        rec = json.load(open(p)).get(kernel_key)
        return (int(rec["bytes"]), rec["source"]) if rec else (None, None)
    except Exception:
        return None, None


class ClockSampler(threading.Thread):
    """Samples SM clock + throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop_evt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            pass

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80): "hw_power_brake",
        }
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.002)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


def physical_gpu_index(local_rank):
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local_rank])
        except Exception:
            return local_rank
    return local_rank


def pin_to_gpu_numa(local_rank):
    """Bind this rank's host threads to the CPUs NVML reports as local to its GPU and, where the
    cpuset confines every rank to one socket (this pool: cpus 0-31 / node 0 for all eight ranks),
    at least bind its MEMORY policy to the GPU's NUMA node (sysfs numa_node of the PCI device,
    set_mempolicy(MPOL_PREFERRED)), so that the pinned staging buffers of ranks 4-7 do not all sit
    behind one socket's memory controller and inter-socket link — what made the end-to-end leg
    collapse at N = 8.  Best effort: returns what it did."""
    did = []
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(physical_gpu_index(local_rank))
        try:
            words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
            cpus = {64 * i + b for i, w in enumerate(words) for b in range(64) if (int(w) >> b) & 1}
            allowed = os.sched_getaffinity(0)
            use = cpus & allowed
            if use and use != allowed:
                os.sched_setaffinity(0, use)
                did.append(f"{len(use)} cpus local to the gpu")
        except Exception as e:  # noqa: BLE001
            did.append(f"cpu affinity unavailable ({type(e).__name__})")
        bus = pynvml.nvmlDeviceGetPciInfo(h).busId
        bus = (bus.decode() if isinstance(bus, bytes) else bus).lower()
        if bus.startswith("00000000:"):
            bus = bus[4:]
        node = int(open(f"/sys/bus/pci/devices/{bus}/numa_node").read())
        if node >= 0:
            libc = ctypes.CDLL(None, use_errno=True)
            mask = ctypes.c_ulong(1 << node)
            MPOL_PREFERRED = 1
            rc = libc.syscall(238, MPOL_PREFERRED, ctypes.byref(mask), ctypes.c_ulong(64))
            did.append(f"memory policy: prefer numa node {node}" if rc == 0
                       else f"set_mempolicy(node {node}) failed, errno {ctypes.get_errno()}")
        else:
            did.append("gpu numa node unknown")
    except Exception as e:  # noqa: BLE001
        did.append(f"numa binding unavailable ({type(e).__name__}: {e})")
    return "; ".join(did) if did else "unchanged"


# =========================================================================== workload description
def parse_layer(name):
    """c5_layer[_nccl][_e<experts>][_t<tokens>]: Mixtral-like dMoE FFN fwd+bwd, expert-sharded
    over ranks.  (`_fused` is accepted and ignored: the fused exchange is the default `value`.)"""
    experts, tokens = 64, None
    for part in name.split("_")[2:]:
        if part in ("fused", "nccl"):
            continue
        if part.startswith("e"):
            experts = int(part[1:])
        elif part.startswith("t"):
            tokens = int(part[1:])
    return experts, tokens if tokens is not None else 1024 * experts, 4096, 14336


def layer_flops(experts, tokens, hidden, ffn):
    """Nonzero-block flops of one forward + backward step (uniform routing, rows already a
    multiple of 128 per expert): 6 products x 2 * nb * 128^2 * hidden, nb = tokens/128 * ffn/128."""
    nb = (tokens // 128) * (ffn // 128)
    return 6 * 2.0 * nb * 128 * 128 * hidden


def layer_config(name, world):
    """The `config` dict — built by ONE function for both arms so the driver's same_config holds."""
    experts, tokens, hidden, ffn = parse_layer(name)
    flops = layer_flops(experts, tokens, hidden, ffn)
    return {
        "workload": f"c5_layer: Mixtral-like dMoE FFN forward+backward, {experts} experts, hidden "
                    f"{hidden}, ffn {ffn}, {tokens} routed token rows (uniform), 6 block-sparse "
                    f"matmuls per step (h=sdd(x,w1^T), y=dsd(h,w2), dh=sdd(dy,w2^T), dw2=dsd(h^T,dy), "
                    f"dx=dsd(dh,w1), dw1=dsd(dh^T,x)), "
                    f"{flops / 1e12:.1f} TFLOP per step over nonzero blocks",
        "w1_layout": "fh: w1 stored [experts*ffn, hidden] (MegaBlocks' storage, sdd(x, w1.t())); the "
                     "[hidden, experts*ffn] storage (sdd nn, dsd nt, dds tn) is timed beside it as "
                     "alt_layout at N = 1",
        "sharding": f"by expert over {world} rank(s): {experts // world} experts and "
                    f"{tokens // world} token rows per rank; one token exchange each way, forward "
                    f"and backward; weight gradients stay local",
        "l2": "per-rank working set (weights, gradients, sparse activations) >> 126 MB L2; no flush needed",
        "metadata": "topology, transposed metadata and row_indices prebuilt (MatmulEx semantics)",
    }


def product_config(wl_desc, note):
    return {"workload": wl_desc, "note": note,
            "l2": "inputs rotate over operand sets larger than the 126 MB L2 in total; no flush needed",
            "metadata": "transposed metadata / row_indices prebuilt (MatmulEx semantics)"}


# =========================================================================== CPU legs (oracle/)
def _sub_topology(topo, block_rows):
    """First `block_rows` block rows of a topology (a bounded sample of the same workload)."""
    from sputnik_b200 import synth
    br = min(block_rows, topo.block_rows)
    nb = int(topo.offsets[br])
    return synth.Topology(br, topo.block_cols, topo.offsets[:br + 1].copy(), topo.indices[:nb].copy())


def cpu_sample(wl, block_rows):
    """Host operands of a bounded sample: the first `block_rows` block rows of the SPARSE matrix
    (as stored) with the full dense operand(s).  Returns (callable running the oracle port,
    effective flops of the sample, description)."""
    from oracle import oracle
    from sputnik_b200 import synth
    sub = _sub_topology(wl.topo, block_rows)
    nb = sub.nnz_blocks
    r = lambda shape, s: oracle.round_to(synth.uniform_values(shape, 1000 + s), wl.dtype)
    if wl.op == "dsd":
        vals = r((nb, 128, 128), 1)
        kdim = sub.rows if wl.ta else sub.cols
        b = r((wl.n, kdim) if wl.tb else (kdim, wl.n), 2)
        fn = lambda: oracle.dsd(vals, sub, wl.ta, b, wl.tb)
        flops = 2.0 * nb * 128 * 128 * wl.n
    elif wl.op == "dds":
        vals = r((nb, 128, 128), 1)
        kdim = sub.cols if wl.tb else sub.rows
        a = r((kdim, wl.m) if wl.ta else (wl.m, kdim), 2)
        fn = lambda: oracle.dds(a, wl.ta, vals, sub, wl.tb)
        flops = 2.0 * nb * 128 * 128 * wl.m
    else:
        a = r((wl.k, sub.rows) if wl.ta else (sub.rows, wl.k), 2)
        b = r((wl.n, wl.k) if wl.tb else (wl.k, wl.n), 3)
        fn = lambda: oracle.sdd(a, wl.ta, b, wl.tb, sub)
        flops = 2.0 * nb * 128 * 128 * wl.k
    desc = (f"first {sub.block_rows} of {wl.topo.block_rows} block rows of the sparse operand "
            f"({nb} of {wl.nb} nonzero blocks), full dense operand")
    return fn, flops, desc


def pick_cpu_rows(wl, budget_s=12.0, gmacs=6.0):
    """Block rows whose oracle run should take about `budget_s` at ~`gmacs` GMAC/s."""
    per_block = 128 * 128 * {"sdd": wl.k, "dsd": wl.n, "dds": wl.m}[wl.op] / 1e9
    blocks = max(1.0, budget_s * gmacs / per_block)
    per_row = max(1.0, wl.nb / wl.topo.block_rows)
    return int(max(1, min(wl.topo.block_rows, round(blocks / per_row))))


def time_cpu_port(wl, budget_s=12.0):
    from oracle import oracle
    fn, flops, desc = cpu_sample(wl, pick_cpu_rows(wl, budget_s))
    t0 = time.perf_counter()
    fn()
    dt = time.perf_counter() - t0
    return {"value": flops / dt / 1e12, "unit": UNIT, "cores": oracle.num_threads(), "kind": "port",
            "sample": desc + f"; oracle block-aware OpenMP restatement of the reference host "
                             f"matmul, {dt:.1f} s"}


def time_cpu_port_layer(experts, tokens, hidden, ffn, budget_s=14.0):
    """The oracle port on the six products of the layer for the first block rows of expert 0
    (block-diagonal: an expert's products are dense [rows, hidden] x [hidden, ffn] shapes walked
    block by block), all host cores."""
    import numpy as np
    from oracle import oracle
    from sputnik_b200 import synth
    fc = ffn // 128
    # 6 products x 128 * ffn * hidden MACs per block row at ~25 GMAC/s on a 16-core box
    per_row = 6 * 128 * ffn * hidden / 1e9
    br = int(max(1, min(tokens // experts // 128, round(budget_s * 25.0 / per_row))))
    topo = synth.moe_topology(1, br, fc)
    r = lambda shape, s: oracle.round_to(synth.uniform_values(shape, 2000 + s), "bf16")
    x, dy = r((br * 128, hidden), 1), r((br * 128, hidden), 2)
    w1, w2 = r((ffn, hidden), 3), r((ffn, hidden), 4)   # w1 stored [ffn, hidden] (layer_config)
    t0 = time.perf_counter()
    h = oracle.round_to(oracle.sdd(x, False, w1, True, topo), "bf16")         # h = sdd(x, w1^T)
    oracle.dsd(h, topo, False, w2, False)                                     # y = dsd(h, w2)
    dh = oracle.round_to(oracle.sdd(dy, False, w2, True, topo), "bf16")       # dh = sdd(dy, w2^T)
    oracle.dsd(h, topo, True, dy, False)                                      # dw2 = dsd(h^T, dy)
    oracle.dsd(dh, topo, False, w1, False)                                    # dx = dsd(dh, w1)
    oracle.dsd(dh, topo, True, x, False)                                      # dw1 = dsd(dh^T, x)
    dt = time.perf_counter() - t0
    flops = 6 * 2.0 * br * fc * 128 * 128 * hidden
    return {"value": flops / dt / 1e12, "unit": UNIT, "cores": oracle.num_threads(), "kind": "port",
            "sample": f"all six products of the layer on the first {br} of {tokens // 128} block "
                      f"rows (expert 0 of {experts}: {br * fc} of {tokens // 128 * fc} nonzero blocks "
                      f"per product, full hidden / ffn extents); oracle block-aware OpenMP "
                      f"restatement of the reference host matmul, {dt:.1f} s"}


# ----------------------------------------------------------------- the reference's own host code
class RefHost:
    """oracle/_ref/libsputnik_ref_host.so: sputnik::Matrix (operator*, T(), Mul()) compiled from the
    reference's sources.  Operands become sputnik::Matrix handles ONCE (outside any timed loop);
    mm / t / mul are the reference's own functions and nothing else."""

    def __init__(self):
        import numpy as np
        path = os.path.join(ROOT, "oracle", "_ref", "libsputnik_ref_host.so")
        self.h = h = ctypes.CDLL(path)
        f = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
        vp, i = ctypes.c_void_p, ctypes.c_int
        h.ref_host_new.argtypes, h.ref_host_new.restype = [f, i, i], vp
        h.ref_host_free.argtypes = [vp]
        h.ref_host_mm.argtypes, h.ref_host_mm.restype = [vp, vp], vp
        h.ref_host_t.argtypes, h.ref_host_t.restype = [vp], vp
        h.ref_host_mul_inplace.argtypes = [vp, vp]
        self.np = np

    @staticmethod
    def available():
        return os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libsputnik_ref_host.so"))

    def new(self, arr):
        a = self.np.ascontiguousarray(arr, dtype=self.np.float32)
        return self.h.ref_host_new(a, a.shape[0], a.shape[1])

    def product(self, a, ta, b, tb, mask=None):
        """The reference tests' expectation (dsd_test.cu:185-190, sdd_test.cu:170-193):
        (ta ? A.T() : A) * (tb ? B.T() : B), then Mul(mask) for sparse outputs.  Frees every
        temporary; returns nothing (the result is discarded, as in a benchmark)."""
        h = self.h
        at = h.ref_host_t(a) if ta else None
        bt = h.ref_host_t(b) if tb else None
        c = h.ref_host_mm(at or a, bt or b)
        if mask is not None:
            h.ref_host_mul_inplace(c, mask)
        for t in (at, bt, c):
            if t:
                h.ref_host_free(t)


def _run_threads(fns):
    ts = [threading.Thread(target=f) for f in fns]
    [t.start() for t in ts]
    [t.join() for t in ts]


def _import_host_only():
    """synth / workloads WITHOUT loading libbsmm.so (the reference arm must not map the product
    library): pre-seed the package so that its __init__ (which loads the library) never runs."""
    import types
    if "sputnik_b200" not in sys.modules:
        pkg = types.ModuleType("sputnik_b200")
        pkg.__path__ = [os.path.join(ROOT, "sputnik_b200")]
        sys.modules["sputnik_b200"] = pkg
    from sputnik_b200 import synth, workloads  # noqa: F401
    return synth, workloads


def reference_line(args, config, dtype, value, dt, cores, kind, sample, scaling):
    return {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT,
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": scaling,
            "vs_baseline": None, "dtype": dtype, "data": "synthetic", "config": config,
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind,
                             "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}


def _timed_steps(args, one_step):
    for _ in range(args.warmup):
        one_step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        one_step()
    return (time.perf_counter() - t0) / max(1, args.steps)


def reference_layer_arm(args, world):
    """--impl reference on the default workload: the six products of the layer through the
    reference's OWN host code, every host thread on its own bounded sample (one 128-row block row
    of one expert x `cb` ffn blocks x `hs` hidden columns), each product formed exactly as the
    reference's tests form their expectation (densified sparse operand, T() for transpose flags,
    operator*, Mul(mask) for the sampled outputs)."""
    import numpy as np
    synth, _ = _import_host_only()
    from oracle import oracle
    experts, tokens, hidden, ffn = parse_layer(args.workload)
    cores = os.cpu_count() or 1
    config = layer_config(args.workload, world)
    budget = min(2.0, 100.0 / max(1, args.steps + args.warmup))
    r, cb = 128, 2
    c = 128 * cb
    rnd = lambda shape, s: oracle.round_to(synth.uniform_values(shape, 3000 + s), "bf16")
    if RefHost.available():
        ref = RefHost()
        kind = "reference"

        def make(hs, seed):
            x, dy = ref.new(rnd((r, hs), seed)), ref.new(rnd((r, hs), seed + 1))
            w1, w2 = ref.new(rnd((c, hs), seed + 2)), ref.new(rnd((c, hs), seed + 3))
            # h / dh: densified block-sparse activations (1 block row x cb blocks = dense [r, c])
            hd, dhd = ref.new(rnd((r, c), seed + 4)), ref.new(rnd((r, c), seed + 5))
            mask = ref.new(np.ones((r, c), np.float32))

            def run():
                ref.product(x, False, w1, True, mask)      # h = sdd(x, w1^T)
                ref.product(hd, False, w2, False)          # y = dsd(h, w2)
                ref.product(dy, False, w2, True, mask)     # dh = sdd(dy, w2^T)
                ref.product(hd, True, dy, False)           # dw2 = dsd(h^T, dy)
                ref.product(dhd, False, w1, False)         # dx = dsd(dh, w1)
                ref.product(dhd, True, x, False)           # dw1 = dsd(dh^T, x)
            return run
        probe = [make(64, 10 * i) for i in range(cores)]
        _run_threads(probe)
        t0 = time.perf_counter()
        _run_threads(probe)
        per_col = (time.perf_counter() - t0) / 64
        hs = int(max(8, min(hidden, (budget / per_col) // 8 * 8)))
        runs = [make(hs, 10 * i) for i in range(cores)]
        one_step = lambda: _run_threads(runs)
        flops = cores * 6 * 2.0 * r * c * hs
        sample = (f"{cores} host threads, each the six products of the layer on its own sample "
                  f"(1 block row of one expert = {r} token rows, {cb} of {ffn // 128} ffn blocks, "
                  f"{hs} of {hidden} hidden columns) through the reference's sputnik::Matrix "
                  f"operator* / T() / Mul() (oracle/_ref, as written: dense, single-threaded per "
                  f"call); operands converted to sputnik::Matrix outside the timed loop")
    else:
        kind, cores = "port", oracle.num_threads()
        topo = synth.moe_topology(1, 1, ffn // 128)
        x, dy = rnd((128, hidden), 1), rnd((128, hidden), 2)
        w1, w2 = rnd((ffn, hidden), 3), rnd((ffn, hidden), 4)
        hv = rnd((ffn // 128, 128, 128), 5)

        def one_step():
            oracle.sdd(x, False, w1, True, topo)
            oracle.dsd(hv, topo, False, w2, False)
            oracle.sdd(dy, False, w2, True, topo)
            oracle.dsd(hv, topo, True, dy, False)
            oracle.dsd(hv, topo, False, w1, False)
            oracle.dsd(hv, topo, True, x, False)
        flops = 6 * 2.0 * 128 * ffn * hidden
        sample = ("oracle port (oracle/_ref host library missing): six products on 1 block row of "
                  "one expert, all host cores")
    dt = _timed_steps(args, one_step)
    return reference_line(args, config, "bf16", flops / dt / 1e12, dt, cores, kind, sample, "strong")


def reference_product_arm(args, wl):
    """--impl reference --workload <one product>: densify the first block row(s) of the sparse
    operand, T() where a transpose flag is set, operator*, Mul(mask) for SDD — every op through
    oracle/_ref; each host thread on its own sample (same shapes, own values)."""
    import numpy as np
    synth, _ = _import_host_only()
    from oracle import oracle
    cores = os.cpu_count() or 1
    config = product_config(wl.describe(), wl.note)
    budget = min(2.0, 100.0 / max(1, args.steps + args.warmup))
    rnd = lambda shape, s: oracle.round_to(synth.uniform_values(shape, 4000 + s), wl.dtype)
    if not RefHost.available():
        fn, flops, desc = cpu_sample(wl, pick_cpu_rows(wl, 4.0))
        dt = _timed_steps(args, fn)
        return reference_line(args, config, wl.dtype, flops / dt / 1e12, dt, oracle.num_threads(),
                              "port", desc + "; oracle port (oracle/_ref host library missing)", "weak")
    ref = RefHost()
    sub = _sub_topology(wl.topo, 1)
    nbs = max(1, sub.nnz_blocks)
    sp_rows, sp_cols = sub.rows, sub.cols

    def make(ext, seed):
        """ext = sampled extent of the free DENSE dimension (n for DSD, m for DDS, k for SDD)."""
        vals = rnd((sub.nnz_blocks, 128, 128), seed)
        dense_sp = oracle.densify(vals, sub.offsets, sub.indices, sub.block_rows, sub.block_cols)
        if wl.op == "dsd":       # op(S)[m, k] x op(B)[k, ext]
            kdim = sp_rows if wl.ta else sp_cols
            a, b, mask = ref.new(dense_sp), ref.new(rnd((ext, kdim) if wl.tb else (kdim, ext), seed + 1)), None
        elif wl.op == "dds":     # op(A)[ext, k] x op(S)[k, n]
            kdim = sp_cols if wl.tb else sp_rows
            a, b, mask = ref.new(rnd((kdim, ext) if wl.ta else (ext, kdim), seed + 1)), ref.new(dense_sp), None
        else:                    # SDD: op(A)[rows, ext] x op(B)[ext, cols], masked
            a = ref.new(rnd((ext, sp_rows) if wl.ta else (sp_rows, ext), seed + 1))
            b = ref.new(rnd((sp_cols, ext) if wl.tb else (ext, sp_cols), seed + 2))
            ones = np.ones((sub.nnz_blocks, 128, 128), np.float32)
            mask = ref.new(oracle.densify(ones, sub.offsets, sub.indices, sub.block_rows, sub.block_cols))
        return lambda: ref.product(a, wl.ta, b, wl.tb, mask)
    full = {"dsd": wl.n, "dds": wl.m, "sdd": wl.k}[wl.op]
    probe = [make(8, 10 * i) for i in range(cores)]
    _run_threads(probe)
    t0 = time.perf_counter()
    _run_threads(probe)
    per_col = (time.perf_counter() - t0) / 8
    ext = int(max(8, min(full, (budget / per_col) // 8 * 8)))
    runs = [make(ext, 10 * i) for i in range(cores)]
    dt = _timed_steps(args, lambda: _run_threads(runs))
    flops = cores * 2.0 * nbs * 128 * 128 * ext
    sample = (f"{cores} host threads, each: block row 0 of the sparse matrix densified "
              f"({sub.nnz_blocks} nonzero blocks, {sp_cols} columns), {ext} of {full} along the free "
              f"dense dimension, through the reference's sputnik::Matrix T() / operator* / Mul() "
              f"(oracle/_ref, as written: dense, single-threaded per call; only nonzero-block flops "
              f"are counted); operands converted outside the timed loop")
    return reference_line(args, config, wl.dtype, flops / dt / 1e12, dt, cores, "reference", sample, "weak")


# =========================================================================== GPU arm helpers
def _events(torch):
    return torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)


IDLE_BEFORE_RECORD_S = 1.0


def time_calls(torch, fn, warmup=10, iters=100, graph=True, replays=3, idle_s=0.0):
    """ms per call: `warmup` calls, then `iters` timed ones (the reference's benchmark method,
    dsd_benchmark.cu:82-107).  graph=True: the `iters` calls are captured into one CUDA graph and
    `replays` replays are timed, which takes the host (Python + ctypes, ~10 us per call) out of
    kernels shorter than that.  idle_s: the GPU is left idle that long first, so that a 5-30 ms
    record of a single product starts from the state the burst peak was measured in ("a kernel
    timed alone") and not from the power-capped state the previous record left behind."""
    if idle_s > 0:
        torch.cuda.synchronize()
        time.sleep(idle_s)
    for i in range(warmup):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = _events(torch)
    if graph:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=side):
                for i in range(iters):
                    fn(i)
            g.replay()
            torch.cuda.synchronize()
            e0.record()
            for _ in range(replays):
                g.replay()
            e1.record()
        torch.cuda.synchronize()
        del g
        return e0.elapsed_time(e1) / (replays * iters)
    e0.record()
    for i in range(iters):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def record_of(wl, ms, peaks, **extra):
    tf = wl.flops / ms / 1e9
    rec = {"desc": wl.describe(), "us": round(ms * 1e3, 2), "tflops": round(tf, 1),
           "frac": round(tf / peaks["bf16_burst"], 4),
           "achieved_hbm_gbs": round(wl.algorithmic_bytes / ms / 1e6, 1),
           "hbm_frac": round(wl.algorithmic_bytes / ms / 1e6 / peaks["hbm_gbs"], 4)}
    rec.update(extra)
    return rec


def config_records(torch, peaks, names, dev):
    """One record per named single-product workload: device-resident, operand sets rotating so no
    call finds its inputs in L2, CUDA-graph replay of 100 calls after 10 warm-up calls."""
    from sputnik_b200 import harness, workloads
    out = {}
    for name in names:
        wl = workloads.get(name)
        nsets = max(2, min(4, -(-2 * L2_BYTES // wl.algorithmic_bytes)))
        prob = harness.DeviceProblem(wl, dev, sets=nsets, seed=7)
        ms = time_calls(torch, lambda i: prob.step(i), idle_s=IDLE_BEFORE_RECORD_S)
        extra = {}
        if name in ("c3a", "c3b"):
            # the reference's Matmul (not MatmulEx) rebuilds the transposed metadata on every
            # call (dsd.cu:9-18 -> transpose.cu:69-125); here that is one more device kernel
            ms_meta = time_calls(torch, lambda i: prob.step(i, with_metadata=True),
                                 idle_s=IDLE_BEFORE_RECORD_S)
            extra = {"us_with_metadata": round(ms_meta * 1e3, 2),
                     "tflops_with_metadata": round(wl.flops / ms_meta / 1e9, 1)}
        if name in ("c2", "c3a", "c3b"):
            # an unmodified caller of the reference API passes no topology hints
            for s in prob.sets:
                s.sparse.twin_rows = s.sparse.twin_cols = s.sparse.even_rows = False
            ms_nh = time_calls(torch, lambda i: prob.step(i), idle_s=IDLE_BEFORE_RECORD_S)
            extra["us_no_hints"] = round(ms_nh * 1e3, 2)
        out[name] = record_of(wl, ms, peaks, **extra)
        del prob
        torch.cuda.empty_cache()
    return out


def swiglu_record(torch, peaks, dev, experts=8, hidden=4096, ffn=14336, rows_per_expert=1024):
    """Fused SwiGLU SDD (BSMM_FLAG_SWIGLU: gate | up panels interleaved by block column, weights
    stored [2 E F, hidden]) on a dMoE shape, beside the unfused sequence it replaces: two SDD
    products (2-SM kernel) + an elementwise silu(g) * u.  flops = both products."""
    import torch.nn.functional as F
    import sputnik_b200 as sb
    from sputnik_b200 import synth
    topo = synth.moe_topology(experts, rows_per_expert // 128, ffn // 128)
    g = torch.Generator(device=dev).manual_seed(0)
    rnd = lambda *s: torch.empty(s, device=dev).uniform_(-1, 1, generator=g).bfloat16()
    x = rnd(topo.rows, hidden)
    w_gu = rnd(2 * topo.cols, hidden) * 0.02
    w4 = w_gu.view(topo.cols // 128, 2, 128, hidden)
    w_g, w_u = (w4[:, i].reshape(topo.cols, hidden).contiguous() for i in (0, 1))

    def blocks():
        m = sb.BlockMatrix(topo.rows, topo.cols, 128, topo.nonzeros,
                           torch.empty((topo.nnz_blocks, 128, 128), dtype=torch.bfloat16, device=dev),
                           torch.from_numpy(topo.offsets).to(dev), torch.from_numpy(topo.indices).to(dev))
        sb.allocate_row_indices_buffer(m)
        sb.row_indices(m)
        m.__dict__.update(topo.hints())
        return m
    h, hg, hu = blocks(), blocks(), blocks()
    out = torch.empty_like(h.data)
    M = sb.Matrix.of

    def fused(i):
        sb.matmul_ex(M(x), False, M(w_gu), True, h, activation="swiglu")

    def unfused(i):
        sb.matmul_ex(M(x), False, M(w_g), True, hg)
        sb.matmul_ex(M(x), False, M(w_u), True, hu)
        torch.mul(F.silu(hg.data), hu.data, out=out)
    ms_f = time_calls(torch, fused, 5, 20, graph=False, idle_s=IDLE_BEFORE_RECORD_S)
    ms_u = time_calls(torch, unfused, 5, 20, graph=False, idle_s=IDLE_BEFORE_RECORD_S)
    flops = 2 * 2.0 * topo.nnz_blocks * 128 * 128 * hidden
    tf = flops / ms_f / 1e9
    return {"desc": f"h = silu(x wg^T) * (x wu^T) sampled at the dMoE topology: {experts} experts, "
                    f"{rows_per_expert} rows each, hidden {hidden}, ffn {ffn}, bf16, one SDD launch",
            "us": round(ms_f * 1e3, 1), "tflops": round(tf, 1), "frac": round(tf / peaks["bf16_burst"], 4),
            "us_unfused": round(ms_u * 1e3, 1), "fused_over_unfused": round(ms_f / ms_u, 3),
            "note": "unfused = two SDD products (2-SM kernel) + torch silu * mul on the block values; "
                    "eager calls, 5 warm-up + 20 timed"}


def legacy_gpu_records(torch, names, dev):
    """The reference's own kernels (CUTLASS 2.5 mma.sync, fp16; oracle/_ref/libsputnik_ref.so,
    compiled for sm_100 from the reference's sources) with the reference's benchmark method:
    10 warm-up + 100 timed back-to-back calls (dsd_benchmark.cu:82-107).  Baseline only."""
    path = os.path.join(ROOT, "oracle", "_ref", "libsputnik_ref.so")
    if not os.path.exists(path):
        return {"unavailable": "oracle/_ref/libsputnik_ref.so not built"}
    from sputnik_b200 import harness, workloads
    lib = ctypes.CDLL(path)
    p = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else None
    cur = lambda: ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    out = {"method": "10 warm-up + 100 timed back-to-back calls, CUDA events (dsd_benchmark.cu:82-107)",
           "dtype": "f16", "library": "oracle/_ref/libsputnik_ref.so (reference sources, sm_100)"}
    for name in names:
        wl = workloads.get(name)
        wl.dtype = "f16"
        prob = harness.DeviceProblem(wl, dev, sets=1, seed=7)
        s, t, sp = prob.sets[0], wl.topo, prob.sets[0].sparse
        if wl.op == "sdd":
            call = lambda i: lib.ref_sdd(s.a.rows, s.a.cols, p(s.a.data), int(wl.ta), s.b.rows,
                                         s.b.cols, p(s.b.data), int(wl.tb), t.rows, t.cols,
                                         t.nonzeros, p(sp.data), p(sp.offsets), p(sp.indices),
                                         p(sp.row_indices), cur())
        elif wl.op == "dsd":
            call = lambda i: lib.ref_dsd(t.rows, t.cols, t.nonzeros, p(sp.data), p(sp.offsets),
                                         p(sp.indices), p(sp.offsets_t), p(sp.indices_t),
                                         p(sp.block_offsets), int(wl.ta), s.b.rows, s.b.cols,
                                         p(s.b.data), int(wl.tb), wl.m, wl.n, p(s.c.data), 0, cur())
        else:
            call = lambda i: lib.ref_dds(s.a.rows, s.a.cols, p(s.a.data), int(wl.ta), t.rows,
                                         t.cols, t.nonzeros, p(sp.data), p(sp.offsets),
                                         p(sp.indices), p(sp.offsets_t), p(sp.indices_t),
                                         p(sp.block_offsets), int(wl.tb), wl.m, wl.n, p(s.c.data),
                                         0, cur())
        ms = time_calls(torch, call, graph=False, idle_s=IDLE_BEFORE_RECORD_S)
        out[name] = {"us": round(ms * 1e3, 2), "tflops": round(wl.flops / ms / 1e9, 1)}
        del prob
        torch.cuda.empty_cache()
    return out


def sharded_records(torch, dist, peaks, names, rank, world, dev):
    """N > 1: ONE unstructured product split into nnz-balanced contiguous block-row slabs
    (sputnik_b200/shard.py), dense operand replicated, outputs disjoint, no collective.  Time =
    max over ranks of the rank's slab product; flops = the whole product's."""
    from sputnik_b200 import harness, shard, workloads
    out = {}
    for name in names:
        parent = workloads.get(name)
        slab = shard.row_slabs(parent.topo, world)[rank]
        wl = workloads.Workload(f"{name}[slab {rank}/{world}]", parent.op, parent.ta, parent.tb,
                                parent.dtype, slab.topo.rows if parent.op != "dds" else parent.m,
                                parent.k, parent.n, slab.topo, parent.note)
        dist.barrier()
        if slab.topo.nnz_blocks == 0:
            ms = 0.0
        else:
            prob = harness.DeviceProblem(wl, dev, sets=3, seed=7)
            ms = time_calls(torch, lambda i: prob.step(i))
            del prob
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t[0])
        tf = parent.flops / ms / 1e9
        out[name + "_sharded"] = {
            "desc": parent.describe() + f", nnz-balanced block-row slabs over {world} ranks",
            "us": round(ms * 1e3, 2), "tflops": round(tf, 1),
            "frac": round(tf / (world * peaks["bf16_burst"]), 4), "scaling": "strong"}
        torch.cuda.empty_cache()
    return out


def ragged_record(torch, dist, args, peaks, rank, world, dev, experts, tokens, hidden, ffn):
    """The layer with a Zipf-like (ragged) routing whose counts exist only on the device:
    DeviceRoutedFFN (bsmm_moe_exchange + bsmm_moe_topology per step, fused peer-memory exchange,
    no host synchronisation).  Every rank holds tokens / world rows; expert popularity ~ 1 / (e + 1)^0.8
    (seeded); flops count the nonzero blocks of the padded topology the counts imply."""
    import numpy as np
    from sputnik_b200 import moe
    rng = np.random.default_rng(1234)
    pop = 1.0 / np.arange(1, experts + 1) ** 0.8
    pop = rng.permutation(pop / pop.sum())
    rows = tokens // world
    counts = np.stack([np.random.default_rng(100 + s).multinomial(rows, pop) for s in range(world)])
    le = experts // world
    per_owner = counts.reshape(world, world, le).sum(axis=0)                     # [owner, e_local]
    blocks = [[-(-int(v) // 256) * 2 for v in per_owner[g]] for g in range(world)]
    cap = max(sum(b) for b in blocks) + 2
    nb_total = sum(sum(b) for b in blocks) * (ffn // 128)
    flops = 6 * 2.0 * nb_total * 128 * 128 * hidden
    layer = moe.DeviceRoutedFFN(world, rank, experts, hidden, ffn, rows, cap, torch.bfloat16, dev, seed=rank)
    g = torch.Generator(device=dev).manual_seed(50 + rank)
    layer.x_in.uniform_(-1, 1, generator=g)
    layer.dy_in.uniform_(-1, 1, generator=g)
    my_counts = torch.from_numpy(counts[rank].astype(np.int32)).to(dev)

    def step():
        layer.route(my_counts)
        layer.forward(rows)
        layer.backward(rows)
    for _ in range(args.warmup):
        step()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = _events(torch)
    torch.cuda.set_sync_debug_mode("error")      # the timed steps must not synchronise with the host
    try:
        e0.record()
        for _ in range(args.steps):
            step()
        e1.record()
    finally:
        torch.cuda.set_sync_debug_mode("default")
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    if world > 1:
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t[0])
    del layer
    torch.cuda.empty_cache()
    tf = flops / ms / 1e9
    worst = max(sum(b) for b in blocks) * world / max(1, sum(sum(b) for b in blocks))
    return {"desc": f"same layer, Zipf-like routing (popularity ~ 1/(e+1)^0.8, seeded): per-expert rows "
                    f"{int(counts.sum(axis=0).min())}..{int(counts.sum(axis=0).max())}; routing counts "
                    f"device-resident, topology + exchange tables rebuilt on the device every step, "
                    f"no host synchronisation (torch sync-debug mode 'error' around the timed steps)",
            "ms_per_step": round(ms, 3), "tflops": round(tf, 1),
            "frac_of_sustained": round(tf / world / peaks["bf16_sustained"], 4),
            "rank_load_imbalance": round(worst, 3), "scaling": "strong"}


# =========================================================================== GPU arm: one product
def ours_arm(args, wl, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    import sputnik_b200 as sb
    from sputnik_b200 import harness
    lib = sb._lib.lib

    torch.cuda.set_device(local_rank)
    dev = f"cuda:{local_rank}"
    peaks = load_peaks()

    # Operand sets rotate so that back-to-back steps never find their inputs in L2.
    set_bytes = wl.algorithmic_bytes
    nsets = max(2, min(8, -(-3 * L2_BYTES // set_bytes)))
    prob = harness.DeviceProblem(wl, dev, sets=nsets, seed=rank)
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident leg: `value`
    for i in range(args.warmup):
        prob.step(i)
    barrier()
    sampler = ClockSampler(physical_gpu_index(local_rank))
    sampler.start()
    ev0, ev1 = _events(torch)
    launches0 = lib.bsmm_launch_count()
    if args.graph:
        G = next(g for g in range(min(args.steps, 200), 0, -1) if args.steps % g == 0)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=side):
                for i in range(G):
                    prob.step(i)
            per_graph = lib.bsmm_launch_count() - launches0
            graph.replay()
            barrier()
            ev0.record()
            for _ in range(args.steps // G):
                graph.replay()
            ev1.record()
        barrier()
        launches = per_graph * (args.steps // G)
    else:
        ev0.record()
        for i in range(args.steps):
            prob.step(i)
        ev1.record()
        barrier()
        launches = lib.bsmm_launch_count() - launches0
    clocks = sampler.stop()
    ms = ev0.elapsed_time(ev1) / args.steps

    # ---- end-to-end leg: host (pinned) inputs -> H2D -> operator -> D2H of the result, per step
    host, result = harness.host_operands(wl, seed=rank)
    s = prob.sets[0]
    dst = {"values": s.sparse.data if wl.op != "sdd" else None,
           "a": s.a.data if s.a else None, "b": s.b.data if s.b else None}
    h2d = sum(t.numel() * t.element_size() for t in host.values())
    d2h = result.numel() * result.element_size()
    e2e_steps = max(3, min(args.steps, 20))
    main_stream = torch.cuda.current_stream()
    out_stream = torch.cuda.Stream()
    computed, drained = torch.cuda.Event(), torch.cuda.Event()
    drained.record(main_stream)

    def e2e_step():
        for key, t in host.items():
            dst[key].copy_(t, non_blocking=True)
        main_stream.wait_event(drained)       # the output buffer is free again
        prob.call(s)
        computed.record(main_stream)
        with torch.cuda.stream(out_stream):
            out_stream.wait_event(computed)
            result.copy_(s.output(), non_blocking=True)
            drained.record(out_stream)
    for _ in range(3):
        e2e_step()
    barrier()
    ee0, ee1 = _events(torch)
    ee0.record()
    for _ in range(e2e_steps):
        e2e_step()
    main_stream.wait_stream(out_stream)       # the last result has landed on the host
    ee1.record()
    barrier()
    e2e_ms = ee0.elapsed_time(ee1) / e2e_steps

    if world > 1:
        t = torch.tensor([ms, e2e_ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, e2e_ms = float(t[0]), float(t[1])
    if rank != 0:
        return None

    hints = wl.topo.hints()
    twin = hints["twin_cols"] if ((wl.op == "dsd" and wl.ta) or (wl.op == "dds" and not wl.tb)) \
        else hints["twin_rows"]
    kernel_name = "bsmm_pair_kernel (2-SM, tcgen05.mma.cta_group::2)" if twin and \
        (wl.op != "sdd" or hints["even_rows"]) else "bsmm_gemm_kernel (1-SM)"
    total_flops = args.parent.flops if args.parent is not None else wl.flops * world
    value = total_flops / (ms * 1e-3) / 1e12
    per_gpu = wl.flops / (ms * 1e-3) / 1e12
    hbm = wl.algorithmic_bytes / (ms * 1e-3) / 1e9
    traffic, traffic_src = ncu_traffic(wl.name)
    config = product_config((args.parent or wl).describe(), wl.note)
    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "strong" if args.parent is not None else "weak",
        "vs_baseline": None, "dtype": wl.dtype, "data": "synthetic", "config": config,
        "notes": {"timing": "CUDA-graph replay of the steps" if args.graph else "one launch per step from the host",
                  "operand_sets": nsets,
                  "sharding": ("one problem, nnz-balanced contiguous block-row slabs, dense operand replicated, "
                               "outputs disjoint; no collective") if args.parent is not None
                              else "each rank runs the workload on its own batch; no collective"},
        "roofline": {"bound": "tensor", "achieved": per_gpu, "peak": peaks["bf16_burst"],
                     "unit": "TFLOP/s", "frac": per_gpu / peaks["bf16_burst"],
                     "frac_of_sustained": per_gpu / peaks["bf16_sustained"],
                     "peak_source": peaks["source"] + " (burst; kernel timed back-to-back alone)",
                     "achieved_hbm_gbs": hbm, "hbm_frac": hbm / peaks["hbm_gbs"],
                     "traffic": traffic, "traffic_source": traffic_src,
                     "kernel": f"{kernel_name}<{wl.op}>: 1 launch per step, "
                               f"{wl.flops / 1e9:.2f} GFLOP, {wl.algorithmic_bytes / 1e6:.1f} MB algorithmic"},
        "e2e": {"value": total_flops / (e2e_ms * 1e-3) / 1e12, "unit": UNIT,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms,
                "steps": e2e_steps},
        "gpu_launches": int(launches),
        "clocks": clocks,
    }
    if world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = time_cpu_port(wl)
    return out


# =========================================================================== GPU arm: the layer
PRODUCTS = [  # name, op kind, the workload whose flops / algorithmic bytes describe one launch
    ("h=sdd(x,w1^T)", "sdd", "sdd_nt"), ("y=dsd(h,w2)", "dsd", "dsd"),
    ("dh=sdd(dy,w2^T)", "sdd", "sdd_nt"), ("dw2=dsd(h^T,dy)", "dsd", "dsd_tn"),
    ("dx=dsd(dh,w1)", "dsd", "dsd"), ("dw1=dsd(dh^T,x)", "dsd", "dsd_tn"),
    # w1 stored [hidden, experts*ffn] (alt_layout)
    ("h=sdd(x,w1)", "sdd", "sdd"), ("dx=dsd(dh,w1^T)", "dsd", "dsd_nt"),
    ("dw1=dds(x^T,dh)", "dds", "dds_tn"),
]


def alt_layout_record(torch, args, plan, dev, peaks, experts, tokens, hidden, ffn):
    """The same layer with w1 stored [hidden, experts*ffn] (x @ w1; weight gradient = the DDS
    product): N = 1, fused exchange path, same steps / warm-up, per-product event timers."""
    from sputnik_b200 import moe, workloads
    layer = moe.ExpertShardedFFN(plan, torch.bfloat16, dev, seed=0, fused_combine=True, w1_layout="hf")
    g = torch.Generator(device=dev).manual_seed(0)
    mk = lambda: torch.empty((plan.rows_in, hidden), device=dev).uniform_(-1, 1, generator=g).bfloat16()
    layer.x_in[:plan.rows_in].copy_(mk())
    layer.dy_in[:plan.rows_in].copy_(mk())
    x_f, dy_f = layer.x_in[:plan.rows_in], layer.dy_in[:plan.rows_in]

    def step():
        layer.forward(x_f, fused=True)
        layer.backward(dy_f, fused=True)
    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    layer.timers = {}
    ev0, ev1 = _events(torch)
    ev0.record()
    for _ in range(args.steps):
        step()
    ev1.record()
    torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1) / args.steps
    kinds = {p[0]: p[2] for p in PRODUCTS}
    prod = {}
    for name, evs in layer.timers.items():
        pms = sum(a.elapsed_time(b) for a, b in evs) / len(evs)
        wl = workloads.get(f"c5_{kinds[name]}_e{experts}")
        prod[name] = {"us": round(pms * 1e3, 1), "tflops": round(wl.flops / pms / 1e9, 1),
                      "frac_of_sustained": round(wl.flops / pms / 1e9 / peaks["bf16_sustained"], 4)}
    layer.timers = None
    del layer, x_f, dy_f
    torch.cuda.empty_cache()
    return {"w1_layout": "hf: w1 stored [hidden, experts*ffn]; h=sdd(x,w1), dx=dsd(dh,w1^T), dw1=dds(x^T,dh)",
            "value": layer_flops(experts, tokens, hidden, ffn) / (ms * 1e-3) / 1e12, "unit": UNIT,
            "ms_per_step": ms, "steps": args.steps, "per_product": prod}


def layer_arm(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    import sputnik_b200 as sb
    from sputnik_b200 import moe
    lib = sb._lib.lib
    torch.cuda.set_device(local_rank)
    dev = torch.device(f"cuda:{local_rank}")
    peaks = load_peaks()
    experts, tokens, hidden, ffn = parse_layer(args.workload)
    numa = pin_to_gpu_numa(local_rank) if world > 1 else "single rank: not bound"

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- parity gate (before anything is timed): ragged routing at small shapes
    parity = {"checked": False}
    if not args.no_parity:
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import moe_check
        parity = moe_check.check(rank, world, dev)
        if not parity["ok"]:
            if rank == 0:
                print(json.dumps({"metric": METRIC, "error": "parity check failed", "parity": parity,
                                  "n_gpus": world}), flush=True)
            sys.exit(1)

    plan = moe.ExpertShardPlan.uniform(world, rank, experts, tokens, hidden, ffn)
    layer = moe.ExpertShardedFFN(plan, torch.bfloat16, dev, seed=rank, fused_combine=True)
    g = torch.Generator(device=dev).manual_seed(rank)
    mk = lambda: torch.empty((plan.rows_in, hidden), device=dev).uniform_(-1, 1, generator=g).bfloat16()
    # fused path: the caller's rows live in the layer's symmetric buffers (no staging copy)
    layer.x_in[:plan.rows_in].copy_(mk())
    layer.dy_in[:plan.rows_in].copy_(mk())
    x_f, dy_f = layer.x_in[:plan.rows_in], layer.dy_in[:plan.rows_in]
    x_n, dy_n = x_f.clone(), dy_f.clone()
    y, dx = torch.empty_like(x_n), torch.empty_like(x_n)

    def step_fused():
        layer.forward(x_f, fused=True)
        layer.backward(dy_f, fused=True)

    def step_nccl():
        layer.forward(x_n, out=y, fused=False)
        layer.backward(dy_n, out=dx, fused=False)

    def timed(step, with_timers):
        for _ in range(args.warmup):
            step()
        barrier()
        sampler = ClockSampler(physical_gpu_index(local_rank))
        sampler.start()
        layer.timers = {} if with_timers else None
        ev0, ev1 = _events(torch)
        launches0 = lib.bsmm_launch_count()
        ev0.record()
        for _ in range(args.steps):
            step()
        ev1.record()
        barrier()
        launches = lib.bsmm_launch_count() - launches0
        clocks = sampler.stop()
        timers, layer.timers = layer.timers, None
        return ev0.elapsed_time(ev1) / args.steps, launches, clocks, timers

    # Three timed regions of K steps each, same warm-up and bracketing: (1) `value`: the fused
    # path with nothing between the launches; (2) `alt`: the NCCL path; (3) the fused path again
    # with a CUDA event pair around every product, which is where roofline.per_product comes
    # from.  (3) is kept apart from (1) because the twelve event records per step cost the step
    # 3-6 % on this pod (tools/layer_order_probe.py, profiles/r02_layer_order_probe.txt); its own
    # step time is reported as roofline.ms_per_step_with_events.
    ms, launches, clocks, _ = timed(step_fused, False)
    ms_nccl, _, _, _ = timed(step_nccl, False)
    ms_events, _, _, timers = timed(step_fused, True)

    # per-product device times from events recorded inside timed region (3)
    per_product = {}
    for name, evs in (timers or {}).items():
        per_product[name] = sum(a.elapsed_time(b) for a, b in evs) / len(evs)

    # ---- end to end: token rows in from pinned host memory, y / dx out to pinned host memory,
    # every step; copies run on their own streams and overlap the other pass's compute.
    hx, hdy = x_n.cpu().pin_memory(), dy_n.cpu().pin_memory()
    hy, hdx = torch.empty_like(hx).pin_memory(), torch.empty_like(hx).pin_memory()
    row_bytes = hx.numel() * hx.element_size()
    e2e_steps = max(3, min(args.steps, 10))
    comp = torch.cuda.current_stream()
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
    E = lambda: torch.cuda.Event()
    ev = {k: E() for k in ("x_rdy", "dy_rdy", "fwd_done", "bwd_done", "y_out", "dx_out")}
    for k in ("fwd_done", "bwd_done", "y_out", "dx_out"):
        ev[k].record(comp)

    def e2e_step():
        with torch.cuda.stream(s_in):
            s_in.wait_event(ev["fwd_done"])          # peers have gathered the previous x rows
            x_f.copy_(hx, non_blocking=True)
            ev["x_rdy"].record(s_in)
            s_in.wait_event(ev["bwd_done"])
            dy_f.copy_(hdy, non_blocking=True)
            ev["dy_rdy"].record(s_in)
        comp.wait_event(ev["x_rdy"])
        comp.wait_event(ev["y_out"])                 # the previous y has left out_y
        yv = layer.forward(x_f, fused=True)
        ev["fwd_done"].record(comp)
        with torch.cuda.stream(s_out):
            s_out.wait_event(ev["fwd_done"])
            hy.copy_(yv, non_blocking=True)
            ev["y_out"].record(s_out)
        comp.wait_event(ev["dy_rdy"])
        comp.wait_event(ev["dx_out"])
        dxv = layer.backward(dy_f, fused=True)
        ev["bwd_done"].record(comp)
        with torch.cuda.stream(s_out):
            s_out.wait_event(ev["bwd_done"])
            hdx.copy_(dxv, non_blocking=True)
            ev["dx_out"].record(s_out)
    for _ in range(2):
        e2e_step()
    comp.wait_stream(s_out)
    barrier()
    ee0, ee1 = _events(torch)
    ee0.record()
    for _ in range(e2e_steps):
        e2e_step()
    comp.wait_stream(s_out)                          # the last results have landed on the host
    ee1.record()
    barrier()
    e2e_ms = ee0.elapsed_time(ee1) / e2e_steps

    if world > 1:
        t = torch.tensor([ms, e2e_ms, ms_nccl, ms_events], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, e2e_ms, ms_nccl, ms_events = float(t[0]), float(t[1]), float(t[2]), float(t[3])

    # free the layer before the sub-records
    del layer, x_f, dy_f, x_n, dy_n, y, dx
    torch.cuda.empty_cache()

    extra = {}
    if world == 1 and not args.no_configs:
        extra["alt_layout"] = alt_layout_record(torch, args, plan, dev, peaks, experts, tokens, hidden, ffn)
    if not args.no_configs:
        extra["ragged"] = ragged_record(torch, dist, args, peaks, rank, world, dev, experts, tokens,
                                        hidden, ffn)
    if world == 1 and not args.no_configs:
        names = ["c1", "c2", "c3a", "c3b"] + [f"c4_{op}_{d}" for op in ("sdd", "dsd", "dds")
                                              for d in (10, 25, 50, 75, 90)]
        extra["configs"] = config_records(torch, peaks, names, str(dev))
        extra["configs"]["_method"] = ("device-resident, rotating operand sets (> L2 in total), 1 s of idle GPU, "
                                       "10 warm-up calls then CUDA-graph replays of 100 calls; frac = tflops "
                                       "/ bf16 burst peak (a kernel timed alone); flops = 2 * nnz_blocks * "
                                       "128^2 * D")
        # FP8 (e4m3 operands, bf16 output): time-to-solution beside bf16
        f8 = config_records(torch, peaks, ["c2@f8", "c5_dsd_e8", "c5_dsd_e8@f8", "c5_sdd_e8",
                                           "c5_sdd_e8@f8"], str(dev))
        for name in ("c2", "c5_dsd_e8", "c5_sdd_e8"):
            base = f8.get(name) or extra["configs"].get(name)
            f8[name + "@f8"]["speedup_vs_bf16"] = round(base["us"] / f8[name + "@f8"]["us"], 3)
            f8[name + "@f8"]["note"] = ("e4m3 operands, fp32 accumulate (tcgen05.mma.kind::f8f6f4), bf16 output; "
                                        "frac is still quoted against the BF16 peak")
        extra["configs"].update(f8)
        extra["configs"]["c5_sdd_e8+swiglu"] = swiglu_record(torch, peaks, str(dev))
        if not args.no_legacy:
            extra["legacy_gpu"] = legacy_gpu_records(
                torch, ["c1", "c2", "c3a", "c3b", "c4_sdd_50", "c4_dsd_50", "c4_dds_50"], str(dev))
    elif world > 1 and not args.no_configs:
        extra["configs"] = sharded_records(torch, dist, peaks, ["c4_dsd_50", "c4_sdd_50", "c4_dsd_90"],
                                           rank, world, str(dev))
    if rank != 0:
        return None

    from sputnik_b200 import workloads
    flops = layer_flops(experts, tokens, hidden, ffn)
    value = flops / (ms * 1e-3) / 1e12
    per_gpu = value / world
    a2a = 4 * row_bytes * (world - 1) // world
    # the dominant kernel of the step: the product with the largest share of the timed region
    prod = {}
    for (pname, kind, wname) in PRODUCTS:
        if pname not in per_product:
            continue
        wl = workloads.get(f"c5_{wname}_e{experts // world}")  # one rank's share of the product
        pms = per_product[pname]
        prod[pname] = {"us": round(pms * 1e3, 1), "tflops": round(wl.flops / pms / 1e9, 1),
                       "frac_of_sustained": round(wl.flops / pms / 1e9 / peaks["bf16_sustained"], 4),
                       "share_of_step": round(pms / ms_events, 4),
                       "algorithmic_mb": round(wl.algorithmic_bytes / 1e6, 1)}
    dom = max(prod, key=lambda k: prod[k]["us"]) if prod else None
    dom_kind = {p[0]: p[2] for p in PRODUCTS}.get(dom)
    traffic, traffic_src = ncu_traffic(f"c5_{dom_kind}_e{experts // world}") if dom else (None, None)
    roof_tf = prod[dom]["tflops"] if dom else per_gpu
    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": layer_config(args.workload, world),
        "notes": {"exchange": "fused: one bsmm_gather_rows pass pulls the token rows from the source ranks' "
                              "symmetric-memory buffers over NVLink, the y / dx DSD epilogues store result rows "
                              "straight into them, stream barriers in between; no collective call on the data path"
                              if world > 1 else "single rank, aligned uniform routing: the exchange is the identity and the layer "
                                               "skips it (the chain reads the caller's rows in place and writes y / dx "
                                               "with the plain TMA-store epilogue)",
                  "exchange_bytes_per_rank_per_step": a2a, "host_affinity": numa},
        "alt": {"exchange": "NCCL all_to_all_single before and after the local chain, forward and backward"
                            if world > 1 else "single rank: the general path (dispatch / combine copies around the chain)",
                "value": flops / (ms_nccl * 1e-3) / 1e12, "unit": UNIT, "ms_per_step": ms_nccl},
        "parity": parity,
        "roofline": {"bound": "tensor", "achieved": roof_tf, "peak": peaks["bf16_sustained"],
                     "unit": "TFLOP/s", "frac": roof_tf / peaks["bf16_sustained"],
                     "frac_of_burst": roof_tf / peaks["bf16_burst"],
                     "step_frac": per_gpu / peaks["bf16_sustained"],
                     "peak_source": peaks["source"] + " (sustained: kernels timed inside a long step)",
                     "traffic": traffic, "traffic_source": traffic_src,
                     "kernel": (f"{dom}: 2-SM tcgen05 kernel, 1 launch per step per rank; achieved = its "
                                f"nonzero-block flops per launch / its mean duration between CUDA events "
                                f"recorded around it inside a timed region of the same K steps") if dom else None,
                     "ms_per_step_with_events": ms_events,
                     "per_product": prod},
        "e2e": {"value": flops / (e2e_ms * 1e-3) / 1e12, "unit": UNIT,
                "h2d_bytes_per_step": 2 * row_bytes, "d2h_bytes_per_step": 2 * row_bytes,
                "ms_per_step": e2e_ms, "steps": e2e_steps,
                "what": "x and dy token rows from pinned host memory, y and dx back to pinned host memory, "
                        "every step; copies on their own streams overlap the other pass's compute"},
        "gpu_launches": int(launches),
        "clocks": clocks,
    }
    out.update(extra)
    if world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = time_cpu_port_layer(experts, tokens, hidden, ffn)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=None)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c5_layer")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the per-config sub-records")
    ap.add_argument("--no-legacy", action="store_true", help="skip the legacy-GPU (reference kernels) leg")
    ap.add_argument("--no-parity", action="store_true", help="skip the in-run parity gate")
    ap.add_argument("--shard", action="store_true",
                    help="single product, N > 1: split ONE problem into nnz-balanced block-row slabs")
    ap.add_argument("--graph", action="store_true",
                    help="single product: time CUDA-graph replays of the steps")
    args = ap.parse_args()
    is_layer = args.workload.startswith("c5_layer")
    if args.steps is None:
        args.steps = 20 if is_layer else (2000 if args.impl == "ours" else 5)
    if args.warmup is None:
        args.warmup = 3 if is_layer else (20 if args.impl == "ours" else 1)
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        if rank != 0:
            return  # rank 0 alone runs and prints the reference arm
        if is_layer:
            out = reference_layer_arm(args, world)
        else:
            _, workloads = _import_host_only()
            out = reference_product_arm(args, workloads.get(args.workload))
        print(json.dumps(out), flush=True)
        return

    from sputnik_b200 import workloads
    wl = None if is_layer else workloads.get(args.workload)
    args.parent = None
    if args.shard and world > 1 and not is_layer:
        # Unstructured topology, strong scaling: rank g owns an nnz-balanced slab of block rows
        # (sputnik_b200/shard.py); dense operand replicated; outputs are disjoint row slabs.
        from sputnik_b200 import shard
        assert wl.op in ("dsd", "sdd") and not wl.ta, "--shard: row-walk products (dsd n*, sdd n*)"
        slab = shard.row_slabs(wl.topo, world)[rank]
        args.parent = wl
        wl = workloads.Workload(f"{wl.name}[slab {rank}/{world}]", wl.op, wl.ta, wl.tb, wl.dtype,
                                slab.topo.rows, wl.k, wl.n, slab.topo, wl.note)

    if world > 1:
        import torch
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local_rank}"))
    out = layer_arm(args, rank, world, local_rank) if is_layer else \
        ours_arm(args, wl, rank, world, local_rank)
    if out is not None:
        print(json.dumps(out), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
