// C-ABI of libbsmm.so (declared in include/bsmm.h): validation, TMA tensor-map construction
// and kernel dispatch.  The host logic mirrors the reference's launchers
// (e.g. sputnik/block/dsd/cutlass/dsd_mixed_b128_128x256x32x3_tn_align8.cu:48-107) and its
// "first compatible kernel" registries (sputnik/block/sdd/cutlass/sdd.cu:55-72), without
// their per-call std::function walk or cudaFuncSetAttribute (block/cutlass/kernel.h:19-38).
#include <algorithm>
#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <mutex>

#include <cuda.h>
#include <cuda_runtime.h>

#include "../../include/bsmm.h"
#include "bsmm_kernel.cuh"
#include "bsmm_pair_kernel.cuh"
#include "bsmm_pair_sk_kernel.cuh"
#include "bsmm_sibling_kernel.cuh"
#include "bsmm_topology.cuh"

namespace bsmm {
namespace {

thread_local int g_last_cuda_error = 0;
thread_local int g_last_cuda_error_line = 0;
std::atomic<unsigned long long> g_launches{0};

inline int cuda_fail(cudaError_t e, int line = 0) {
  g_last_cuda_error = int(e);
  g_last_cuda_error_line = line;
  return BSMM_ERR_CUDA;
}
#define BSMM_CUDA(expr)                                             \
  do {                                                              \
    cudaError_t e_ = (expr);                                        \
    if (e_ != cudaSuccess) return ::bsmm::cuda_fail(e_, __LINE__);  \
  } while (0)

using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                   const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                   const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// libcuda is not linked: the entry point is resolved through the runtime so the library
// still loads (and exports its symbols) on a machine without a driver.
EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

struct DeviceInfo {
  int sms = 0;
};
int current_device_sms(int* sms) {
  static DeviceInfo info[64];
  int dev = 0;
  BSMM_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return cuda_fail(cudaErrorInvalidDevice);
  if (info[dev].sms == 0) {
    int n = 0;
    BSMM_CUDA(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
    info[dev].sms = n;
  }
  *sms = info[dev].sms;
  return BSMM_OK;
}

// Per-device scratch of the library: tile counters of the dynamic scheduler and gate words of the
// device-side kernel selection (32 KiB + 16 KiB per device; the only allocations the library
// ever makes, never on the call path afterwards).  Created by bsmm_init() or at first use:
// cudaMalloc + cudaMemset + cudaDeviceSynchronize, so the zeros are in place before ANY stream
// (including cudaStreamNonBlocking ones, which do not order against the legacy stream) can run a
// kernel that draws from them.  First use inside a stream capture cannot allocate: that launch
// falls back to the static schedule and NOTHING is latched — the next call outside a capture
// initialises normally.
//
// Every launch takes the next counter slot round-robin; a counter returns to 0 by itself with the
// last draw of its launch (atomicInc wrap), so slots are reusable without a memset.  Launches
// recorded into a CUDA graph bake their slot pointer in and may be replayed at any later time, so
// they draw from a pool of their own that eager launches never cycle onto.
constexpr unsigned kTileCounterSlots = 4096;  // per pool (eager / captured)
constexpr unsigned kGateSlots = 4096;
struct DeviceScratch {
  std::atomic<unsigned int*> counters{nullptr};  // [2][kTileCounterSlots]
  std::atomic<unsigned int*> gates{nullptr};     // [2][kGateSlots]
  std::mutex mu;
};
DeviceScratch g_scratch[64];

bool stream_is_capturing(cudaStream_t stream) {
  cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(stream, &st) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return st != cudaStreamCaptureStatusNone;
}

// Allocates and zeroes the scratch of device `dev` (current).  Must not be called while `stream`
// is capturing.  Returns false (and latches nothing) on failure.
bool init_scratch(int dev) {
  DeviceScratch& d = g_scratch[dev];
  std::lock_guard<std::mutex> lock(d.mu);
  if (d.counters.load(std::memory_order_acquire)) return true;
  unsigned int *c = nullptr, *g = nullptr;
  const size_t cbytes = 2 * kTileCounterSlots * sizeof(unsigned int);
  const size_t gbytes = 2 * kGateSlots * sizeof(unsigned int);
  if (cudaMalloc(&c, cbytes) != cudaSuccess || cudaMalloc(&g, gbytes) != cudaSuccess ||
      cudaMemset(c, 0, cbytes) != cudaSuccess || cudaMemset(g, 0, gbytes) != cudaSuccess ||
      cudaDeviceSynchronize() != cudaSuccess) {
    cudaGetLastError();
    if (c) cudaFree(c);
    if (g) cudaFree(g);
    return false;
  }
  d.gates.store(g, std::memory_order_release);
  d.counters.store(c, std::memory_order_release);
  return true;
}

DeviceScratch* scratch_for(cudaStream_t stream, bool* capturing) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  DeviceScratch& d = g_scratch[dev];
  *capturing = stream_is_capturing(stream);
  if (!d.counters.load(std::memory_order_acquire)) {
    if (*capturing || !init_scratch(dev)) return nullptr;
  }
  return &d;
}

unsigned int* next_tile_counter(cudaStream_t stream) {
  static std::atomic<unsigned> seq[2];
  static const bool enabled = [] {
    const char* e = getenv("BSMM_SCHED");
    return !(e && e[0] == 's');  // BSMM_SCHED=static: round-robin (for A/B runs)
  }();
  if (!enabled) return nullptr;
  bool capturing = false;
  DeviceScratch* d = scratch_for(stream, &capturing);
  if (!d) return nullptr;  // static schedule for this launch
  const int pool = capturing ? 1 : 0;
  return d->counters.load(std::memory_order_acquire) + pool * kTileCounterSlots +
         (seq[pool].fetch_add(1, std::memory_order_relaxed) % kTileCounterSlots);
}

// Device-side kernel selection for hint-less callers ("auto pair").  The reference's API has no
// topology hints, so an unmodified caller would never reach the 2-SM kernels.  For products
// large enough to hide ~5 us, the entry points instead (1) run topology_gate_kernel, which
// writes the hint bits the topology really has into a gate word, (2) launch the 2-SM kernel
// gated on those bits and (3) the 1-SM kernel gated on their absence: exactly one of the two
// runs, the other one's CTAs return immediately; no host synchronisation, the result does not
// depend on which ran.  BSMM_AUTO_PAIR=0 disables it, =1 applies it to every size (tests).
struct Gate {
  const unsigned int* word = nullptr;
  unsigned int mask = 0;
  int want = 1;
};
// Whole-operand L2 warm-up of the 1-SM kernel (GemmParams::warm_*).  EXPERIMENT, off by default:
// BSMM_WARM=<MB> (read on every call) turns it on for calls whose two input operands together are
// at most that large.  Measured with a limit of 80 MB (profiles/r02_warm_ab.txt): neutral on C4
// DSD / DDS up to 50 % density, 3-10 % SLOWER on SDD and above 75 % — the compulsory first touches
// are not what the short launches wait for, and the burst competes with the pipeline's own loads.
void set_warm(GemmParams& p, const void* a, unsigned long long a_bytes, const void* b,
              unsigned long long b_bytes) {
  unsigned long long limit = 0;
  if (const char* e = getenv("BSMM_WARM")) limit = (unsigned long long)(atoll(e)) << 20;
  if (!p.prefetch || a_bytes + b_bytes > limit) return;
  p.warm_ptr[0] = a; p.warm_bytes[0] = a_bytes & ~15ull;
  p.warm_ptr[1] = b; p.warm_bytes[1] = b_bytes & ~15ull;
}
// BSMM_PREFETCH=0 (read on every call) turns the pre-wait L2 prefetch off (A/B runs).
bool prefetch_enabled() {
  const char* e = getenv("BSMM_PREFETCH");
  return !(e && atoi(e) == 0);
}
void apply_gate(GemmParams& p, const Gate& g) {
  p.gate = g.word;
  p.gate_mask = g.mask;
  p.gate_want = g.want;
  // a gated launch waits for the previous grid before anything else: nothing to overlap with
  p.prefetch = (!g.word && prefetch_enabled()) ? 1 : 0;
  p.raster = 0;  // set by the 2-SM DSD path (run_dsd_like)
}
unsigned int* next_gate_slot(cudaStream_t stream) {
  static std::atomic<unsigned> seq[2];
  bool capturing = false;
  DeviceScratch* d = scratch_for(stream, &capturing);
  if (!d) return nullptr;
  const int pool = capturing ? 1 : 0;
  return d->gates.load(std::memory_order_acquire) + pool * kGateSlots +
         (seq[pool].fetch_add(1, std::memory_order_relaxed) % kGateSlots);
}
// flops of the call decide: below ~5e10 the two extra launches (~4 us) cost what the faster
// kernel saves.  (C2 / C3 are 6.9e10: without hints the transposed-walk products run 61 vs 74 us
// and DDS 61 vs 69 us with the device-side selection; DSD nn is a wash, 64 vs 60 us.)
bool auto_pair_wanted(double flops) {
  if (const char* e = getenv("BSMM_AUTO_PAIR")) return atoi(e) != 0;
  return flops >= 5e10;
}
bool pdl_enabled();
void add_pdl(cudaLaunchConfig_t& cfg, cudaLaunchAttribute* attrs);
// Writes the gate word for a walk (rows: offsets / indices; columns: offsets_t / indices_t).
int launch_gate(const int* offsets, const short* indices, int walk_rows, unsigned twin_bit,
                unsigned even_bit, unsigned int* slot, cudaStream_t stream) {
  cudaLaunchConfig_t cfg{};
  cudaLaunchAttribute attr[1];
  cfg.gridDim = dim3(1);
  cfg.blockDim = dim3(1024);
  cfg.stream = stream;
  add_pdl(cfg, attr);
  BSMM_CUDA(cudaLaunchKernelEx(&cfg, topology_gate_kernel, offsets, indices, walk_rows, twin_bit,
                               even_bit, slot));
  g_launches.fetch_add(1, std::memory_order_relaxed);
  BSMM_CUDA(cudaGetLastError());
  return BSMM_OK;
}

// 2-D row-major view [outer][inner] of 16-bit elements with a 64-element inner box and the
// 128 B swizzle the UMMA descriptors in ptx.cuh assume.  Out-of-bounds box elements read 0.
// Element type of a tensor map: fp16 / bf16 (2 B) or e4m3 (1 B, BSMM_F8E4M3 operands).
// `bk()` = elements per 128 B = K elements per pipeline stage = edge of an MN-major slab.
struct ElemType {
  bool bf16, f8;
  ElemType(bool bf16_) : bf16(bf16_), f8(false) {}
  ElemType(bool bf16_, bool f8_) : bf16(bf16_), f8(f8_) {}
  uint32_t size() const { return f8 ? 1u : 2u; }
  uint32_t bk() const { return 128u / size(); }
};

// box_inner = 0: 128 B worth of elements (what the 128 B swizzle and the UMMA descriptors assume).
int make_map(CUtensorMap* map, ElemType t, const void* base, uint64_t inner, uint64_t outer,
             uint64_t ld_elems, uint32_t box_outer, uint32_t box_inner = 0, bool swizzle = true) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return cuda_fail(cudaErrorNotSupported);
  if (box_inner == 0) box_inner = t.bk();
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {ld_elems * t.size()};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  auto encode = [&] {
    return fn(map,
              t.f8 ? CU_TENSOR_MAP_DATA_TYPE_UINT8
                   : (t.bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16),
              2,
              const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
              swizzle ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
              CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  };
  CUresult r = encode();
  if (r == CUDA_ERROR_INVALID_CONTEXT) {
    // A host thread that has only used the runtime lazily (e.g. PyTorch's autograd workers) may
    // have no driver context bound yet: bind the runtime's current device and retry.
    cudaFree(nullptr);
    r = encode();
  }
  if (r != CUDA_SUCCESS) {
    if (getenv("BSMM_VERBOSE"))
      fprintf(stderr, "bsmm: cuTensorMapEncodeTiled failed (%d): base=%p inner=%llu outer=%llu ld=%llu box=%ux%u\n",
              int(r), base, (unsigned long long)inner, (unsigned long long)outer,
              (unsigned long long)ld_elems, box_inner, box_outer);
    return cuda_fail(cudaErrorInvalidValue, __LINE__);
  }
  return BSMM_OK;
}

// Output tensor map of the TMA-store epilogue: the dense (or BCSR value) output as a 2-D array
// [rows][cols] with leading dimension `ld`; box 64 x 32 in the 128 B swizzle for row-major
// stores, 32 x 32 unswizzled for the transposed (DDS as DSD^T) store.  `out` may be null
// (bsmm_dsd_scatter: per-row pointers, no map): then `fallback` is copied, never used.
int make_out_map(CUtensorMap* map, bool bf16, const void* out, uint64_t rows, uint64_t cols,
                 uint64_t ld, bool transposed_tile, const CUtensorMap& fallback) {
  if (!out || rows == 0 || cols == 0) {
    *map = fallback;
    return BSMM_OK;
  }
  return transposed_tile ? make_map(map, bf16, out, cols, rows, ld, 32, 32, false)
                         : make_map(map, bf16, out, cols, rows, ld, 32, 64, true);
}

// Programmatic dependent launch: the kernel's prologue (barrier init, TMEM allocation, tensor-map
// prefetch) may overlap the tail of the previous kernel in the stream; the kernels block in
// griddepcontrol.wait before their first global access.  BSMM_PDL=0 disables it.
bool pdl_enabled() {
  static const bool on = [] {
    const char* e = getenv("BSMM_PDL");
    return !(e && atoi(e) == 0);
  }();
  return on;
}
void add_pdl(cudaLaunchConfig_t& cfg, cudaLaunchAttribute* attrs) {
  if (!pdl_enabled()) return;
  cudaLaunchAttribute& a = attrs[cfg.numAttrs];
  a.id = cudaLaunchAttributeProgrammaticStreamSerialization;
  a.val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attrs;
  cfg.numAttrs += 1;
}

// Launch one kernel variant.  CL > 1 launches thread-block clusters of CL CTAs; `units` is then
// the number of tile GROUPS (clusters' worth of work) and the grid is sized to the number of
// clusters that can be co-resident (cudaOccupancyMaxActiveClusters, cached per device).
template <int OP, bool A_MN, bool B_MN, int BN, bool TRANS_OUT, int CL>
int launch_one(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& mc,
               const GemmParams& p, long long units, int sms, cudaStream_t stream) {
  auto* kern = &bsmm_gemm_kernel<OP, A_MN, B_MN, BN, TRANS_OUT, CL>;
  constexpr int smem = Tiling<BN>::kSmemBytes;
  static std::atomic<unsigned long long> configured{0};  // bit per device
  static int max_clusters[64] = {0};
  int dev = 0;
  BSMM_CUDA(cudaGetDevice(&dev));
  const unsigned long long bit = 1ull << (dev & 63);
  cudaLaunchConfig_t cfg{};
  cudaLaunchAttribute attr[2];
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  if (CL > 1) {
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CL;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
  }
  if (!(configured.load(std::memory_order_acquire) & bit)) {
    BSMM_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    if (CL > 1) {
      int n = 0;
      cfg.gridDim = dim3(unsigned(sms / CL * CL));
      BSMM_CUDA(cudaOccupancyMaxActiveClusters(&n, kern, &cfg));
      max_clusters[dev & 63] = n > 0 ? n : 1;
    }
    configured.fetch_or(bit, std::memory_order_release);
  }
  long long groups = units;
  const long long cap = CL > 1 ? max_clusters[dev & 63] : sms;
  if (groups > cap) groups = cap;
  cfg.gridDim = dim3(unsigned(groups * CL));
  add_pdl(cfg, attr);
  BSMM_CUDA(cudaLaunchKernelEx(&cfg, kern, ma, mb, mc, p));
  g_launches.fetch_add(1, std::memory_order_relaxed);
  BSMM_CUDA(cudaGetLastError());
  return BSMM_OK;
}

// 2-SM kernels (static cluster dims (2,1,1), PairTiling shared memory): per (kernel, device) state.
struct ClusterKernelState {
  std::atomic<unsigned long long> configured{0};  // bit per device
  int max_clusters[64] = {0};
};

// First use per (kernel, device): opt in to the dynamic shared-memory size and ask how many
// clusters can be co-resident.  Returns that capacity.
template <typename Kern>
int pair_capacity(Kern* kern, ClusterKernelState& st, int sms, int* cap) {
  constexpr int smem = PairTiling::kSmemBytes;
  int dev = 0;
  BSMM_CUDA(cudaGetDevice(&dev));
  const unsigned long long bit = 1ull << (dev & 63);
  if (!(st.configured.load(std::memory_order_acquire) & bit)) {
    BSMM_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    cudaLaunchConfig_t cfg{};
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.gridDim = dim3(unsigned(sms / 2 * 2));
    int n = 0;
    BSMM_CUDA(cudaOccupancyMaxActiveClusters(&n, kern, &cfg));
    st.max_clusters[dev & 63] = n > 0 ? n : 1;
    st.configured.fetch_or(bit, std::memory_order_release);
  }
  *cap = st.max_clusters[dev & 63];
  return BSMM_OK;
}

template <typename Kern, typename... Maps>
int launch_pair_grid(Kern* kern, long long clusters, const GemmParams& p, cudaStream_t stream,
                     const Maps&... maps) {
  cudaLaunchConfig_t cfg{};
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = PairTiling::kSmemBytes;
  cfg.stream = stream;
  cfg.gridDim = dim3(unsigned(clusters * 2));
  cudaLaunchAttribute pdl_attr[1];
  add_pdl(cfg, pdl_attr);
  BSMM_CUDA(cudaLaunchKernelEx(&cfg, kern, maps..., p));
  g_launches.fetch_add(1, std::memory_order_relaxed);
  BSMM_CUDA(cudaGetLastError());
  return BSMM_OK;
}

// grid = 2 x min(tiles, co-resident clusters)
template <typename Kern>
int launch_pair_tiles(Kern* kern, ClusterKernelState& st, const CUtensorMap& ma,
                      const CUtensorMap& mb, const CUtensorMap& mc, const GemmParams& p,
                      long long tiles, int sms, cudaStream_t stream) {
  int cap = 0, rc;
  if ((rc = pair_capacity(kern, st, sms, &cap))) return rc;
  return launch_pair_grid(kern, std::min<long long>(tiles, cap), p, stream, ma, mb, mc);
}

template <bool A_MN, bool B_MN, bool TRANS_OUT>
int launch_pair_one(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& mc,
                    const GemmParams& p, long long tiles, int sms, cudaStream_t stream) {
  static ClusterKernelState st;
  return launch_pair_tiles(&bsmm_pair_kernel<A_MN, B_MN, TRANS_OUT>, st, ma, mb, mc, p, tiles, sms,
                           stream);
}

template <bool A_MN, bool B_MN>
int launch_pair_sdd_one(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& mc,
                        const GemmParams& p, long long tiles, int sms, cudaStream_t stream) {
  static ClusterKernelState st;
  return launch_pair_tiles(&bsmm_pair_sdd_kernel<A_MN, B_MN>, st, ma, mb, mc, p, tiles, sms, stream);
}

template <bool A_MN, bool B_MN>
int launch_pair_dds_one(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& mc,
                        const GemmParams& p, long long tiles, int sms, cudaStream_t stream) {
  static ClusterKernelState st;
  return launch_pair_tiles(&bsmm_pair_dds_kernel<A_MN, B_MN>, st, ma, mb, mc, p, tiles, sms, stream);
}

// Split-K workspace: kSkRegions regions of (co-resident clusters) x 2 CTAs x 128 KiB fp32 slots
// plus their flags, allocated per device at first use (with the scheduler counters, the only
// memory the library owns).  Launches take regions round-robin and stamp their flags with a
// fresh epoch, so nothing is ever reset; two launches would have to be kSkRegions launches
// apart and still overlap in time to share a region.
constexpr unsigned kSkRegions = 4;
struct SkWorkspace {
  float* ws = nullptr;
  unsigned int* flags = nullptr;
  int slots = 0;  // clusters per region
};
bool sk_workspace(int dev, int clusters, GemmParams* p) {
  static SkWorkspace w[64];
  static std::once_flag once[64];
  static std::atomic<unsigned> epoch{0};
  if (dev < 0 || dev >= 64) return false;
  std::call_once(once[dev], [dev, clusters] {
    const size_t slot = 2 * kSkSlotFloats * sizeof(float);
    float* a = nullptr;
    unsigned int* f = nullptr;
    if (cudaMalloc(&a, kSkRegions * clusters * slot) != cudaSuccess) { cudaGetLastError(); return; }
    if (cudaMalloc(&f, kSkRegions * clusters * 8 * sizeof(unsigned int)) != cudaSuccess ||
        cudaMemset(f, 0, kSkRegions * clusters * 8 * sizeof(unsigned int)) != cudaSuccess) {
      cudaGetLastError();
      cudaFree(a);
      return;
    }
    w[dev].ws = a;
    w[dev].flags = f;
    w[dev].slots = clusters;
  });
  if (!w[dev].ws || clusters > w[dev].slots) return false;
  unsigned e = epoch.fetch_add(1, std::memory_order_relaxed) + 1u;
  if (e == 0) e = epoch.fetch_add(1, std::memory_order_relaxed) + 1u;  // 0 = "never written"
  const unsigned region = e % kSkRegions;
  p->sk_ws = w[dev].ws + size_t(region) * w[dev].slots * 2 * kSkSlotFloats;
  p->sk_flags = w[dev].flags + size_t(region) * w[dev].slots * 8;
  p->sk_epoch = e;
  return true;
}

// Split-K of the remainder wave (bsmm_pair_sk_kernel.cuh): opt-in with BSMM_SPLITK=1 (read on
// every call), never chosen automatically.  Measured on B200 (profiles/r01_streamk_ab.txt) the
// hand-off — parking 128 KiB of fp32 accumulators per CTA, the flag round trip, folding them back
// through TMEM — costs 13-18 us per launch, more than the 7 us the better balance saves on C2
// (54.2 -> 60.3 us).  Kept as a tested experiment (tests/test_gpu_splitk.py).
bool split_k_requested() {
  const char* e = getenv("BSMM_SPLITK");
  return e && atoi(e) != 0;
}

template <bool A_MN, bool B_MN, bool TRANS_OUT>
int launch_pair_sk_one(const CUtensorMap& ma, const CUtensorMap& mb, GemmParams p,
                       long long tiles, long long units, int sms, cudaStream_t stream,
                       bool* launched) {
  static ClusterKernelState st;
  auto* kern = &bsmm_pair_sk_kernel<A_MN, B_MN, TRANS_OUT>;
  *launched = false;
  if (!split_k_requested() || tiles <= 0 || units <= 0) return BSMM_OK;
  int cap = 0, dev = 0, rc;
  if ((rc = pair_capacity(kern, st, sms, &cap))) return rc;
  if (tiles % cap == 0) return BSMM_OK;  // no remainder wave
  BSMM_CUDA(cudaGetDevice(&dev));
  if (!sk_workspace(dev, cap, &p)) return BSMM_OK;  // no workspace: classic schedule
  // every co-resident cluster is launched: those without a tile of their own take tails
  if ((rc = launch_pair_grid(kern, cap, p, stream, ma, mb))) return rc;
  *launched = true;
  return BSMM_OK;
}

template <bool TRANS_OUT>
int launch_pair_sk(bool a_mn, bool b_mn, const CUtensorMap& ma, const CUtensorMap& mb,
                   const GemmParams& p, long long tiles, long long units, int sms, cudaStream_t s,
                   bool* launched) {
  if (a_mn) {
    return b_mn ? launch_pair_sk_one<true, true, TRANS_OUT>(ma, mb, p, tiles, units, sms, s, launched)
                : launch_pair_sk_one<true, false, TRANS_OUT>(ma, mb, p, tiles, units, sms, s, launched);
  }
  return b_mn ? launch_pair_sk_one<false, true, TRANS_OUT>(ma, mb, p, tiles, units, sms, s, launched)
              : launch_pair_sk_one<false, false, TRANS_OUT>(ma, mb, p, tiles, units, sms, s, launched);
}

template <bool TRANS_OUT>
int launch_pair(bool a_mn, bool b_mn, const CUtensorMap& ma, const CUtensorMap& mb,
                const CUtensorMap& mc, const GemmParams& p, long long tiles, int sms,
                cudaStream_t s) {
  if (a_mn) {
    return b_mn ? launch_pair_one<true, true, TRANS_OUT>(ma, mb, mc, p, tiles, sms, s)
                : launch_pair_one<true, false, TRANS_OUT>(ma, mb, mc, p, tiles, sms, s);
  }
  return b_mn ? launch_pair_one<false, true, TRANS_OUT>(ma, mb, mc, p, tiles, sms, s)
              : launch_pair_one<false, false, TRANS_OUT>(ma, mb, mc, p, tiles, sms, s);
}

bool pair_enabled() {
  static const bool on = [] {
    const char* e = getenv("BSMM_PAIR");
    return !(e && atoi(e) == 0);
  }();
  return on;
}

template <int OP, int BN, bool TRANS_OUT, int CL>
int launch_variant(bool a_mn, bool b_mn, const CUtensorMap& ma, const CUtensorMap& mb,
                   const CUtensorMap& mc, const GemmParams& p, long long units, int sms,
                   cudaStream_t s) {
  if (a_mn) {
    return b_mn ? launch_one<OP, true, true, BN, TRANS_OUT, CL>(ma, mb, mc, p, units, sms, s)
                : launch_one<OP, true, false, BN, TRANS_OUT, CL>(ma, mb, mc, p, units, sms, s);
  }
  return b_mn ? launch_one<OP, false, true, BN, TRANS_OUT, CL>(ma, mb, mc, p, units, sms, s)
              : launch_one<OP, false, false, BN, TRANS_OUT, CL>(ma, mb, mc, p, units, sms, s);
}

constexpr char kEpilogueDefault = 't';  // 's'taged (ld/st), 'd'irect (registers), 't'ma store
// Output format word of the kernels (kFmt* in bsmm_kernel.cuh) from dtype + epilogue flags.
inline int fmt_word(bool bf16, unsigned flags) {
  int act = kActNone;
  if (flags & BSMM_FLAG_GELU) act = kActGelu;
  else if (flags & BSMM_FLAG_RELU) act = kActRelu;
  // BSMM_EPI=staged|direct (read on every call; A/B runs): how row-major epilogues store.
  const char* e = getenv("BSMM_EPI");
  const char mode = e ? e[0] : kEpilogueDefault;
  return (bf16 ? kFmtBf16 : 0) | (act << kFmtActShift) |
         ((flags & BSMM_FLAG_ACCUMULATE) ? kFmtAccum : 0) | (mode == 'd' ? kFmtDirect : 0) |
         (mode == 't' ? kFmtTma : 0);
}
// Scattered output rows (bsmm_dsd_scatter) cannot use the tiled TMA store.  BSMM_SCATTER_EPI=direct
// (A/B switch, read on every call) makes them use the register-direct stores instead of the
// staged ld.shared -> st.global path.
inline int scatter_fmt(void* const* c_rows) {
  if (!c_rows) return 0;
  const char* e = getenv("BSMM_SCATTER_EPI");
  return (e && e[0] == 'd') ? kFmtDirect : 0;
}
// What an output without any contribution becomes: zeros — act(0) = 0 for both activations — or,
// when accumulating, C itself (nothing to do).
inline bool zero_fill_needed(unsigned flags) { return !(flags & BSMM_FLAG_ACCUMULATE); }

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

struct Shape { int m, n, k; };

int check_common(int dtype, int block_size, bool allow_f8 = false) {
  if (dtype != BSMM_F16 && dtype != BSMM_BF16 && !(allow_f8 && dtype == BSMM_F8E4M3))
    return BSMM_ERR_DTYPE;
  if (block_size != 128) return BSMM_ERR_BLOCK_SIZE;
  return BSMM_OK;
}

int check_sparse(const bsmm_block_matrix* s) {
  if (s->rows < 0 || s->cols < 0 || s->nonzeros < 0) return BSMM_ERR_SHAPE;
  if (s->rows % 128 || s->cols % 128 || s->nonzeros % (128 * 128)) return BSMM_ERR_MISALIGNED;
  if (s->rows / 128 > 32767 || s->cols / 128 > 32767) return BSMM_ERR_RANGE;
  if (s->nonzeros > 0 && (!s->data || !s->offsets || !s->indices)) return BSMM_ERR_MISSING_META;
  if (!s->offsets) return BSMM_ERR_MISSING_META;
  if (!aligned16(s->data)) return BSMM_ERR_MISALIGNED;
  return BSMM_OK;
}

// `align` = 8 elements for 16-bit matrices (the reference's align8 kernels), 16 for e4m3 operands
// (rows of a TMA view must be multiples of 16 B).
int check_dense(const bsmm_matrix* d, int align = 8) {
  if (d->rows < 0 || d->cols < 0) return BSMM_ERR_SHAPE;
  if (d->rows % align || d->cols % align) return BSMM_ERR_MISALIGNED;
  if (d->rows > 0 && d->cols > 0 && !d->data) return BSMM_ERR_MISSING_META;
  if (!aligned16(d->data)) return BSMM_ERR_MISALIGNED;
  return BSMM_OK;
}

template <class A, class B, class C>
int check_shapes(const A* a, bool ta, const B* b, bool tb, const C* c, Shape* sh) {
  // sputnik/block/arguments.h:164-231 (MatmulShape + ValidMatmul)
  sh->m = ta ? a->cols : a->rows;
  sh->k = ta ? a->rows : a->cols;
  sh->n = tb ? b->rows : b->cols;
  const int kb = tb ? b->cols : b->rows;
  if (kb != sh->k || c->rows != sh->m || c->cols != sh->n) return BSMM_ERR_SHAPE;
  return BSMM_OK;
}

int transposed_meta_present(const bsmm_block_matrix* s) {
  return (s->offsets_t && s->indices_t && s->block_offsets) ? BSMM_OK : BSMM_ERR_MISSING_META;
}

// BSMM_VALIDATE_HINTS=1 (debug): check on the device that the hint flags a call is about to act on
// hold for the walk it will use; synchronous.  `claimed` = the hint bits of the call restricted to
// the ones this entry point consumes; `by_cols`: the walk follows offsets_t / indices_t.
bool validate_hints_requested() {
  const char* e = getenv("BSMM_VALIDATE_HINTS");
  return e && atoi(e) != 0;
}
int validate_hints(const bsmm_block_matrix* s, unsigned claimed, bool by_cols, cudaStream_t stream) {
  if (!claimed || !validate_hints_requested() || stream_is_capturing(stream)) return BSMM_OK;
  const int walk_rows = (by_cols ? s->cols : s->rows) / 128;
  const int* offsets = static_cast<const int*>(by_cols ? s->offsets_t : s->offsets);
  const short* indices = static_cast<const short*>(by_cols ? s->indices_t : s->indices);
  const unsigned twin_bit = by_cols ? BSMM_FLAG_TWIN_COLS : BSMM_FLAG_TWIN_ROWS;
  const unsigned even_bit = by_cols ? 0u : BSMM_FLAG_EVEN_ROWS;
  if (!offsets || walk_rows == 0) return (claimed & twin_bit) ? BSMM_ERR_HINT : BSMM_OK;
  unsigned* d = nullptr;
  unsigned init = twin_bit | even_bit, out = 0;
  BSMM_CUDA(cudaMallocAsync(&d, sizeof(unsigned), stream));
  BSMM_CUDA(cudaMemcpyAsync(d, &init, sizeof(unsigned), cudaMemcpyHostToDevice, stream));
  topology_hints_kernel<<<(walk_rows / 2 + 256) / 256, 256, 0, stream>>>(offsets, indices, walk_rows,
                                                                        twin_bit, even_bit, d);
  g_launches.fetch_add(1, std::memory_order_relaxed);
  BSMM_CUDA(cudaGetLastError());
  BSMM_CUDA(cudaMemcpyAsync(&out, d, sizeof(unsigned), cudaMemcpyDeviceToHost, stream));
  BSMM_CUDA(cudaFreeAsync(d, stream));
  BSMM_CUDA(cudaStreamSynchronize(stream));
  return (claimed & ~out) ? BSMM_ERR_HINT : BSMM_OK;
}

int do_transpose(const bsmm_block_matrix* a, cudaStream_t stream) {
  const int br = a->rows / 128, bc = a->cols / 128, nb = a->nonzeros / (128 * 128);
  static std::atomic<unsigned long long> configured{0};
  int dev = 0;
  BSMM_CUDA(cudaGetDevice(&dev));
  const unsigned long long bit = 1ull << (dev & 63);
  constexpr int kMaxSmem = 200 * 1024;
  constexpr int kMcSmem = kTrListCap * 4;
  if (!(configured.load(std::memory_order_acquire) & bit)) {
    BSMM_CUDA(cudaFuncSetAttribute(transpose_metadata_kernel,
                                   cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxSmem));
    BSMM_CUDA(cudaFuncSetAttribute(transpose_metadata_mc_kernel,
                                   cudaFuncAttributeMaxDynamicSharedMemorySize, kMcSmem));
    configured.fetch_or(bit, std::memory_order_release);
  }
  // BSMM_TRANSPOSE=single (read on every call): the one-CTA kernel (A/B runs, tests)
  const char* mode = getenv("BSMM_TRANSPOSE");
  const bool single = mode && mode[0] == 's';
  int sms = 0, st;
  if ((st = current_device_sms(&sms))) return st;
  // Multi-CTA kernel: CTA b owns `cpc` block columns.  nb < 2^17 (int nonzeros) and
  // cpc <= kTrMaxOwnCols always hold on a device with >= 128 SMs; otherwise fall back.
  if (!single && bc > 0 && nb > 0 && nb < (1 << 17)) {
    int grid = std::min(sms, bc);
    const int cpc = (bc + grid - 1) / grid;
    if (cpc <= kTrMaxOwnCols) {
      grid = (bc + cpc - 1) / cpc;
      cudaLaunchConfig_t cfg{};
      cudaLaunchAttribute attr[1];
      cfg.gridDim = dim3(unsigned(grid));
      cfg.blockDim = dim3(kTrThreads);
      cfg.dynamicSmemBytes = kMcSmem;
      cfg.stream = stream;
      add_pdl(cfg, attr);
      BSMM_CUDA(cudaLaunchKernelEx(
          &cfg, transpose_metadata_mc_kernel, static_cast<const int*>(a->offsets),
          static_cast<const short*>(a->indices), br, bc, nb, cpc, static_cast<int*>(a->offsets_t),
          static_cast<short*>(a->indices_t), static_cast<int*>(a->block_offsets)));
      g_launches.fetch_add(1, std::memory_order_relaxed);
      BSMM_CUDA(cudaGetLastError());
      return BSMM_OK;
    }
  }
  const int hist_bytes = (bc + 1) * 4;
  int seg_cap = (kMaxSmem - hist_bytes) / 4;
  if (seg_cap > br) seg_cap = br;
  if (seg_cap < 0) seg_cap = 0;
  const int smem = hist_bytes + seg_cap * 4;
  transpose_metadata_kernel<<<1, kTransposeThreads, smem, stream>>>(
      static_cast<const int*>(a->offsets), static_cast<const short*>(a->indices), br, bc, nb,
      seg_cap, static_cast<int*>(a->offsets_t), static_cast<short*>(a->indices_t),
      static_cast<int*>(a->block_offsets));
  g_launches.fetch_add(1, std::memory_order_relaxed);
  BSMM_CUDA(cudaGetLastError());
  return BSMM_OK;
}

}  // namespace
}  // namespace bsmm

using namespace bsmm;

#ifdef BSMM_TRACE
extern "C" unsigned long long* g_trace_buffer;
#define BSMM_SET_TRACE(p) (p).trace = g_trace_buffer
#else
#define BSMM_SET_TRACE(p)
#endif

extern "C" {

int bsmm_sdd_check(const bsmm_matrix* a, int ta, const bsmm_matrix* b, int tb,
                   const bsmm_block_matrix* c, int dtype) {
  if (!a || !b || !c) return BSMM_ERR_MISSING_META;
  int st = check_common(dtype, c->block_size, true);
  if (st) return st;
  Shape sh;
  const int al = dtype == BSMM_F8E4M3 ? 16 : 8;   // operand alignment in elements (16 B rows)
  if ((st = check_shapes(a, ta != 0, b, tb != 0, c, &sh))) return st;
  if ((st = check_dense(a, al)) || (st = check_dense(b, al)) || (st = check_sparse(c))) return st;
  if (c->nonzeros > 0 && !c->row_indices) return BSMM_ERR_MISSING_META;  // sdd_..._nn_align8.cu:75
  return BSMM_OK;
}

int bsmm_dsd_check(const bsmm_block_matrix* a, int ta, const bsmm_matrix* b, int tb,
                   const bsmm_matrix* c, int dtype, unsigned flags) {
  if (flags & BSMM_FLAG_SWIGLU) return BSMM_ERR_SHAPE;  // an epilogue of bsmm_sdd only
  if (!a || !b || !c) return BSMM_ERR_MISSING_META;
  int st = check_common(dtype, a->block_size, true);
  if (st) return st;
  Shape sh;
  const int al = dtype == BSMM_F8E4M3 ? 16 : 8;
  if ((st = check_shapes(a, ta != 0, b, tb != 0, c, &sh))) return st;
  if ((st = check_sparse(a)) || (st = check_dense(b, al)) || (st = check_dense(c))) return st;
  if (ta && (st = transposed_meta_present(a))) return st;  // dsd_..._tn_align8.cu:73-75
  return BSMM_OK;
}

int bsmm_dds_check(const bsmm_matrix* a, int ta, const bsmm_block_matrix* b, int tb,
                   const bsmm_matrix* c, int dtype, unsigned flags) {
  if (flags & BSMM_FLAG_SWIGLU) return BSMM_ERR_SHAPE;  // an epilogue of bsmm_sdd only
  if (!a || !b || !c) return BSMM_ERR_MISSING_META;
  int st = check_common(dtype, b->block_size, true);
  if (st) return st;
  Shape sh;
  const int al = dtype == BSMM_F8E4M3 ? 16 : 8;
  if ((st = check_shapes(a, ta != 0, b, tb != 0, c, &sh))) return st;
  if ((st = check_dense(a, al)) || (st = check_sparse(b)) || (st = check_dense(c))) return st;
  if (!tb && (st = transposed_meta_present(b))) return st;  // dds_..._nn_align8.cu:74-84
  return BSMM_OK;
}

namespace {
int sdd_impl(const bsmm_matrix* a, int ta, const bsmm_matrix* b, int tb,
             const bsmm_block_matrix* c, int dtype, unsigned flags, cudaStream_t stream,
             Gate gate) {
  int st;
  const int nb = c->nonzeros / (128 * 128);
  const int m = c->rows, n = c->cols, k = ta ? a->rows : a->cols;
  if (nb == 0) return BSMM_OK;
  // BSMM_F8E4M3: e4m3 operands, bf16 output
  const bool f8 = dtype == BSMM_F8E4M3, bf16 = dtype != BSMM_F16, a_mn = ta != 0, b_mn = tb == 0;
  const ElemType ot(bf16, f8);
  const int bk = int(ot.bk());
  if (k == 0) {
    if (zero_fill_needed(flags))
      BSMM_CUDA(cudaMemsetAsync(c->data, 0, size_t(c->nonzeros) * 2, stream));
    return BSMM_OK;
  }
  CUtensorMap ma, mb;
  const bool swiglu = (flags & BSMM_FLAG_SWIGLU) != 0;
  const int nbw = swiglu ? 2 * n : n;  // columns of op(B): gate and up panels interleaved
  // A stored [m][k] (K-major) or [k][m] (MN-major); B stored [k][n] (MN-major) or [n][k].
  if ((st = a_mn ? make_map(&ma, ot, a->data, m, k, m, bk)
                 : make_map(&ma, ot, a->data, k, m, k, 128)))
    return st;
  if ((st = b_mn ? make_map(&mb, ot, b->data, nbw, k, nbw, bk)
                 : make_map(&mb, ot, b->data, k, nbw, k, 128)))
    return st;
  CUtensorMap mc;  // the BCSR values as [nb * 128][128]
  if ((st = make_out_map(&mc, bf16, c->data, uint64_t(nb) * 128, 128, 128, false, ma))) return st;
  GemmParams p{};
  p.sdd_rows = static_cast<const short*>(c->row_indices);
  p.sdd_cols = static_cast<const short*>(c->indices);
  p.c = c->data;
  p.ldc = 128;
  p.sparse_lim = m; p.dense_lim = n;
  p.nb = nb;
  if (flags & BSMM_FLAG_DEVICE_NNZ) p.nnz_dev = static_cast<const int*>(c->offsets) + m / 128;
  int sms = 0;
  if ((st = current_device_sms(&sms))) return st;
  // Tile ids are block ids.  An even block and its odd neighbour in the same block row form one
  // 128 x 256 tile (the odd id is skipped when drawn); every other block is a 128 x 128 tile of
  // its own — so a tile never spans two block rows and the dynamic scheduler can place the
  // small ones freely (C4 SDD 50 %: the two-pass straddling tiles used to set the critical
  // path).  With few blocks, one block per tile so that all of them run concurrently.
  static const bool allow_single = [] {
    const char* e = getenv("BSMM_SDD_SINGLE");
    return !(e && atoi(e) == 0);
  }();
  p.sdd_single = allow_single && nb <= sms;
  p.tiles_major = nb; p.tiles_minor = 1;
  p.k_steps = (k + bk - 1) / bk;
  p.bk = bk;
  p.f8 = f8;
  p.idesc = ptx::make_idesc(bf16 && !f8, a_mn, b_mn, 128, 256);
  p.idesc_half = ptx::make_idesc(bf16 && !f8, a_mn, b_mn, 128, 128);
  p.bf16 = fmt_word(bf16, flags);
  BSMM_SET_TRACE(p);
  apply_gate(p, gate);
  if (swiglu) {   // one output block per CTA, 128 x 256 accumulator (gate | up)
    p.swiglu = 1;
    p.sdd_single = 0;
    p.prefetch = 0;
    if ((flags & BSMM_FLAG_TWIN_ROWS) && pair_enabled() && nb % 2 == 0 && (m / 128) % 2 == 0) {
      // twin block rows (hint): 2-SM kernel, tile = rows 2p, 2p + 1 of one output block column;
      // CTA r loads its block row of op(A) and panel 2j + r of op(B) (gate / up)
      p.walk_offsets = static_cast<const int*>(c->offsets);
      p.tiles_major = nb / 2;
      p.idesc = ptx::make_idesc(bf16 && !f8, a_mn, b_mn, 256, 256);
      if (a_mn) {
        return b_mn ? launch_pair_sdd_one<true, true>(ma, mb, mc, p, p.tiles_major, sms, stream)
                    : launch_pair_sdd_one<true, false>(ma, mb, mc, p, p.tiles_major, sms, stream);
      }
      return b_mn ? launch_pair_sdd_one<false, true>(ma, mb, mc, p, p.tiles_major, sms, stream)
                  : launch_pair_sdd_one<false, false>(ma, mb, mc, p, p.tiles_major, sms, stream);
    }
    if (p.tiles_major > sms) p.tile_counter = next_tile_counter(stream);
    return launch_variant<kOpSDD, 256, false, 1>(a_mn, b_mn, ma, mb, mc, p, p.tiles_major, sms, stream);
  }
  constexpr unsigned kClique = BSMM_FLAG_TWIN_ROWS | BSMM_FLAG_EVEN_ROWS;
  const bool clique_shape = pair_enabled() && nb % 4 == 0 && (m / 128) % 2 == 0;
  if ((flags & kClique) != kClique && !gate.word && clique_shape &&
      auto_pair_wanted(2.0 * nb * 128.0 * 128.0 * k)) {
    // No hints: let the device decide between the clique kernel and the 1-SM kernel (see Gate).
    if (unsigned int* slot = next_gate_slot(stream)) {
      if ((st = launch_gate(static_cast<const int*>(c->offsets),
                            static_cast<const short*>(c->indices), m / 128, BSMM_FLAG_TWIN_ROWS,
                            BSMM_FLAG_EVEN_ROWS, slot, stream)))
        return st;
      if ((st = sdd_impl(a, ta, b, tb, c, dtype, flags | kClique, stream, Gate{slot, kClique, 1})))
        return st;
      return sdd_impl(a, ta, b, tb, c, dtype, flags & ~kClique, stream, Gate{slot, kClique, 0});
    }
  }
  if ((flags & kClique) == kClique && clique_shape) {
    // 2-SM kernel over 2 x 2 cliques of output blocks (bsmm_pair_kernel.cuh)
    p.walk_offsets = static_cast<const int*>(c->offsets);
    p.tiles_major = nb / 4;
    p.idesc = ptx::make_idesc(bf16 && !f8, a_mn, b_mn, 256, 256);
    if (a_mn) {
      return b_mn ? launch_pair_sdd_one<true, true>(ma, mb, mc, p, p.tiles_major, sms, stream)
                  : launch_pair_sdd_one<true, false>(ma, mb, mc, p, p.tiles_major, sms, stream);
    }
    return b_mn ? launch_pair_sdd_one<false, true>(ma, mb, mc, p, p.tiles_major, sms, stream)
                : launch_pair_sdd_one<false, false>(ma, mb, mc, p, p.tiles_major, sms, stream);
  }
  if (p.tiles_major > sms) p.tile_counter = next_tile_counter(stream);
  set_warm(p, a->data, 1ull * a->rows * a->cols * ot.size(), b->data,
           1ull * b->rows * b->cols * ot.size());
  return launch_variant<kOpSDD, 256, false, 1>(a_mn, b_mn, ma, mb, mc, p, p.tiles_major, sms, stream);
}
}  // namespace

int bsmm_sdd(const bsmm_matrix* a, int ta, const bsmm_matrix* b, int tb,
             const bsmm_block_matrix* c, int dtype, unsigned flags, void* stream_) {
  int st;
  if (flags & BSMM_FLAG_SWIGLU) {
    // op(B) carries gate and up panels: validate against the view with half its columns
    if (!b) return BSMM_ERR_MISSING_META;
    if (flags & (BSMM_FLAG_ACCUMULATE | BSMM_FLAG_RELU | BSMM_FLAG_GELU)) return BSMM_ERR_SHAPE;
    bsmm_matrix half = *b;
    int& nb2 = tb ? half.rows : half.cols;
    if (nb2 % 256) return nb2 % 2 ? BSMM_ERR_SHAPE : BSMM_ERR_MISALIGNED;
    nb2 /= 2;
    st = bsmm_sdd_check(a, ta, &half, tb, c, dtype);
  } else {
    st = bsmm_sdd_check(a, ta, b, tb, c, dtype);
  }
  if (st) return st;
  constexpr unsigned kClique = BSMM_FLAG_TWIN_ROWS | BSMM_FLAG_EVEN_ROWS;
  if ((flags & kClique) == kClique &&
      (st = validate_hints(c, kClique, false, static_cast<cudaStream_t>(stream_))))
    return st;
  return sdd_impl(a, ta, b, tb, c, dtype, flags, static_cast<cudaStream_t>(stream_), Gate());
}

namespace {
// DSD, and DDS as the transposed DSD problem C^T = op(B)^T op(A)^T (trans_out).
//   sp / sp_t : the sparse operand and whether it is used transposed (column walk);
//   dn / dn_t : the dense operand as stored and whether it is stored [dense dim][k] ("K-major");
//   dense_ext : extent of the output along the dense operand's free dimension.
int run_dsd_like(const bsmm_block_matrix* sp, bool sp_t, const bsmm_matrix* dn, bool dn_kmajor,
                 int dense_ext, int k, void* cdata, long long ldc, bool trans_out, bool bf16,
                 unsigned flags, cudaStream_t stream, void* const* c_rows = nullptr,
                 Gate gate = Gate(), bool f8 = false) {
  // `bf16`: the OUTPUT type; `f8`: e4m3 operands (then bf16 is true)
  const ElemType ot(bf16, f8);
  const int bk = int(ot.bk());
  const int nb = sp->nonzeros / (128 * 128);
  const int sparse_ext = sp_t ? sp->cols : sp->rows;
  int st;
  // C is [sparse_ext][dense_ext] (DSD), or its transpose [dense_ext][sparse_ext] (DDS: trans_out),
  // leading dimension ldc; scattered rows (c_rows) have no map.
  auto out_map = [&](CUtensorMap* mc, const CUtensorMap& fallback) {
    return make_out_map(mc, bf16, c_rows ? nullptr : cdata, trans_out ? dense_ext : sparse_ext,
                        trans_out ? sparse_ext : dense_ext, uint64_t(ldc), trans_out, fallback);
  };
  const int bn = dense_ext > 128 ? 256 : 128;
  const int tiles_minor = (dense_ext + bn - 1) / bn;
  // Twin walk rows (caller's hint for this walk direction): 2-SM kernel over CTA pairs.
  const unsigned need = sp_t ? BSMM_FLAG_TWIN_COLS : BSMM_FLAG_TWIN_ROWS;
  const bool twins = (flags & need) != 0;
  const bool pair_shape = pair_enabled() && bn == 256 && (sparse_ext / 128) % 2 == 0;
  if (!twins && !gate.word && pair_shape &&
      auto_pair_wanted(2.0 * nb * 128.0 * 128.0 * dense_ext)) {
    // No hint: let the device decide (see Gate).  Both launches go through this function again.
    if (unsigned int* slot = next_gate_slot(stream)) {
      if ((st = launch_gate(static_cast<const int*>(sp_t ? sp->offsets_t : sp->offsets),
                            static_cast<const short*>(sp_t ? sp->indices_t : sp->indices),
                            sparse_ext / 128, need, 0u, slot, stream)))
        return st;
      if ((st = run_dsd_like(sp, sp_t, dn, dn_kmajor, dense_ext, k, cdata, ldc, trans_out, bf16,
                             flags | need, stream, c_rows, Gate{slot, need, 1}, f8)))
        return st;
      return run_dsd_like(sp, sp_t, dn, dn_kmajor, dense_ext, k, cdata, ldc, trans_out, bf16, flags,
                          stream, c_rows, Gate{slot, need, 0}, f8);
    }
  }
  if (twins && pair_shape) {
    CUtensorMap ma, mb;
    if ((st = make_map(&ma, ot, sp->data, 128, uint64_t(nb) * 128, 128, sp_t ? bk : 128)))
      return st;
    if ((st = dn_kmajor ? make_map(&mb, ot, dn->data, k, dense_ext, k, 128)
                        : make_map(&mb, ot, dn->data, dense_ext, k, dense_ext, bk)))
      return st;
    CUtensorMap mc;
    if ((st = out_map(&mc, ma))) return st;
    GemmParams p{};
    p.walk_offsets = static_cast<const int*>(sp_t ? sp->offsets_t : sp->offsets);
    p.walk_indices = static_cast<const short*>(sp_t ? sp->indices_t : sp->indices);
    p.walk_blocks = sp_t ? static_cast<const int*>(sp->block_offsets) : nullptr;
    p.c = cdata;
    p.c_rows = c_rows;
    apply_gate(p, gate);
    p.ldc = ldc;
    p.sparse_lim = sparse_ext; p.dense_lim = dense_ext;
    p.tiles_major = sparse_ext / 128;
    p.tiles_minor = tiles_minor;
    // Group-rasterised tile order (pair_dsd_tile) for long walks, where an expert's panel of the
    // dense operand is what gets re-streamed; short walks (transposed dMoE walks: 8 blocks) keep
    // the dense-tile-fastest order, which shares the sparse blocks.  BSMM_RASTER=0|1 overrides.
    {
      const char* r = getenv("BSMM_RASTER");
      p.raster = r ? atoi(r) != 0 : (1ll * nb >= 64ll * (sparse_ext / 128));
    }
    p.idesc = ptx::make_idesc(bf16 && !f8, sp_t, !dn_kmajor, 256, 256);
    p.bf16 = fmt_word(bf16, flags) | scatter_fmt(c_rows);
    p.bk = bk;
    p.f8 = f8;
    BSMM_SET_TRACE(p);
    int sms = 0;
    if ((st = current_device_sms(&sms))) return st;
    const long long tiles = 1ll * (p.tiles_major / 2) * p.tiles_minor;
    // work in sparse blocks (per walk-row pair x dense tiles), for the split-K estimate
    const long long units = 1ll * (nb / 2) * p.tiles_minor;
    bool launched = f8;  // the split-K experiment is 16-bit only
    if (!f8)
    st = trans_out ? launch_pair_sk<true>(sp_t, !dn_kmajor, ma, mb, p, tiles, units, sms, stream, &launched)
                   : launch_pair_sk<false>(sp_t, !dn_kmajor, ma, mb, p, tiles, units, sms, stream, &launched);
    if (st || (launched && !f8)) return st;
    return trans_out ? launch_pair<true>(sp_t, !dn_kmajor, ma, mb, mc, p, tiles, sms, stream)
                     : launch_pair<false>(sp_t, !dn_kmajor, ma, mb, mc, p, tiles, sms, stream);
  }
  // Cluster size along the dense dimension: the CTAs of a cluster TMA-multicast the sparse block
  // to each other.  Measured on B200 (profiles/r01_cluster_ab.txt) it is neutral within +-2 % on
  // most shapes (lockstep CTAs already share L2 lines; the L2 still serves one sector stream per
  // destination SM) and helps only shallow transposed walks, so the default stays 1 and
  // BSMM_CLUSTER=2|4 selects it for experiments.
  int cl = 1;
  if (const char* e = f8 ? nullptr : getenv("BSMM_CLUSTER")) {
    const int want = atoi(e);
    if (want == 4 && bn == 256 && tiles_minor % 4 == 0) cl = 4;
    if (want == 2 && bn == 256 && tiles_minor % 2 == 0) cl = 2;
  }
  CUtensorMap ma, mb;
  // Sparse values as a 2-D [nb * 128][128] array.  K-major reads 128 / cl block rows per TMA;
  // the transposed (MN-major) read takes 64-row slabs (cl <= 2) or half slabs (cl = 4).
  const uint32_t sp_box = sp_t ? (cl == 4 ? 32 : uint32_t(bk)) : uint32_t(128 / cl);
  if ((st = make_map(&ma, ot, sp->data, 128, uint64_t(nb) * 128, 128, sp_box))) return st;
  // K-major dense operand: boxes of 128 rows (a 256-wide tile takes two, a halved tile one)
  if ((st = dn_kmajor ? make_map(&mb, ot, dn->data, k, dense_ext, k, 128)
                      : make_map(&mb, ot, dn->data, dense_ext, k, dense_ext, bk)))
    return st;
  CUtensorMap mc;
  if ((st = out_map(&mc, ma))) return st;
  GemmParams p{};
  p.walk_offsets = static_cast<const int*>(sp_t ? sp->offsets_t : sp->offsets);
  p.walk_indices = static_cast<const short*>(sp_t ? sp->indices_t : sp->indices);
  p.walk_blocks = sp_t ? static_cast<const int*>(sp->block_offsets) : nullptr;
  p.c = cdata;
  p.c_rows = c_rows;
  apply_gate(p, gate);
  p.ldc = ldc;
  p.sparse_lim = sparse_ext; p.dense_lim = dense_ext;
  p.tiles_major = sparse_ext / 128;
  p.tiles_minor = tiles_minor;
  p.idesc = ptx::make_idesc(bf16 && !f8, sp_t, !dn_kmajor, 128, bn);
  p.bf16 = fmt_word(bf16, flags) | scatter_fmt(c_rows);
  p.bk = bk;
  p.f8 = f8;
  BSMM_SET_TRACE(p);
  int sms = 0;
  if ((st = current_device_sms(&sms))) return st;
  const long long units = 1ll * p.tiles_major * (p.tiles_minor / cl);
  if (cl == 1 && units > sms) p.tile_counter = next_tile_counter(stream);
  // Tail-wave split (EXPERIMENT, off by default; BSMM_TAIL_SPLIT=1, read on every call): when the
  // last wave of 128 x 256 tiles would fill at most half of the SMs, hand those tiles out as
  // twice as many 128 x 128 half tiles (C4: 512 tiles on 148 SMs = 3.46 waves -> 3.5 instead of
  // 4).  Measured (profiles/r02_tail_split_ab.txt): a half tile still streams the whole sparse
  // block row, and the 1-SM kernel is bound by that L2 -> SM feed, so it takes nearly as long as a
  // full tile: -4 % at 10 % density, neutral at 25 %, +3..10 % (slower) from 50 % up.
  if (cl == 1 && bn == 256 && p.tile_counter) {
    const char* e = getenv("BSMM_TAIL_SPLIT");
    const long long r = units % sms;
    if (e && atoi(e) != 0 && r > 0 && 2 * r <= sms) {
      p.split_from = int(units - r);
      p.idesc_half = ptx::make_idesc(bf16 && !f8, sp_t, !dn_kmajor, 128, 128);
    }
  }
  if (cl == 1)
    set_warm(p, sp->data, 1ull * nb * 128 * 128 * ot.size(), dn->data,
             1ull * dn->rows * dn->cols * ot.size());
  const bool a_mn = sp_t, b_mn = !dn_kmajor;
#define BSMM_DSD(BN_, CL_)                                                                      \
  (trans_out ? launch_variant<kOpDSD, BN_, true, CL_>(a_mn, b_mn, ma, mb, mc, p, units, sms, stream) \
             : launch_variant<kOpDSD, BN_, false, CL_>(a_mn, b_mn, ma, mb, mc, p, units, sms, stream))
  if (bn == 128) return BSMM_DSD(128, 1);
  if (cl == 4) return BSMM_DSD(256, 4);
  if (cl == 2) return BSMM_DSD(256, 2);
  return BSMM_DSD(256, 1);
#undef BSMM_DSD
}
}  // namespace

int bsmm_dsd(const bsmm_block_matrix* a, int ta, const bsmm_matrix* b, int tb,
             const bsmm_matrix* c, int dtype, unsigned flags, void* stream_) {
  int st = bsmm_dsd_check(a, ta, b, tb, c, dtype, flags);
  if (st) return st;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const int nb = a->nonzeros / (128 * 128);
  const int m = c->rows, n = c->cols, k = ta ? a->rows : a->cols;
  if (m == 0 || n == 0) return BSMM_OK;
  if (nb == 0) {  // every block row is empty: the output is all zeros
    if (zero_fill_needed(flags))
      BSMM_CUDA(cudaMemsetAsync(c->data, 0, size_t(m) * size_t(n) * 2, stream));
    return BSMM_OK;
  }
  if (ta && (flags & BSMM_FLAG_CREATE_METADATA)) {
    if ((st = do_transpose(a, stream))) return st;
  }
  if ((st = validate_hints(a, flags & (ta ? BSMM_FLAG_TWIN_COLS : BSMM_FLAG_TWIN_ROWS), ta != 0,
                           stream)))
    return st;
  // B stored [k][n] (tb = 0, MN-major) or [n][k] (tb = 1, K-major).
  return run_dsd_like(a, ta != 0, b, tb != 0, n, k, c->data, n, false, dtype != BSMM_F16, flags,
                      stream, nullptr, Gate(), dtype == BSMM_F8E4M3);
}

int bsmm_gather_rows(void* const* src_row_ptrs, int rows, int cols, void* dst, void* stream_) {
  if (rows < 0 || cols < 0) return BSMM_ERR_SHAPE;
  if (cols % 8 || !aligned16(dst)) return BSMM_ERR_MISALIGNED;
  if (rows == 0 || cols == 0) return BSMM_OK;
  if (!src_row_ptrs || !dst) return BSMM_ERR_MISSING_META;
  int sms = 0, st;
  if ((st = current_device_sms(&sms))) return st;
  const int blocks = std::min((rows + 7) / 8, sms * 8);  // 8 warps (rows) per CTA
  gather_rows_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      src_row_ptrs, rows, cols / 8, static_cast<uint4*>(dst));
  g_launches.fetch_add(1, std::memory_order_relaxed);
  BSMM_CUDA(cudaGetLastError());
  return BSMM_OK;
}

int bsmm_dsd_scatter(const bsmm_block_matrix* a, int ta, const bsmm_matrix* b, int tb, int c_rows,
                     int c_cols, void* const* c_row_ptrs, int dtype, unsigned flags,
                     void* stream_) {
  alignas(16) static char placeholder[16];  // shape validation only; never dereferenced
  const bsmm_matrix c{c_rows, c_cols, placeholder};
  int st = bsmm_dsd_check(a, ta, b, tb, &c, dtype, flags);
  if (st) return st;
  if (c_rows == 0 || c_cols == 0) return BSMM_OK;
  if (!c_row_ptrs) return BSMM_ERR_MISSING_META;
  flags &= ~BSMM_FLAG_ACCUMULATE;  // scattered rows are written, never accumulated
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const int k = ta ? a->rows : a->cols;
  if (a->nonzeros == 0) {
    scatter_zero_rows_kernel<<<296, 256, 0, stream>>>(c_row_ptrs, c_rows, c_cols / 8);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    BSMM_CUDA(cudaGetLastError());
    return BSMM_OK;
  }
  if (ta && (flags & BSMM_FLAG_CREATE_METADATA)) {
    if ((st = do_transpose(a, stream))) return st;
  }
  if ((st = validate_hints(a, flags & (ta ? BSMM_FLAG_TWIN_COLS : BSMM_FLAG_TWIN_ROWS), ta != 0,
                           stream)))
    return st;
  return run_dsd_like(a, ta != 0, b, tb != 0, c_cols, k, nullptr, c_cols, false,
                      dtype != BSMM_F16, flags, stream, c_row_ptrs, Gate(), dtype == BSMM_F8E4M3);
}

namespace {
int dds_impl(const bsmm_matrix* a, int ta, const bsmm_block_matrix* b, int tb,
             const bsmm_matrix* c, int dtype, unsigned flags, cudaStream_t stream, Gate gate) {
  int st;
  const int nb = b->nonzeros / (128 * 128);
  const int m = c->rows, n = c->cols, k = tb ? b->cols : b->rows;
  // Twin walk rows (hint) and more than one 128-row tile of op(A): direct 2-SM DDS kernel.
  {
    const bool by_cols = tb == 0;
    const unsigned need = by_cols ? BSMM_FLAG_TWIN_COLS : BSMM_FLAG_TWIN_ROWS;
    const bool twins = (flags & need) != 0;
    if (!twins && !gate.word && pair_enabled() && m > 128 && (n / 128) % 2 == 0 &&
        auto_pair_wanted(2.0 * nb * 128.0 * 128.0 * m)) {
      // No hint: let the device decide (see Gate).
      if (unsigned int* slot = next_gate_slot(stream)) {
        if ((st = launch_gate(static_cast<const int*>(by_cols ? b->offsets_t : b->offsets),
                              static_cast<const short*>(by_cols ? b->indices_t : b->indices),
                              n / 128, need, 0u, slot, stream)))
          return st;
        if ((st = dds_impl(a, ta, b, tb, c, dtype, flags | need, stream, Gate{slot, need, 1})))
          return st;
        return dds_impl(a, ta, b, tb, c, dtype, flags, stream, Gate{slot, need, 0});
      }
    }
    static const bool direct = [] {
      const char* e = getenv("BSMM_DDS_DIRECT");
      return !(e && atoi(e) == 0);
    }();
    if (twins && direct && pair_enabled() && m > 128 && (n / 128) % 2 == 0) {
      const bool f8 = dtype == BSMM_F8E4M3, bf16 = dtype != BSMM_F16;
      const bool a_mn = ta != 0, b_mn = by_cols;
      const ElemType ot(bf16, f8);
      const int bk = int(ot.bk());
      CUtensorMap ma, mb;
      if ((st = a_mn ? make_map(&ma, ot, a->data, m, k, m, bk)
                     : make_map(&ma, ot, a->data, k, m, k, 128)))
        return st;
      if ((st = make_map(&mb, ot, b->data, 128, uint64_t(nb) * 128, 128, b_mn ? bk : 128)))
        return st;
      CUtensorMap mc;
      if ((st = make_out_map(&mc, bf16, c->data, uint64_t(m), uint64_t(n), uint64_t(n), false, ma)))
        return st;
      GemmParams p{};
      p.walk_offsets = static_cast<const int*>(by_cols ? b->offsets_t : b->offsets);
      p.walk_indices = static_cast<const short*>(by_cols ? b->indices_t : b->indices);
      p.walk_blocks = by_cols ? static_cast<const int*>(b->block_offsets) : nullptr;
      p.c = c->data;
      p.ldc = n;
      p.sparse_lim = n; p.dense_lim = m;
      p.tiles_major = n / 128;
      p.tiles_minor = (m + 255) / 256;
      p.idesc = ptx::make_idesc(bf16 && !f8, a_mn, b_mn, 256, 256);
      p.bf16 = fmt_word(bf16, flags);
      p.bk = bk;
      p.f8 = f8;
      apply_gate(p, gate);
      int sms = 0;
      if ((st = current_device_sms(&sms))) return st;
      const long long tiles = 1ll * (p.tiles_major / 2) * p.tiles_minor;
      if (a_mn) {
        return b_mn ? launch_pair_dds_one<true, true>(ma, mb, mc, p, tiles, sms, stream)
                    : launch_pair_dds_one<true, false>(ma, mb, mc, p, tiles, sms, stream);
      }
      return b_mn ? launch_pair_dds_one<false, true>(ma, mb, mc, p, tiles, sms, stream)
                  : launch_pair_dds_one<false, false>(ma, mb, mc, p, tiles, sms, stream);
    }
  }
  // C^T = op(B)^T op(A)^T.  op(B)^T walks the stored B by block column unless tb (then by its
  // block rows); op(A)^T has A stored [m][k] (ta = 0: K-major for the N side) or [k][m].
  return run_dsd_like(b, tb == 0, a, ta == 0, m, k, c->data, n, true, dtype != BSMM_F16, flags,
                      stream, nullptr, gate, dtype == BSMM_F8E4M3);
}
}  // namespace

int bsmm_dds(const bsmm_matrix* a, int ta, const bsmm_block_matrix* b, int tb,
             const bsmm_matrix* c, int dtype, unsigned flags, void* stream_) {
  int st = bsmm_dds_check(a, ta, b, tb, c, dtype, flags);
  if (st) return st;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const int m = c->rows, n = c->cols;
  if (m == 0 || n == 0) return BSMM_OK;
  if (b->nonzeros == 0) {
    if (zero_fill_needed(flags))
      BSMM_CUDA(cudaMemsetAsync(c->data, 0, size_t(m) * size_t(n) * 2, stream));
    return BSMM_OK;
  }
  if (!tb && (flags & BSMM_FLAG_CREATE_METADATA)) {
    if ((st = do_transpose(b, stream))) return st;
  }
  if ((st = validate_hints(b, flags & (tb ? BSMM_FLAG_TWIN_ROWS : BSMM_FLAG_TWIN_COLS), tb == 0,
                           stream)))
    return st;
  return dds_impl(a, ta, b, tb, c, dtype, flags, stream, Gate());
}

// ------------------------------------------------------------------ SSD / SDS / DSS siblings
namespace {
extern "C++" {
template <bool A_MN, bool B_MN>
int launch_sibling_one(const CUtensorMap& ma, const CUtensorMap& mb, const SiblingParams& p,
                       int sms, cudaStream_t stream) {
  auto* kern = &bsmm_sibling_kernel<A_MN, B_MN>;
  constexpr int smem = SiblingTiling::kSmemBytes;
  static std::atomic<unsigned long long> configured{0};  // bit per device
  int dev = 0;
  BSMM_CUDA(cudaGetDevice(&dev));
  const unsigned long long bit = 1ull << (dev & 63);
  if (!(configured.load(std::memory_order_acquire) & bit)) {
    BSMM_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured.fetch_or(bit, std::memory_order_release);
  }
  cudaLaunchConfig_t cfg{};
  cudaLaunchAttribute attr[1];
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cfg.gridDim = dim3(unsigned(std::min(p.num_tiles, sms)));
  add_pdl(cfg, attr);
  BSMM_CUDA(cudaLaunchKernelEx(&cfg, kern, ma, mb, p));
  g_launches.fetch_add(1, std::memory_order_relaxed);
  BSMM_CUDA(cudaGetLastError());
  return BSMM_OK;
}
}  // extern "C++"

// Row view or column view of a stored sparse matrix, as the walk of one block row of op(X).
void set_walk(const bsmm_block_matrix* s, bool by_cols, const int** offsets, const short** indices,
              const int** blocks) {
  *offsets = static_cast<const int*>(by_cols ? s->offsets_t : s->offsets);
  *indices = static_cast<const short*>(by_cols ? s->indices_t : s->indices);
  *blocks = by_cols ? static_cast<const int*>(s->block_offsets) : nullptr;
}

// Exactly one of (sa, da), (sb, db), (sc, dc) is non-null per operand.
int sibling_impl(int op, const bsmm_block_matrix* sa, const bsmm_matrix* da, bool ta,
                 const bsmm_block_matrix* sb, const bsmm_matrix* db, bool tb,
                 const bsmm_block_matrix* sc, const bsmm_matrix* dc, int dtype, unsigned flags,
                 cudaStream_t stream) {
  int st;
  const bool bf16 = dtype == BSMM_BF16;
  const int m = sc ? sc->rows : dc->rows, n = sc ? sc->cols : dc->cols;
  const int k = sa ? (ta ? sa->rows : sa->cols) : (ta ? da->rows : da->cols);
  const size_t out_elems = sc ? size_t(sc->nonzeros) : size_t(m) * size_t(n);
  void* cdata = sc ? sc->data : dc->data;
  if (out_elems == 0) return BSMM_OK;
  if (k == 0 || (sa && sa->nonzeros == 0) || (sb && sb->nonzeros == 0)) {
    if (zero_fill_needed(flags))  // nothing to contract
      BSMM_CUDA(cudaMemsetAsync(cdata, 0, out_elems * 2, stream));
    return BSMM_OK;
  }
  // op(A)'s block rows are the stored A's block columns when transposed; op(B)'s block columns
  // are the stored B's block columns unless transposed.
  const bool a_cols = sa && ta, b_cols = sb && !tb;
  if (flags & BSMM_FLAG_CREATE_METADATA) {
    if (a_cols && (st = do_transpose(sa, stream))) return st;
    if (b_cols && (st = do_transpose(sb, stream))) return st;
  }
  const bool a_mn = ta, b_mn = !tb;
  CUtensorMap ma, mb;
  if (sa) {
    const int nb = sa->nonzeros / (128 * 128);
    if ((st = make_map(&ma, bf16, sa->data, 128, uint64_t(nb) * 128, 128, a_mn ? 64 : 128)))
      return st;
  } else if ((st = a_mn ? make_map(&ma, bf16, da->data, m, k, m, 64)
                        : make_map(&ma, bf16, da->data, k, m, k, 128))) {
    return st;
  }
  if (sb) {
    const int nb = sb->nonzeros / (128 * 128);
    if ((st = make_map(&mb, bf16, sb->data, 128, uint64_t(nb) * 128, 128, b_mn ? 64 : 128)))
      return st;
  } else if ((st = b_mn ? make_map(&mb, bf16, db->data, n, k, n, 64)
                        : make_map(&mb, bf16, db->data, k, n, k, 128))) {
    return st;
  }
  SiblingParams p{};
  p.op = op;
  if (sa) set_walk(sa, a_cols, &p.a_offsets, &p.a_indices, &p.a_blocks);
  if (sb) set_walk(sb, b_cols, &p.b_offsets, &p.b_indices, &p.b_blocks);
  p.c = cdata;
  int sms = 0;
  if ((st = current_device_sms(&sms))) return st;
  if (sc) {
    p.c_rows = static_cast<const short*>(sc->row_indices);
    p.c_cols = static_cast<const short*>(sc->indices);
    p.ldc = 128;
    p.num_tiles = sc->nonzeros / (128 * 128);
    p.single = p.num_tiles <= sms;
  } else {
    p.ldc = n;
    p.tiles_n = n / 128;
    p.num_tiles = (m / 128) * (n / 128);
  }
  p.idesc = ptx::make_idesc(bf16, a_mn, b_mn, 128, 256);
  p.idesc_half = ptx::make_idesc(bf16, a_mn, b_mn, 128, 128);
  p.bf16 = fmt_word(bf16, flags) & ~(kFmtTma | kFmtDirect);
  if (p.num_tiles > sms) p.tile_counter = next_tile_counter(stream);
  if (a_mn) {
    return b_mn ? launch_sibling_one<true, true>(ma, mb, p, sms, stream)
                : launch_sibling_one<true, false>(ma, mb, p, sms, stream);
  }
  return b_mn ? launch_sibling_one<false, true>(ma, mb, p, sms, stream)
              : launch_sibling_one<false, false>(ma, mb, p, sms, stream);
}
}  // namespace

int bsmm_ssd_check(const bsmm_block_matrix* a, int ta, const bsmm_matrix* b, int tb,
                   const bsmm_block_matrix* c, int dtype, unsigned flags) {
  if (flags & BSMM_FLAG_SWIGLU) return BSMM_ERR_SHAPE;  // an epilogue of bsmm_sdd only
  if (!a || !b || !c) return BSMM_ERR_MISSING_META;
  int st = check_common(dtype, c->block_size);  // ssd.cu:11-14
  if (st) return st;
  if (a->block_size != 128) return BSMM_ERR_BLOCK_SIZE;
  Shape sh;
  if ((st = check_shapes(a, ta != 0, b, tb != 0, c, &sh))) return st;
  if ((st = check_sparse(a)) || (st = check_dense(b)) || (st = check_sparse(c))) return st;
  if (ta && (st = transposed_meta_present(a))) return st;  // ssd_..._tn_align8.cu:75-77
  if (c->nonzeros > 0 && !c->row_indices) return BSMM_ERR_MISSING_META;  // ssd_..._tn_align8.cu:88
  return BSMM_OK;
}

int bsmm_sds_check(const bsmm_matrix* a, int ta, const bsmm_block_matrix* b, int tb,
                   const bsmm_block_matrix* c, int dtype, unsigned flags) {
  if (flags & BSMM_FLAG_SWIGLU) return BSMM_ERR_SHAPE;  // an epilogue of bsmm_sdd only
  if (!a || !b || !c) return BSMM_ERR_MISSING_META;
  int st = check_common(dtype, c->block_size);  // sds.cu:11-14
  if (st) return st;
  if (b->block_size != 128) return BSMM_ERR_BLOCK_SIZE;
  Shape sh;
  if ((st = check_shapes(a, ta != 0, b, tb != 0, c, &sh))) return st;
  if ((st = check_dense(a)) || (st = check_sparse(b)) || (st = check_sparse(c))) return st;
  if (!tb && (st = transposed_meta_present(b))) return st;
  if (c->nonzeros > 0 && !c->row_indices) return BSMM_ERR_MISSING_META;
  return BSMM_OK;
}

int bsmm_dss_check(const bsmm_block_matrix* a, int ta, const bsmm_block_matrix* b, int tb,
                   const bsmm_matrix* c, int dtype, unsigned flags) {
  if (flags & BSMM_FLAG_SWIGLU) return BSMM_ERR_SHAPE;  // an epilogue of bsmm_sdd only
  if (!a || !b || !c) return BSMM_ERR_MISSING_META;
  int st = check_common(dtype, a->block_size);
  if (st) return st;
  if (b->block_size != 128) return BSMM_ERR_BLOCK_SIZE;
  Shape sh;
  if ((st = check_shapes(a, ta != 0, b, tb != 0, c, &sh))) return st;
  if ((st = check_sparse(a)) || (st = check_sparse(b)) || (st = check_dense(c))) return st;
  if (ta && (st = transposed_meta_present(a))) return st;   // dss_..._tn_align8.cu:76-83
  if (!tb && (st = transposed_meta_present(b))) return st;
  if (sh.k / 128 > kSibMaxKBlocks) return BSMM_ERR_RANGE;   // reference: 32768 (index_merge.h:103)
  return BSMM_OK;
}

int bsmm_ssd(const bsmm_block_matrix* a, int ta, const bsmm_matrix* b, int tb,
             const bsmm_block_matrix* c, int dtype, unsigned flags, void* stream) {
  int st = bsmm_ssd_check(a, ta, b, tb, c, dtype, flags);
  if (st) return st;
  return sibling_impl(kSibSSD, a, nullptr, ta != 0, nullptr, b, tb != 0, c, nullptr, dtype, flags,
                      static_cast<cudaStream_t>(stream));
}

int bsmm_sds(const bsmm_matrix* a, int ta, const bsmm_block_matrix* b, int tb,
             const bsmm_block_matrix* c, int dtype, unsigned flags, void* stream) {
  int st = bsmm_sds_check(a, ta, b, tb, c, dtype, flags);
  if (st) return st;
  return sibling_impl(kSibSDS, nullptr, a, ta != 0, b, nullptr, tb != 0, c, nullptr, dtype, flags,
                      static_cast<cudaStream_t>(stream));
}

int bsmm_dss(const bsmm_block_matrix* a, int ta, const bsmm_block_matrix* b, int tb,
             const bsmm_matrix* c, int dtype, unsigned flags, void* stream) {
  int st = bsmm_dss_check(a, ta, b, tb, c, dtype, flags);
  if (st) return st;
  return sibling_impl(kSibDSS, a, nullptr, ta != 0, b, nullptr, tb != 0, nullptr, c, dtype, flags,
                      static_cast<cudaStream_t>(stream));
}

int bsmm_bitmask_bytes(const bsmm_block_matrix* m, size_t* bytes) {
  if (!m || !bytes) return BSMM_ERR_MISSING_META;
  if (m->block_size != 128) return BSMM_ERR_BLOCK_SIZE;
  if (m->rows % 128 || m->cols % 128) return BSMM_ERR_MISALIGNED;
  const bool trans = m->offsets_t != nullptr;  // bitmask.h:16-23
  const int walk_rows = (trans ? m->cols : m->rows) / 128;
  const int walk_cols = (trans ? m->rows : m->cols) / 128;
  *bytes = size_t((walk_cols + 63) / 64) * size_t(walk_rows) * sizeof(unsigned long long);
  return BSMM_OK;
}

int bsmm_bitmask(const bsmm_block_matrix* m, void* bitmask, void* stream_) {
  size_t bytes = 0;
  int st = bsmm_bitmask_bytes(m, &bytes);
  if (st) return st;
  if (m->nonzeros % (128 * 128)) return BSMM_ERR_MISALIGNED;
  if (bytes == 0) return BSMM_OK;
  const bool trans = m->offsets_t != nullptr;  // bitmask.cu:8-10: the transposed view wins
  if (!bitmask || !m->offsets || (trans && m->nonzeros > 0 && !m->indices_t) ||
      (!trans && m->nonzeros > 0 && !m->indices))
    return BSMM_ERR_MISSING_META;
  const int walk_rows = (trans ? m->cols : m->rows) / 128;
  const int walk_cols = (trans ? m->rows : m->cols) / 128;
  const int blocks = std::min((walk_rows + 7) / 8, 1184);
  bitmask_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      static_cast<const int*>(trans ? m->offsets_t : m->offsets),
      static_cast<const short*>(trans ? m->indices_t : m->indices), walk_rows,
      (walk_cols + 63) / 64, static_cast<unsigned long long*>(bitmask));
  g_launches.fetch_add(1, std::memory_order_relaxed);
  BSMM_CUDA(cudaGetLastError());
  return BSMM_OK;
}

int bsmm_row_indices(const bsmm_block_matrix* a, short* row_indices, void* stream_) {
  if (!a) return BSMM_ERR_MISSING_META;
  if (a->block_size != 128) return BSMM_ERR_BLOCK_SIZE;
  if (a->rows % 128 || a->nonzeros % (128 * 128)) return BSMM_ERR_MISALIGNED;
  const int nb = a->nonzeros / (128 * 128);
  if (nb == 0) return BSMM_OK;
  if (!a->offsets || !row_indices) return BSMM_ERR_MISSING_META;
  const int threads = 256;
  int blocks = (nb + threads - 1) / threads;
  if (blocks > 1184) blocks = 1184;
  row_indices_kernel<<<blocks, threads, 0, static_cast<cudaStream_t>(stream_)>>>(
      static_cast<const int*>(a->offsets), a->rows / 128, nb, row_indices);
  g_launches.fetch_add(1, std::memory_order_relaxed);
  BSMM_CUDA(cudaGetLastError());
  return BSMM_OK;
}

int bsmm_transpose_metadata(const bsmm_block_matrix* a, void* stream_) {
  if (!a) return BSMM_ERR_MISSING_META;
  if (a->block_size != 128) return BSMM_ERR_BLOCK_SIZE;
  // metadata only: `data` is not needed (the reference's Transpose never touches it either)
  if (a->rows < 0 || a->cols < 0 || a->nonzeros < 0) return BSMM_ERR_SHAPE;
  if (a->rows % 128 || a->cols % 128 || a->nonzeros % (128 * 128)) return BSMM_ERR_MISALIGNED;
  if (a->rows / 128 > 32767 || a->cols / 128 > 32767) return BSMM_ERR_RANGE;
  if (!a->offsets || (a->nonzeros > 0 && !a->indices)) return BSMM_ERR_MISSING_META;
  int st;
  if ((st = transposed_meta_present(a))) return st;
  return do_transpose(a, static_cast<cudaStream_t>(stream_));
}

int bsmm_moe_topology(const int* tokens_per_expert, int num_experts, int ffn_blocks,
                      int row_align_blocks, const bsmm_block_matrix* out, int* padded_bins,
                      void* stream_) {
  if (!tokens_per_expert || !out) return BSMM_ERR_MISSING_META;
  if (out->block_size != 128) return BSMM_ERR_BLOCK_SIZE;
  if (num_experts <= 0 || num_experts > kMoeMaxExperts || ffn_blocks <= 0 ||
      (row_align_blocks != 1 && row_align_blocks != 2) || out->rows < 0)
    return BSMM_ERR_SHAPE;
  if (out->rows % (128 * row_align_blocks)) return BSMM_ERR_MISALIGNED;
  const long long cap_rows = out->rows / 128, bcols = 1ll * num_experts * ffn_blocks;
  if (cap_rows > 32767 || bcols > 32767) return BSMM_ERR_RANGE;
  if (out->cols != bcols * 128 || cap_rows * ffn_blocks * 128 * 128 != out->nonzeros)
    return BSMM_ERR_SHAPE;  // also rejects capacities whose element count overflows `nonzeros`
  if (!out->offsets || (cap_rows > 0 && (!out->indices || !out->row_indices)))
    return BSMM_ERR_MISSING_META;
  const int have_t = (out->offsets_t != nullptr) + (out->indices_t != nullptr) +
                     (out->block_offsets != nullptr);
  if (have_t != 0 && have_t != 3 && cap_rows > 0) return BSMM_ERR_MISSING_META;
  if (have_t != 0 && !out->offsets_t) return BSMM_ERR_MISSING_META;
  const long long work = std::max<long long>(cap_rows * ffn_blocks, bcols + 1);
  int blocks = int(std::min<long long>((work + kMoeThreads - 1) / kMoeThreads, 296));
  if (blocks < 1) blocks = 1;
  moe_topology_kernel<<<blocks, kMoeThreads, 0, static_cast<cudaStream_t>(stream_)>>>(
      tokens_per_expert, num_experts, ffn_blocks, row_align_blocks, int(cap_rows),
      static_cast<int*>(out->offsets), static_cast<short*>(out->indices),
      static_cast<short*>(out->row_indices), static_cast<int*>(out->offsets_t),
      static_cast<short*>(out->indices_t), static_cast<int*>(out->block_offsets), padded_bins);
  g_launches.fetch_add(1, std::memory_order_relaxed);
  BSMM_CUDA(cudaGetLastError());
  return BSMM_OK;
}

int bsmm_moe_exchange(const int* counts, int world, int num_experts, int rank, int row_align_blocks,
                      int cap_block_rows, size_t row_pitch_bytes, int num_tables,
                      const void* const* peer_bases, void* const* tables, int* tokens_per_expert,
                      void* stream_) {
  if (!counts || !peer_bases || !tables) return BSMM_ERR_MISSING_META;
  if (world <= 0 || world > kMoeExchangeMaxRanks || num_experts <= 0 ||
      num_experts > kMoeMaxExperts || num_experts % world || rank < 0 || rank >= world ||
      (row_align_blocks != 1 && row_align_blocks != 2) || cap_block_rows < 0 || num_tables < 1 ||
      num_tables > kMoeExchangeMaxTables)
    return BSMM_ERR_SHAPE;
  if (row_pitch_bytes % 16) return BSMM_ERR_MISALIGNED;
  const int el = num_experts / world;
  MoeExchangeBases b{};
  for (int t = 0; t < num_tables; ++t) {
    if (!tables[t]) return BSMM_ERR_MISSING_META;
    b.table[t] = static_cast<unsigned long long*>(tables[t]);
    for (int s = 0; s < world; ++s)
      b.base[t][s] = reinterpret_cast<unsigned long long>(peer_bases[t * world + s]);
  }
  const size_t smem = (size_t(el) + 1 + 2 * size_t(el) * world) * sizeof(int);
  if (smem > 48 * 1024) return BSMM_ERR_RANGE;
  const long long rows = 128ll * cap_block_rows;
  int blocks = int(std::min<long long>((rows + kMoeThreads - 1) / kMoeThreads, 296));
  if (blocks < 1) blocks = 1;
  moe_exchange_kernel<<<blocks, kMoeThreads, smem, static_cast<cudaStream_t>(stream_)>>>(
      counts, world, num_experts, rank * el, el, row_align_blocks, cap_block_rows,
      static_cast<unsigned long long>(row_pitch_bytes), num_tables, b, tokens_per_expert);
  g_launches.fetch_add(1, std::memory_order_relaxed);
  BSMM_CUDA(cudaGetLastError());
  return BSMM_OK;
}

int bsmm_analyze_topology(const bsmm_block_matrix* a, unsigned* hint_flags, void* stream_) {
  if (!a || !hint_flags) return BSMM_ERR_MISSING_META;
  if (a->block_size != 128) return BSMM_ERR_BLOCK_SIZE;
  if (a->rows % 128 || a->cols % 128 || a->nonzeros % (128 * 128)) return BSMM_ERR_MISALIGNED;
  if (!a->offsets || (a->nonzeros > 0 && !a->indices)) return BSMM_ERR_MISSING_META;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const int br = a->rows / 128, bc = a->cols / 128;
  const bool cols = a->offsets_t && a->indices_t;
  unsigned init = BSMM_FLAG_TWIN_ROWS | BSMM_FLAG_EVEN_ROWS | (cols ? BSMM_FLAG_TWIN_COLS : 0u);
  if (br == 0) init &= ~BSMM_FLAG_TWIN_ROWS;
  if (bc == 0) init &= ~BSMM_FLAG_TWIN_COLS;
  unsigned* d = nullptr;
  BSMM_CUDA(cudaMallocAsync(&d, sizeof(unsigned), stream));
  BSMM_CUDA(cudaMemcpyAsync(d, &init, sizeof(unsigned), cudaMemcpyHostToDevice, stream));
  if (br > 0) {
    topology_hints_kernel<<<(br / 2 + 256) / 256, 256, 0, stream>>>(
        static_cast<const int*>(a->offsets), static_cast<const short*>(a->indices), br,
        BSMM_FLAG_TWIN_ROWS, BSMM_FLAG_EVEN_ROWS, d);
    g_launches.fetch_add(1, std::memory_order_relaxed);
  }
  if (cols && bc > 0) {
    topology_hints_kernel<<<(bc / 2 + 256) / 256, 256, 0, stream>>>(
        static_cast<const int*>(a->offsets_t), static_cast<const short*>(a->indices_t), bc,
        BSMM_FLAG_TWIN_COLS, 0u, d);
    g_launches.fetch_add(1, std::memory_order_relaxed);
  }
  BSMM_CUDA(cudaGetLastError());
  unsigned out = 0;
  BSMM_CUDA(cudaMemcpyAsync(&out, d, sizeof(unsigned), cudaMemcpyDeviceToHost, stream));
  BSMM_CUDA(cudaFreeAsync(d, stream));
  BSMM_CUDA(cudaStreamSynchronize(stream));
  *hint_flags = out;
  return BSMM_OK;
}

int bsmm_transpose_buffer_bytes(const bsmm_block_matrix* a, size_t* offsets_t_bytes,
                                size_t* indices_t_bytes, size_t* block_offsets_bytes) {
  if (!a) return BSMM_ERR_MISSING_META;
  if (a->block_size != 128) return BSMM_ERR_BLOCK_SIZE;
  const size_t nb = size_t(a->nonzeros) / (128 * 128);
  if (offsets_t_bytes) *offsets_t_bytes = (size_t(a->cols) / 128 + 1) * sizeof(int);
  if (indices_t_bytes) *indices_t_bytes = nb * sizeof(short);
  if (block_offsets_bytes) *block_offsets_bytes = nb * sizeof(int);
  return BSMM_OK;
}

const char* bsmm_status_string(int status) {
  switch (status) {
    case BSMM_OK: return "ok";
    case BSMM_ERR_BLOCK_SIZE: return "unsupported block size (only 128)";
    case BSMM_ERR_MISALIGNED: return "misaligned dimension or pointer";
    case BSMM_ERR_SHAPE: return "operand shapes do not agree";
    case BSMM_ERR_MISSING_META: return "required pointer or metadata buffer is null";
    case BSMM_ERR_CUDA: return "CUDA call failed";
    case BSMM_ERR_DTYPE: return "unsupported dtype (fp16 / bf16; e4m3 operands for SDD / DSD / DDS only)";
    case BSMM_ERR_RANGE: return "dimension exceeds int16 block index range";
    case BSMM_ERR_HINT: return "a topology hint flag does not hold for this topology (BSMM_VALIDATE_HINTS)";
    case BSMM_ERR_STATIC_LIMIT:
      return "bsmm_dsd_static: sparse operand exceeds BSMM_STATIC_MAX_BLOCKS (16384 blocks) or "
             "BSMM_STATIC_MAX_BLOCK_COLS (2048 block columns)";
    default: return "unknown status";
  }
}

int bsmm_init(void) {
  int dev = 0;
  BSMM_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return cuda_fail(cudaErrorInvalidDevice, __LINE__);
  if (!encode_fn()) return cuda_fail(cudaErrorNotSupported, __LINE__);
  if (!init_scratch(dev)) return cuda_fail(cudaErrorMemoryAllocation, __LINE__);
  return BSMM_OK;
}

int bsmm_last_cuda_error(void) { return g_last_cuda_error; }
int bsmm_last_cuda_error_line(void) { return g_last_cuda_error_line; }
unsigned long long bsmm_launch_count(void) { return g_launches.load(); }
#ifdef BSMM_TRACE
unsigned long long* g_trace_buffer = nullptr;
void bsmm_set_trace(unsigned long long* device_buffer) { g_trace_buffer = device_buffer; }
void bsmm_set_dbg(int flags) { cudaMemcpyToSymbol(bsmm::g_bsmm_dbg, &flags, sizeof(int)); }
#endif
const char* bsmm_version(void) { return "bsmm 0.2 (sm_100a, tcgen05+TMA)"; }

}  // extern "C"

// ======================================================================== allocation-free DSD
// bsmm_dsd_static (include/bsmm.h): no counterpart in the reference; for callers that cannot
// allocate.  All scratch is this file-scope device array.  In HOST code `g_static_scratch` is only
// a stand-in for the symbol — not a device address — so the address each device has for it is
// resolved once per device with cudaGetSymbolAddress (below) and only that pointer is handed to
// kernels.
namespace bsmm {
namespace {
struct StaticScratch {
  unsigned int tile_counter[64];                       // one counter on 256 B of its own
  int offsets_t[BSMM_STATIC_MAX_BLOCK_COLS + 64];      // transposed metadata of the last call
  int block_offsets[BSMM_STATIC_MAX_BLOCKS];
  short indices_t[BSMM_STATIC_MAX_BLOCKS];
};
static_assert(sizeof(StaticScratch) <= 112 * 1024, "keep the resident scratch modest");
static __device__ __align__(256) unsigned char g_static_scratch[sizeof(StaticScratch)];

std::mutex g_static_mu;
StaticScratch* g_static_ptr[64];  // per device, resolved at first use with that device current

int static_scratch_of_current_device(StaticScratch** out, bool refresh = false) {
  int dev = 0;
  BSMM_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return cuda_fail(cudaErrorInvalidDevice, __LINE__);
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (!g_static_ptr[dev] || refresh) {
    void* ptr = nullptr;
    BSMM_CUDA(cudaGetSymbolAddress(&ptr, g_static_scratch));
    if (!ptr) return cuda_fail(cudaErrorInvalidSymbol, __LINE__);
    g_static_ptr[dev] = static_cast<StaticScratch*>(ptr);
  }
  *out = g_static_ptr[dev];
  return BSMM_OK;
}

// make_map without its context-binding retry (which goes through cudaFree(nullptr)): a thread with
// no context bound yet gets one from cudaGetSymbolAddress instead.
int encode_map_static(CUtensorMap* map, ElemType t, const void* base, uint64_t inner,
                      uint64_t outer, uint64_t ld_elems, uint32_t box_outer, uint32_t box_inner,
                      bool swizzle) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return cuda_fail(cudaErrorNotSupported, __LINE__);
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {ld_elems * t.size()};
  cuuint32_t box[2] = {box_inner ? box_inner : t.bk(), box_outer};
  cuuint32_t estr[2] = {1, 1};
  auto encode = [&] {
    return fn(map,
              t.f8 ? CU_TENSOR_MAP_DATA_TYPE_UINT8
                   : (t.bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16),
              2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
              swizzle ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
              CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  };
  CUresult r = encode();
  if (r == CUDA_ERROR_INVALID_CONTEXT) {
    StaticScratch* unused = nullptr;
    int st = static_scratch_of_current_device(&unused, /*refresh=*/true);
    if (st) return st;
    r = encode();
  }
  return r == CUDA_SUCCESS ? BSMM_OK : cuda_fail(cudaErrorInvalidValue, __LINE__);
}
}  // namespace
}  // namespace bsmm

extern "C" int bsmm_dsd_static(const bsmm_block_matrix* a, int ta, const bsmm_matrix* b, int tb,
                               const bsmm_matrix* c, int dtype, unsigned flags, void* stream_) {
  if (!a || !b || !c) return BSMM_ERR_MISSING_META;
  if (flags & BSMM_FLAG_SWIGLU) return BSMM_ERR_SHAPE;  // an epilogue of bsmm_sdd only
  int st = check_common(dtype, a->block_size, true);
  if (st) return st;
  Shape sh;
  const int al = dtype == BSMM_F8E4M3 ? 16 : 8;
  if ((st = check_shapes(a, ta != 0, b, tb != 0, c, &sh))) return st;
  if ((st = check_sparse(a)) || (st = check_dense(b, al)) || (st = check_dense(c))) return st;
  const int nb = a->nonzeros / (128 * 128);
  if (nb > BSMM_STATIC_MAX_BLOCKS || a->cols / 128 > BSMM_STATIC_MAX_BLOCK_COLS)
    return BSMM_ERR_STATIC_LIMIT;  // before any CUDA call
  const int m = c->rows, n = c->cols, k = ta ? a->rows : a->cols;
  if (m == 0 || n == 0) return BSMM_OK;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  flags &= ~(BSMM_FLAG_TWIN_ROWS | BSMM_FLAG_TWIN_COLS | BSMM_FLAG_EVEN_ROWS |
             BSMM_FLAG_CREATE_METADATA);
  if (nb == 0) {  // every block row is empty: the output is all zeros
    if (zero_fill_needed(flags))
      BSMM_CUDA(cudaMemsetAsync(c->data, 0, size_t(m) * size_t(n) * 2, stream));
    return BSMM_OK;
  }
  StaticScratch* scratch = nullptr;
  if ((st = static_scratch_of_current_device(&scratch))) return st;
  // The array keeps the last call's contents.  The counter must start at zero (it also returns
  // to zero by itself with the last draw of a launch, unless that launch was torn down); the
  // metadata arrays are fully rewritten below for the entries this call reads.
  BSMM_CUDA(cudaMemsetAsync(scratch->tile_counter, 0, sizeof(scratch->tile_counter), stream));
  bsmm_block_matrix sp = *a;
  if (ta) {
    sp.offsets_t = scratch->offsets_t;
    sp.indices_t = scratch->indices_t;
    sp.block_offsets = scratch->block_offsets;
    if ((st = do_transpose(&sp, stream))) return st;  // kernel launches only
  }
  const bool f8 = dtype == BSMM_F8E4M3, bf16 = dtype != BSMM_F16, sp_t = ta != 0, dn_kmajor = tb != 0;
  const ElemType ot(bf16, f8);
  const int bk = int(ot.bk());
  const int sparse_ext = m, dense_ext = n;
  const int bn = dense_ext > 128 ? 256 : 128;
  CUtensorMap ma, mb, mc;
  // operand boxes as in run_dsd_like's 1-SM path (cluster size 1), output box as make_out_map
  if ((st = encode_map_static(&ma, ot, sp.data, 128, uint64_t(nb) * 128, 128,
                              sp_t ? uint32_t(bk) : 128u, 0, true)))
    return st;
  if ((st = dn_kmajor ? encode_map_static(&mb, ot, b->data, k, dense_ext, k, 128, 0, true)
                      : encode_map_static(&mb, ot, b->data, dense_ext, k, dense_ext, uint32_t(bk),
                                          0, true)))
    return st;
  if ((st = encode_map_static(&mc, ElemType(bf16), c->data, dense_ext, sparse_ext, dense_ext, 32, 64,
                              true)))
    return st;
  GemmParams p{};
  p.walk_offsets = static_cast<const int*>(sp_t ? sp.offsets_t : sp.offsets);
  p.walk_indices = static_cast<const short*>(sp_t ? sp.indices_t : sp.indices);
  p.walk_blocks = sp_t ? static_cast<const int*>(sp.block_offsets) : nullptr;
  p.c = c->data;
  p.c_rows = nullptr;
  apply_gate(p, Gate());
  if (sp_t) p.prefetch = 0;  // the walk is being rebuilt by the launch in front: nothing to peek at
  p.ldc = dense_ext;
  p.sparse_lim = sparse_ext; p.dense_lim = dense_ext;
  p.tiles_major = sparse_ext / 128;
  p.tiles_minor = (dense_ext + bn - 1) / bn;
  p.idesc = ptx::make_idesc(bf16 && !f8, sp_t, !dn_kmajor, 128, bn);
  p.bf16 = fmt_word(bf16, flags);
  p.bk = bk;
  p.f8 = f8;
  BSMM_SET_TRACE(p);
  int sms = 0;
  if ((st = current_device_sms(&sms))) return st;
  const long long units = 1ll * p.tiles_major * p.tiles_minor;
  if (units > sms) p.tile_counter = scratch->tile_counter;
  return bn == 128
             ? launch_variant<kOpDSD, 128, false, 1>(sp_t, !dn_kmajor, ma, mb, mc, p, units, sms, stream)
             : launch_variant<kOpDSD, 256, false, 1>(sp_t, !dn_kmajor, ma, mb, mc, p, units, sms, stream);
}
