// Microbenchmark (developer tool): L2 -> shared-memory bulk-copy throughput of persistent CTAs
// as a function of how many CTAs read the SAME data in lockstep.  Motivates the tile schedule /
// multicast design of the block-sparse kernels (DESIGN.md "L2 -> SM feed").
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tma_bw tools/tma_bw.cu && ./tma_bw
//
// Each CTA runs a 4-stage ring of 48 KiB stages (the GEMM kernel's stage size) filled by
// cp.async.bulk (the TMA engine) and drained by a consumer thread that only flips barriers.
// CTAs are grouped: the g CTAs of a group walk the same region; groups walk disjoint regions.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(b)), "r"(c) : "memory"); }
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(b)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t* b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(b)) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
  uint32_t ok = 0;
  while (!ok) {
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.b32 %0, 1, 0, p; }"
                 : "=r"(ok) : "r"(s32(b)), "r"(parity) : "memory");
  }
}
__device__ __forceinline__ void bulk_load(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(s32(dst)), "l"(src), "r"(bytes), "r"(s32(bar)) : "memory");
}

#ifndef STAGES
#define STAGES 4
#endif
#ifndef STAGE_KB
#define STAGE_KB 48
#endif
constexpr int kStages = STAGES;
constexpr int kStageBytes = STAGE_KB * 1024;
constexpr int kChunk = 16 * 1024;

__global__ void __launch_bounds__(64, 1)
feed_kernel(const uint8_t* base, size_t region_bytes, int group, int iters, int skew) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t full[kStages], empty[kStages];
  if (threadIdx.x == 0) {
    for (int i = 0; i < kStages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const uint8_t* region = base + size_t(blockIdx.x / group) * region_bytes;
  // skew: members of a group start at different offsets of the shared region (no lockstep)
  size_t off = skew ? (size_t(blockIdx.x % group) * (region_bytes / group)) % region_bytes : 0;
  off -= off % kStageBytes;
  if (threadIdx.x == 0) {
    int st = 0; uint32_t ph = 0;
    for (int i = 0; i < iters; ++i) {
      mbar_wait(&empty[st], ph ^ 1);
      mbar_expect(&full[st], kStageBytes);
      for (int c = 0; c < kStageBytes / kChunk; ++c)
        bulk_load(smem + st * kStageBytes + c * kChunk, region + off + c * kChunk, kChunk, &full[st]);
      off += kStageBytes;
      if (off + kStageBytes > region_bytes) off = 0;
      if (++st == kStages) { st = 0; ph ^= 1; }
    }
  } else if (threadIdx.x == 32) {
    int st = 0; uint32_t ph = 0;
    for (int i = 0; i < iters; ++i) {
      mbar_wait(&full[st], ph);
      mbar_arrive(&empty[st]);
      if (++st == kStages) { st = 0; ph ^= 1; }
    }
  }
}

int main(int argc, char** argv) {
  int sms = 0;
  const int grid_arg = argc > 1 ? atoi(argv[1]) : 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  const size_t total = size_t(96) << 20;  // footprint kept below L2
  uint8_t* buf;
  CK(cudaMalloc(&buf, total));
  CK(cudaMemset(buf, 1, total));
  CK(cudaFuncSetAttribute(feed_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kStages * kStageBytes));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  const int grid = grid_arg > 0 ? grid_arg : sms;
  printf("SMs %d, grid %d, stage %d KiB x %d\n", sms, grid, kStageBytes / 1024, kStages);
  printf("%6s %5s %10s %12s %10s\n", "group", "skew", "region", "TB/s (SM)", "B/clk/SM");
  int clk_khz = 0;
  CK(cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0));
  const int groups[] = {1, 2, 4};
  for (int skew = 0; skew < 2; ++skew)
    for (int g : groups) {
      if (g > grid) continue;
      if (skew && g == 1) continue;
      const int ngroups = (grid + g - 1) / g;
      size_t region = total / ngroups;
      region -= region % kStageBytes;
      if (region > (size_t(8) << 20)) region = size_t(8) << 20;  // keep regions comparable
      const int iters = 2000;
      for (int rep = 0; rep < 2; ++rep) {
        CK(cudaEventRecord(e0));
        feed_kernel<<<grid, 64, kStages * kStageBytes>>>(buf, region, g, iters, skew);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
      }
      CK(cudaGetLastError());
      float ms = 0;
      CK(cudaEventElapsedTime(&ms, e0, e1));
      const double bytes = double(grid) * iters * kStageBytes;
      printf("%6d %5d %8zuKB %12.2f %10.1f\n", g, skew, region >> 10, bytes / ms / 1e9,
             bytes / grid / (ms * 1e-3 * clk_khz * 1e3));
    }
  return 0;
}
