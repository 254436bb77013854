// Microbenchmark (developer tool): issue rate of tcgen05.mma 128 x N x 16 (bf16, cta_group::1) on
// operands resident in shared memory — the tensor-pipe ceiling of the block-sparse mainloop with
// the memory system taken out.  Prints cycles per MMA and the implied chip TFLOP/s.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I sputnik_b200/csrc -o tools/umma_rate tools/umma_rate.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#include "ptx.cuh"

using namespace bsmm;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

// kmaj: 1 = both operands K-major, 0 = both MN-major (descriptor only; data is garbage)
__global__ void __launch_bounds__(128, 1) rate_kernel(int n, int iters, int kmaj, long long* cycles) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_ptr;
  for (int i = threadIdx.x; i < 48 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); ptx::fence_mbar_init(); }
  ptx::fence_proxy_async_smem();
  if (threadIdx.x < 32) ptx::tmem_alloc(&tmem_ptr, 512);
  ptx::tc_fence_before_sync();
  __syncthreads();
  ptx::tc_fence_after_sync();
  const uint32_t tmem = tmem_ptr;
  if (threadIdx.x == 0) {
    const uint32_t sa = ptx::smem_u32(smem), sb = sa + 16384;
    const uint64_t da = ptx::make_smem_desc(sa, kmaj ? 16 : 8192, 1024);
    const uint64_t db = ptx::make_smem_desc(sb, kmaj ? 16 : 8192, 1024);
    const uint32_t idesc = ptx::make_idesc(true, !kmaj, !kmaj, 128, n);
    const uint32_t adv = kmaj ? 2 : 128;
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
      for (int k = 0; k < 4; ++k)
        ptx::umma_f16(tmem + uint32_t((i & 1) * 256), da + uint64_t(k * adv), db + uint64_t(k * adv), idesc, 1u);
    }
    ptx::umma_commit(&bar);
    ptx::mbar_wait(&bar, 0);
    const long long t1 = clock64();
    cycles[blockIdx.x] = t1 - t0;
  }
  __syncthreads();
  if (threadIdx.x < 32) ptx::tmem_dealloc(tmem, 512);
}

int main() {
  int sms = 0, khz = 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  CK(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0));
  long long* d;
  CK(cudaMalloc(&d, sizeof(long long) * sms));
  CK(cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  printf("%5s %5s %6s %12s %12s %14s\n", "grid", "N", "kmaj", "cyc/MMA", "wall us", "chip TFLOP/s");
  const int iters = 20000;
  for (int grid : {1, sms})
    for (int n : {128, 192, 208, 224, 256})
      for (int kmaj : {1, 0}) {
        cudaEvent_t e0, e1;
        CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
        rate_kernel<<<grid, 128, 64 * 1024>>>(n, 100, kmaj, d);
        CK(cudaEventRecord(e0));
        rate_kernel<<<grid, 128, 64 * 1024>>>(n, iters, kmaj, d);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        CK(cudaGetLastError());
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        long long c;
        CK(cudaMemcpy(&c, d, sizeof(c), cudaMemcpyDeviceToHost));
        const double flops = 2.0 * 128 * n * 16 * 4.0 * iters * grid;
        printf("%5d %5d %6d %12.1f %12.1f %14.1f\n", grid, n, kmaj, double(c) / (4.0 * iters), ms * 1e3,
               flops / (ms * 1e-3) / 1e12);
      }
  return 0;
}
