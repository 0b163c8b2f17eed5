// Micro-benchmark 2: does the per-tile structure (commit, fences, barrier waits) around a short
// burst of tcgen05.mma cost pipeline bubbles?  Timing only, operands are zeros.
#include <cstdio>
#include <cuda_runtime.h>
#include "tc_ptx.cuh"
using namespace abt::ptx;

__device__ __forceinline__ uint64_t desc_any(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= (uint64_t)((addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  return d;
}

// variant bits: 1 = commit to bar A after each burst, 2 = second commit (bar B), 4 = fence::after + elect per burst,
// 8 = wait on bar A of the burst two iterations back (true pipeline dependence), 16 = first MMA of a burst has accumulate=0
template <int BURST>
__global__ void __launch_bounds__(128, 1) bench(int n, int tiles, int variant, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint32_t tslot;
  __shared__ __align__(8) uint64_t bars[8];
  const uint32_t sb = smem_u32(smem);
  if (threadIdx.x < 32) { tmem_alloc(smem_u32(&tslot), 512); tmem_relinquish(); }
  if (threadIdx.x == 0) { for (int i = 0; i < 8; ++i) mbar_init(smem_u32(&bars[i]), 1); fence_barrier_init(); }
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  fence_proxy_async();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  if (threadIdx.x < 32) {
    const uint32_t tm = __shfl_sync(0xffffffffu, tslot, 0);
    const uint32_t idesc = umma_idesc(128, n, 1);
    const uint32_t plane = 232 * 16;
    const uint64_t ad0 = desc_any(sb, plane, 128);
    const uint64_t bd0 = desc_any(sb + 96 * 1024, n * 16, 128);
    const uint32_t barA = smem_u32(&bars[0]), barB = smem_u32(&bars[4]), barE = smem_u32(&bars[7]);
    uint32_t phase[4] = {0, 0, 0, 0};
    long long t0 = clock64();
    for (int it = 0; it < tiles; ++it) {
      const int slot = it & 3;
      if ((variant & 8) && it >= 4) {  // wait for the burst 4 iterations back (same slot)
        mbar_wait(barA + 8 * slot, phase[slot]);
        phase[slot] ^= 1u;
      }
      if (variant & 4) tc_fence_after();
      if (elect_one()) {
        const uint32_t d = tm + (it & 1) * n;
#pragma unroll
        for (int u = 0; u < BURST; ++u)
          umma_f16(d, ad0 + (u % 9) + ((u & 1) ? 2 * (plane >> 4) : 0), bd0 + u * (n >> 2), idesc, (variant & 16) ? (u != 0) : 1u);
        if (variant & 1) umma_commit(barA + 8 * slot);
        if (variant & 2) umma_commit(barB + 8 * (it & 1));
      }
      __syncwarp();
    }
    if (elect_one()) umma_commit(barE);
    __syncwarp();
    mbar_wait(barE, 0);
    long long t1 = clock64();
    if (blockIdx.x == 0 && threadIdx.x == 0) out[0] = t1 - t0;
  }
  tc_fence_before(); __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc(tslot, 512);
}

int main() {
  long long* d; cudaMalloc(&d, 8);
  cudaFuncSetAttribute(bench<18>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaFuncSetAttribute(bench<36>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaFuncSetAttribute(bench<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const int tiles = 200;
  for (int n : {32, 64, 128})
    for (int variant : {0, 1, 3, 4, 7, 15, 31, 16}) {
      for (int burst : {4, 18, 36}) {
        if (burst == 18) bench<18><<<148, 128, 200 * 1024>>>(n, tiles, variant, d);
        else if (burst == 36) bench<36><<<148, 128, 200 * 1024>>>(n, tiles, variant, d);
        else bench<4><<<148, 128, 200 * 1024>>>(n, tiles, variant, d);
        cudaError_t e = cudaDeviceSynchronize();
        long long h = 0; cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
        printf("n %3d variant %2d burst %2d: %8.1f cycles/burst  %6.1f cycles/MMA  %s\n", n, variant, burst, (double)h / tiles,
               (double)h / tiles / burst, e == cudaSuccess ? "" : cudaGetErrorString(e));
      }
    }
  return 0;
}
