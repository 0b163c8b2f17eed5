// Micro-benchmark of tcgen05.mma issue/execute rate on B200 for the operand layouts the encoder
// could use.  Data is garbage (timing only).  Build: nvcc -gencode arch=compute_100a,code=sm_100a
// -O3 -Iabsolutetrack_b200/csrc -Iinclude scripts/umma_bench.cu -o scripts/umma_bench
#include <cstdio>
#include <cuda_runtime.h>
#include "tc_ptx.cuh"
using namespace abt::ptx;

__device__ __forceinline__ uint64_t desc_any(uint32_t addr, uint32_t lbo, uint32_t sbo, uint32_t layout) {
  uint64_t d = 0;
  d |= (uint64_t)((addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)layout << 61;
  return d;
}

// mode 0: no-swizzle planar A (LBO = plane, SBO = 128), aligned start
// mode 1: same, start shifted by 16 B per tap (unaligned core matrices)
// mode 2: SW128 K-major A and B (SBO = 1024), K advance 32 B
// mode 3: no-swizzle A with 128B-aligned shifts only (shift by 8 pixels)
__global__ void __launch_bounds__(128, 1) bench(int mode, int n, int iters, int nacc, int style, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint32_t tslot;
  __shared__ __align__(8) uint64_t bar;
  const uint32_t sb = smem_u32(smem);
  if (threadIdx.x < 32) { tmem_alloc(smem_u32(&tslot), 512); tmem_relinquish(); }
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); fence_barrier_init(); }
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  fence_proxy_async();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  uint32_t tm = tslot;
  if (style == 1 && threadIdx.x < 32) {
    // converged warp, warp-uniform operands, elect.sync around the issue only
    tm = __shfl_sync(0xffffffffu, tm, 0);
    const uint32_t idesc = umma_idesc(128, n, 1);
    const uint32_t plane = 232 * 16;
    const uint32_t a0 = sb, w0 = sb + 96 * 1024;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      const int tap = i % 9, j = (i / 9) & 1;
      uint64_t ad, bd;
      if (mode == 2) {
        ad = desc_any(a0 + (tap & 3) * 16384 + j * 32, 16, 1024, 2);
        bd = desc_any(w0 + (tap & 1) * 32768 + j * 32, 16, 1024, 2);
      } else {
        const uint32_t shift = mode == 1 ? tap * 16 + (tap / 3) * 49 * 16 : (mode == 3 ? tap * 128 : 0);
        ad = desc_any(a0 + shift + 2 * j * plane, plane, 128, 0);
        bd = desc_any(w0 + tap * 2048 + 2 * j * n * 16, n * 16, 128, 0);
      }
      if (elect_one()) umma_f16(tm + (i % nacc) * n, ad, bd, idesc, 1);
      __syncwarp();
    }
    if (elect_one()) umma_commit(smem_u32(&bar));
    __syncwarp();
    mbar_wait(smem_u32(&bar), 0);
    long long t1 = clock64();
    if (blockIdx.x == 0 && threadIdx.x == 0) out[0] = t1 - t0;
  }
  if (style >= 2 && threadIdx.x < 32) {
    // constant descriptors, unrolled: pure hardware rate.  style 3: M = 64.  style 4: two CTAs' worth? no: K-major SW128 consts
    tm = __shfl_sync(0xffffffffu, tm, 0);
    const uint32_t idesc = umma_idesc(style == 3 ? 64 : 128, n, 1);
    const uint32_t plane = 232 * 16;
    const uint64_t ad = mode == 2 ? desc_any(sb, 16, 1024, 2) : desc_any(sb, plane, 128, 0);
    const uint64_t bd = mode == 2 ? desc_any(sb + 96 * 1024, 16, 1024, 2) : desc_any(sb + 96 * 1024, n * 16, 128, 0);
    long long t0 = clock64();
    if (elect_one()) {
      for (int i = 0; i < iters; i += 8) {
#pragma unroll
        for (int u = 0; u < 8; ++u) umma_f16(tm + (u & (nacc - 1)) * n, ad, bd, idesc, 1);
      }
      umma_commit(smem_u32(&bar));
    }
    __syncwarp();
    mbar_wait(smem_u32(&bar), 0);
    long long t1 = clock64();
    if (blockIdx.x == 0 && threadIdx.x == 0) out[0] = t1 - t0;
  }
  if (style == 0 && threadIdx.x == 0) {
    const uint32_t idesc = umma_idesc(128, n, 1);
    const uint32_t plane = 232 * 16;          // like layer 1
    const uint32_t a0 = sb, w0 = sb + 96 * 1024;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      const int tap = i % 9, j = (i / 9) & 1;
      uint64_t ad, bd;
      if (mode == 2) {
        ad = desc_any(a0 + (tap & 3) * 16384 + j * 32, 16, 1024, 2);
        bd = desc_any(w0 + (tap & 1) * 32768 + j * 32, 16, 1024, 2);
      } else {
        const uint32_t shift = mode == 1 ? tap * 16 + (tap / 3) * 49 * 16 : (mode == 3 ? tap * 128 : 0);
        ad = desc_any(a0 + shift + 2 * j * plane, plane, 128, 0);
        bd = desc_any(w0 + tap * 2048 + 2 * j * n * 16, n * 16, 128, 0);
      }
      umma_f16(tm + (i % nacc) * n, ad, bd, idesc, 1);
    }
    umma_commit(smem_u32(&bar));
    mbar_wait(smem_u32(&bar), 0);
    long long t1 = clock64();
    if (blockIdx.x == 0) out[0] = t1 - t0;
  }
  tc_fence_before(); __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc(tm, 512);
}

int main() {
  long long* d; cudaMalloc(&d, 8);
  cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const int iters = 1800;
  for (int mode : {0})
    for (int n : {32, 64, 128, 256})
      for (int nacc : {1, 2, 4})
      for (int style : {2, 3}) {
        if (nacc * n > 512) continue;
        bench<<<148, 128, 200 * 1024>>>(mode, n, iters, nacc, style, d);
        cudaError_t e = cudaDeviceSynchronize();
        long long h = 0; cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
        printf("nacc %d mode %d n %3d style %d: %8.1f cycles/MMA (floor %d)  %s\n", nacc, mode, n, style, (double)h / iters, n / 2,
               e == cudaSuccess ? "" : cudaGetErrorString(e));
      }
  return 0;
}
