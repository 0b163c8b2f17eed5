// tcgen05.mma rate against the ALIGNMENT of the no-swizzle A operand: the shift-GEMM convs start their A descriptors
// at tap shifts of +-1 pixel (16 B) and +-Wp pixels, so most core matrices (8 rows x 16 B = 128 B) straddle two
// 128-byte shared-memory lines.  Timing only (garbage data).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -Iabsolutetrack_b200/csrc -Iinclude scripts/umma_bench3.cu -o scripts/umma_bench3
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

// offs[8]: byte offsets of the A start of the 8 MMAs of the unrolled group
__global__ void __launch_bounds__(128, 1) bench(int n, int iters, const int* offs, int b_rot, long long* out) {
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
  if (threadIdx.x < 32) {
    tm = __shfl_sync(0xffffffffu, tm, 0);
    const uint32_t idesc = umma_idesc(128, n, 1);
    const uint32_t plane = 320 * 16;
    uint64_t ad[8], bd[8];
    for (int u = 0; u < 8; ++u) {
      ad[u] = desc_any(sb + 1024 + offs[u], plane, 128);
      bd[u] = desc_any(sb + 96 * 1024 + (b_rot ? u * 4096 : 0), n * 16, 128);
    }
    long long t0 = clock64();
    if (elect_one()) {
      for (int i = 0; i < iters; i += 8) {
#pragma unroll
        for (int u = 0; u < 8; ++u) umma_f16(tm + (u & 1) * n, ad[u], bd[u], idesc, 1);
      }
      umma_commit(smem_u32(&bar));
    }
    __syncwarp();
    mbar_wait(smem_u32(&bar), 0);
    long long t1 = clock64();
    if (blockIdx.x == 0 && threadIdx.x == 0) out[0] = t1 - t0;
  }
  tc_fence_before(); __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc(tm, 512);
}

int main() {
  long long* d; cudaMalloc(&d, 8);
  int* doffs; cudaMalloc(&doffs, 32);
  cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const int iters = 8000;
  struct Case { const char* name; int o[8]; } cases[] = {
    {"aligned, one address           ", {0, 0, 0, 0, 0, 0, 0, 0}},
    {"aligned, 8 addresses (x128 B)  ", {0, 128, 256, 384, 512, 640, 768, 896}},
    {"+16 B, one address             ", {16, 16, 16, 16, 16, 16, 16, 16}},
    {"+64 B, one address             ", {64, 64, 64, 64, 64, 64, 64, 64}},
    {"+112 B, one address            ", {112, 112, 112, 112, 112, 112, 112, 112}},
    {"3x3 taps Wp=49 (layer 1 form)  ", {0, 16, 32, 784, 800, 816, 1568, 1584}},
    {"3x3 taps Wp=25 (S2D form)      ", {0, 16, 32, 400, 416, 432, 800, 816}},
    {"3x3 taps Wp=24 (rows aligned)  ", {0, 16, 32, 384, 400, 416, 768, 784}},
    {"dy taps only, Wp=32 (aligned)  ", {0, 512, 1024, 0, 512, 1024, 0, 512}},
  };
  for (int n : {32, 64, 128, 256})
    for (auto& c : cases)
      for (int b_rot : {0, 1}) {
        cudaMemcpy(doffs, c.o, 32, cudaMemcpyHostToDevice);
        bench<<<148, 128, 200 * 1024>>>(n, iters, doffs, b_rot, d);
        long long h = 0;
        cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
        cudaError_t e = cudaGetLastError();
        printf("n %3d  A %s B %s: %7.1f cycles/MMA (math floor %d, 128 B/clk operand model %d)  %s\n", n, c.name, b_rot ? "8 blocks" : "1 block ",
               (double)h / iters, n / 2, (4096 + n * 32) / 128, e == cudaSuccess ? "" : cudaGetErrorString(e));
      }
  return 0;
}
