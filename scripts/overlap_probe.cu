// What does a tiny co-resident side kernel cost the persistent tcgen05 encoder?  Side kernels of
// different kinds (sleep / ALU spin / PCIe reads via LDG / PCIe reads via cp.async / HBM reads via
// cp.async), all 1 warp x 32 registers per CTA, each running for a fixed wall time.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -shared -Xcompiler -fPIC -o scripts/liboverlap_probe.so scripts/overlap_probe.cu
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }

__global__ void __maxnreg__(32) k_sleep(unsigned long long dur) {
  const unsigned long long t0 = gtime();
  while (gtime() - t0 < dur) __nanosleep(2000);
}
__global__ void __maxnreg__(32) k_alu(unsigned long long dur, unsigned* out) {
  const unsigned long long t0 = gtime();
  unsigned x = threadIdx.x;
  while (gtime() - t0 < dur) {
#pragma unroll
    for (int i = 0; i < 64; ++i) x = x * 1664525u + 1013904223u;
  }
  if (x == 12345u) *out = x;
}
// streaming 16-byte reads from src (host pinned or device), results XOR-folded
__global__ void __maxnreg__(32) k_ldg(unsigned long long dur, const uint4* __restrict__ src, size_t nwords, unsigned* out) {
  const unsigned long long t0 = gtime();
  size_t i = ((size_t)blockIdx.x * 32 + threadIdx.x);
  const size_t stride = (size_t)gridDim.x * 32;
  unsigned acc = 0;
  while (gtime() - t0 < dur) {
    uint4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) { v[u] = __ldg(src + i); i += stride; if (i >= nwords) i -= nwords; }
#pragma unroll
    for (int u = 0; u < 4; ++u) acc ^= v[u].x ^ v[u].w;
  }
  if (acc == 12345u) *out = acc;
}
__global__ void __maxnreg__(32) k_cpasync(unsigned long long dur, const uint4* __restrict__ src, size_t nwords, unsigned* out, int ca, int scatter) {
  __shared__ __align__(16) uint4 ring[2 * 4 * 32];
  const unsigned long long t0 = gtime();
  size_t i = ((size_t)blockIdx.x * 32 + threadIdx.x);
  // scatter: every warp request lands in another 4 KB page (no page reuse between requests)
  const size_t stride = scatter ? (size_t)(gridDim.x * 32) * 131 + 32 * 8 : (size_t)gridDim.x * 32;
  const uint32_t rs = (uint32_t)__cvta_generic_to_shared(ring) + threadIdx.x * 16;
  unsigned acc = 0; int st = 0;
  while (gtime() - t0 < dur) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (ca) asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(rs + (st & 1) * 2048 + u * 512), "l"(src + i) : "memory");
      else asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(rs + (st & 1) * 2048 + u * 512), "l"(src + i) : "memory");
      i += stride; if (i >= nwords) i -= nwords;
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    acc ^= ring[((st + 1) & 1) * 128 + threadIdx.x].x;
    ++st;
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  if (acc == 12345u) *out = acc;
}
extern "C" int probe_launch(int kind, int ctas, unsigned long long dur_ns, const void* src, size_t bytes, void* scratch, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  const int carve = (kind >> 8) - 1;  // (carveout percent + 1) << 8, 0 = leave the default
  kind &= 255;
  const void* fns[] = {(const void*)k_sleep, (const void*)k_alu, (const void*)k_ldg, (const void*)k_cpasync, (const void*)k_cpasync};
  cudaFuncSetAttribute(fns[kind], cudaFuncAttributePreferredSharedMemoryCarveout, carve);
  switch (kind) {
    case 0: k_sleep<<<ctas, 32, 0, s>>>(dur_ns); break;
    case 1: k_alu<<<ctas, 32, 0, s>>>(dur_ns, (unsigned*)scratch); break;
    case 2: k_ldg<<<ctas, 32, 0, s>>>(dur_ns, (const uint4*)src, bytes / 16, (unsigned*)scratch); break;
    case 3: k_cpasync<<<ctas, 32, 0, s>>>(dur_ns, (const uint4*)src, bytes / 16, (unsigned*)scratch, 1, 0); break;
    case 4: k_cpasync<<<ctas, 32, 0, s>>>(dur_ns, (const uint4*)src, bytes / 16, (unsigned*)scratch, 1, 1); break;
  }
  return (int)cudaGetLastError();
}
