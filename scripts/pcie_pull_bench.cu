// Microbenchmark: SM-initiated reads of many small 2-D boxes from PINNED HOST memory over PCIe.
// Which load instruction / span alignment / CTA shape gives the best useful bandwidth?
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/pcie_pull_bench scripts/pcie_pull_bench.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
struct Box { uint32_t base; int bw, bh; };
constexpr uint32_t kRow = 2544;

template <int KIND> __device__ __forceinline__ uint4 load16(const uint8_t* p) {
  uint4 v;
  if (KIND == 0) v = __ldg(reinterpret_cast<const uint4*>(p));
  else if (KIND == 1) asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  else asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}

template <int KIND, int UNROLL>
__global__ void pull_ldg(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, const Box* __restrict__ boxes, int n, uint32_t amask, unsigned long long* words) {
  uint32_t mine = 0;
  for (int b = blockIdx.x; b < n; b += gridDim.x) {
    const Box bx = boxes[b];
    const int wpr = (bx.bw + 2 * (int)amask + 15) >> 4;
    const int total = wpr * bx.bh;
    for (int i0 = threadIdx.x; i0 < total; i0 += blockDim.x * UNROLL) {
      uint4 v[UNROLL]; uint32_t a[UNROLL];
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) {
        const int i = i0 + u * blockDim.x;
        a[u] = 0xFFFFFFFFu;
        if (i < total) {
          const int r = i / wpr, c = i - r * wpr;
          const uint32_t row0 = bx.base + r * kRow;
          const uint32_t addr = (row0 & ~amask) + (c << 4);
          if (addr < ((row0 + bx.bw + amask) & ~amask)) { a[u] = addr; v[u] = load16<KIND>(src + addr); }
        }
      }
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) if (a[u] != 0xFFFFFFFFu) { *reinterpret_cast<uint4*>(dst + a[u]) = v[u]; ++mine; }
    }
  }
  atomicAdd(words, (unsigned long long)mine);
}

// one warp per CTA, cp.async ring (registers hold no data)
template <int CA>
__global__ void pull_cpasync(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, const Box* __restrict__ boxes, int n, uint32_t amask, unsigned long long* words) {
  __shared__ __align__(16) uint8_t ring[2 * 4 * 512];
  const int lane = threadIdx.x;
  const uint32_t ring_s = (uint32_t)__cvta_generic_to_shared(ring) + lane * 16;
  uint32_t mine = 0;
  for (int b = blockIdx.x; b < n; b += gridDim.x) {
    const Box bx = boxes[b];
    const int wpr = (bx.bw + 2 * (int)amask + 15) >> 4;
    const int total = wpr * bx.bh;
    const int nst = (total + 127) / 128;
    auto off = [&](int i) -> uint32_t {
      if (i >= total) return 0xFFFFFFFFu;
      const int r = i / wpr, c = i - r * wpr;
      const uint32_t row0 = bx.base + r * kRow;
      const uint32_t addr = (row0 & ~amask) + (c << 4);
      return addr < ((row0 + bx.bw + amask) & ~amask) ? addr : 0xFFFFFFFFu;
    };
    for (int st = 0; st < nst + 1; ++st) {
      if (st < nst) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const uint32_t a = off(st * 128 + u * 32 + lane);
          if (a != 0xFFFFFFFFu) {
            if (CA) asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(ring_s + (st & 1) * 2048 + u * 512), "l"(src + a) : "memory");
            else asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(ring_s + (st & 1) * 2048 + u * 512), "l"(src + a) : "memory");
          }
        }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
      if (st >= 1) {
        asm volatile("cp.async.wait_group 1;" ::: "memory");
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const uint32_t a = off((st - 1) * 128 + u * 32 + lane);
          if (a != 0xFFFFFFFFu) { *reinterpret_cast<uint4*>(dst + a) = *reinterpret_cast<const uint4*>(ring + lane * 16 + ((st - 1) & 1) * 2048 + u * 512); ++mine; }
        }
      }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
  }
  atomicAdd(words, (unsigned long long)mine);
}

int main() {
  const int S = 512, n = 2048;
  const size_t bytes = (size_t)S * 480 * kRow;
  uint8_t *h, *d; Box* dbox; unsigned long long* dw;
  cudaHostAlloc(&h, bytes, cudaHostAllocDefault);
  for (size_t i = 0; i < bytes; i += 4096) h[i] = (uint8_t)i;
  cudaMalloc(&d, bytes); cudaMalloc(&dbox, n * sizeof(Box)); cudaMalloc(&dw, 8);
  std::vector<Box> hb(n);
  srand(1);
  size_t useful = 0;
  for (int i = 0; i < n; ++i) {
    const int bw = 150 + rand() % 60, bh = 180 + rand() % 50;
    const int x0 = rand() % (636 - bw), y0 = rand() % (480 - bh);
    const int frame = i / 4, view = 1 + (i % 2);
    hb[i] = Box{(uint32_t)((size_t)frame * 480 * kRow + view * 636 + (size_t)y0 * kRow + x0), bw, bh};
    useful += (size_t)bw * bh;
  }
  cudaMemcpy(dbox, hb.data(), n * sizeof(Box), cudaMemcpyHostToDevice);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  auto run = [&](const char* name, auto launch) {
    for (uint32_t al : {16u, 32u, 64u, 128u}) {
      launch(al - 1); cudaDeviceSynchronize();
      cudaMemset(dw, 0, 8);
      cudaEventRecord(e0);
      for (int r = 0; r < 3; ++r) launch(al - 1);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 3;
      unsigned long long w; cudaMemcpy(&w, dw, 8, cudaMemcpyDeviceToHost);
      const cudaError_t err = cudaGetLastError();
      printf("%-34s align %3u: %6.3f ms  useful %5.1f GB/s  wire %5.1f GB/s (%5.1f MB) %s\n", name, al, ms, useful / ms / 1e6, w / 3 * 16 / ms / 1e6, w / 3 * 16 / 1e6, err ? cudaGetErrorString(err) : "");
    }
  };
  run("ldg(L1) 148x128 u4", [&](uint32_t m) { pull_ldg<0, 4><<<148, 128>>>(h, d, dbox, n, m, dw); });
  run("ldg(L1) 2048x256 u4", [&](uint32_t m) { pull_ldg<0, 4><<<2048, 256>>>(h, d, dbox, n, m, dw); });
  run("ldg(L1) 148x32 u4", [&](uint32_t m) { pull_ldg<0, 4><<<148, 32>>>(h, d, dbox, n, m, dw); });
  run("ldg(L1) 148x32 u8", [&](uint32_t m) { pull_ldg<0, 8><<<148, 32>>>(h, d, dbox, n, m, dw); });
  run("ldg(L1) 37x32 u4", [&](uint32_t m) { pull_ldg<0, 4><<<37, 32>>>(h, d, dbox, n, m, dw); });
  run("ld.nc.no_allocate 148x128 u4", [&](uint32_t m) { pull_ldg<1, 4><<<148, 128>>>(h, d, dbox, n, m, dw); });
  run("ld.nc.no_allocate 148x32 u4", [&](uint32_t m) { pull_ldg<1, 4><<<148, 32>>>(h, d, dbox, n, m, dw); });
  run("ld.cg 148x128 u4", [&](uint32_t m) { pull_ldg<2, 4><<<148, 128>>>(h, d, dbox, n, m, dw); });
  run("ld.cg 148x32 u4", [&](uint32_t m) { pull_ldg<2, 4><<<148, 32>>>(h, d, dbox, n, m, dw); });
  run("cp.async.cg 148x32", [&](uint32_t m) { pull_cpasync<0><<<148, 32>>>(h, d, dbox, n, m, dw); });
  run("cp.async.ca 148x32", [&](uint32_t m) { pull_cpasync<1><<<148, 32>>>(h, d, dbox, n, m, dw); });
  run("cp.async.ca 592x32", [&](uint32_t m) { pull_cpasync<1><<<592, 32>>>(h, d, dbox, n, m, dw); });
  // reference: plain DMA of the whole buffer
  cudaEventRecord(e0); cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice); cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  printf("cudaMemcpy H2D %zu MB: %.3f ms  %.1f GB/s\n", bytes >> 20, ms, bytes / ms / 1e6);
  return 0;
}
