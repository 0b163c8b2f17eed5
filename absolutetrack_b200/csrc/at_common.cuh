// Shared helpers for the absolutetrack_b200 CUDA sources (sm_100a only).
#pragma once
#include <atomic>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "absolutetrack_b200.h"

namespace abt {

void set_error(const char* fmt, ...);
void count_launch();  // every kernel launch of this library is counted (at_launch_count)

inline int check_launch(const char* what) {
  count_launch();
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: %s", what, cudaGetErrorString(e));
    return AT_ERR_CUDA;
  }
  return AT_OK;
}

#define AT_REQUIRE(cond, ...)            \
  do {                                   \
    if (!(cond)) {                       \
      abt::set_error(__VA_ARGS__);        \
      return AT_ERR_INVALID_ARGUMENT;    \
    }                                    \
  } while (0)

constexpr int kNumSMs = 148;  // B200: 2 dies x 74 SMs

// Kernels that must co-reside with a long-lived CTA of ANOTHER kernel (the encoder's conv launches
// and the source-box pull of the next step) all ask for the SAME shared-memory carve-out (the
// maximum).  The driver otherwise picks the smallest configuration that fits each launch, and an SM
// can only switch configuration when it is empty: with a pull CTA resident, the first conv launch
// that wants a different carve-out stalls until that CTA exits (measured: the encoder blocked for
// the whole pull; scripts/overlap_probe.cu).  Kernels outside that window keep the default (more L1).
// Function attributes belong to a device's context: "once" means once per device, and callers may
// come from several threads.  `mask` holds one bit per device ordinal.
inline uint64_t this_device_bit() {
  int dev = 0;
  cudaGetDevice(&dev);
  return 1ull << (dev & 63);
}
// the statements are idempotent attribute calls: two racing first callers may both run them, nobody
// launches before they have run
#define AT_ONCE_PER_DEVICE(...)                                    \
  do {                                                             \
    static std::atomic<uint64_t> at_once_mask_{0};                 \
    const uint64_t at_bit_ = abt::this_device_bit();                \
    if ((at_once_mask_.load(std::memory_order_acquire) & at_bit_) == 0) { \
      __VA_ARGS__;                                                 \
      at_once_mask_.fetch_or(at_bit_, std::memory_order_release);  \
    }                                                              \
  } while (0)

// Programmatic dependent launch for the chains of short kernels (the tail: 18 launches of 10-30 us): launched with the
// stream-serialization attribute a kernel becomes resident while its predecessor drains and runs up to its
// pdl_prologue() - which every such kernel calls before it touches global memory - so the launch latency and the
// CTA ramp-up overlap the previous kernel's tail.  A predecessor without the trigger simply completes first; a kernel
// launched without the attribute passes through pdl_prologue() at once.  AT_TAIL_PDL=0: plain launches (A/B).
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args&&... args) {
  static const bool on = getenv("AT_TAIL_PDL") == nullptr || atoi(getenv("AT_TAIL_PDL")) != 0;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = on ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_prologue() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
}
#endif

template <typename K>
inline void prefer_max_smem_carveout(K kernel) {
  cudaFuncSetAttribute((const void*)kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
}
#define AT_CARVEOUT_ONCE(kernel) AT_ONCE_PER_DEVICE(abt::prefer_max_smem_carveout(kernel))

__host__ __device__ inline int64_t round_up(int64_t x, int64_t m) { return (x + m - 1) / m * m; }

}  // namespace abt
