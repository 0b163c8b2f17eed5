// Shared helpers for the absolutetrack_b200 CUDA sources (sm_100a only).
#pragma once
#include <atomic>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>

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

template <typename K>
inline void prefer_max_smem_carveout(K kernel) {
  cudaFuncSetAttribute((const void*)kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
}
#define AT_CARVEOUT_ONCE(kernel) AT_ONCE_PER_DEVICE(abt::prefer_max_smem_carveout(kernel))

__host__ __device__ inline int64_t round_up(int64_t x, int64_t m) { return (x + m - 1) / m * m; }

}  // namespace abt
