// Shared helpers for the absolutetrack_b200 CUDA sources (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>

#include "absolutetrack_b200.h"

namespace at {

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
      at::set_error(__VA_ARGS__);        \
      return AT_ERR_INVALID_ARGUMENT;    \
    }                                    \
  } while (0)

constexpr int kNumSMs = 148;  // B200: 2 dies x 74 SMs

__host__ __device__ inline int64_t round_up(int64_t x, int64_t m) { return (x + m - 1) / m * m; }

}  // namespace at
