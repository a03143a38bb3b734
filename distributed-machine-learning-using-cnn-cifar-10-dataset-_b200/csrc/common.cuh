// Shared helpers for the b200cnn CUDA sources (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/b200cnn.h"

namespace b200 {

void set_error(const char* fmt, ...);

#define B200_CUDA_TRY(expr)                                                              \
  do {                                                                                   \
    cudaError_t e__ = (expr);                                                            \
    if (e__ != cudaSuccess) {                                                            \
      ::b200::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e__)); \
      return B200_E_CUDA;                                                                \
    }                                                                                    \
  } while (0)

#define B200_LAUNCH_CHECK() B200_CUDA_TRY(cudaGetLastError())

#define B200_REQUIRE(cond, ...)                 \
  do {                                          \
    if (!(cond)) {                              \
      ::b200::set_error(__VA_ARGS__);           \
      return B200_E_INVALID;                    \
    }                                           \
  } while (0)

static inline int64_t round_up(int64_t x, int64_t m) { return (x + m - 1) / m * m; }
static inline int cdiv(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

constexpr int kNumSMs = 148;   // B200: 2 dies x 74 SMs; grids below are sized in multiples of it

}  // namespace b200
