// Shared host/device helpers for the mappo_b200 C-ABI library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include "../../include/mappo_b200.h"

namespace mappo {

void set_error(const char* fmt, ...);
void count_launch(int n = 1);

#define MAPPO_REQUIRE(cond, ...)                      \
  do {                                                \
    if (!(cond)) {                                    \
      ::mappo::set_error(__VA_ARGS__);                \
      return MAPPO_ERR_INVALID_ARG;                   \
    }                                                 \
  } while (0)

#define MAPPO_UNSUPPORTED(...)                        \
  do {                                                \
    ::mappo::set_error(__VA_ARGS__);                  \
    return MAPPO_ERR_UNSUPPORTED;                     \
  } while (0)

// after a launch: cudaGetLastError only (no sync)
#define MAPPO_CHECK_LAUNCH(what)                                                          \
  do {                                                                                    \
    cudaError_t e__ = cudaGetLastError();                                                 \
    if (e__ != cudaSuccess) {                                                             \
      ::mappo::set_error("%s: %s", what, cudaGetErrorString(e__));                        \
      return MAPPO_ERR_CUDA;                                                              \
    }                                                                                     \
    ::mappo::count_launch();                                                              \
  } while (0)

#define MAPPO_CUDA(call)                                                                  \
  do {                                                                                    \
    cudaError_t e__ = (call);                                                             \
    if (e__ != cudaSuccess) {                                                             \
      ::mappo::set_error("%s: %s", #call, cudaGetErrorString(e__));                       \
      return MAPPO_ERR_CUDA;                                                              \
    }                                                                                     \
  } while (0)

static inline int ceil_div(long long a, long long b) { return (int)((a + b - 1) / b); }

}  // namespace mappo
