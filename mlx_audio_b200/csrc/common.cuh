// Shared helpers for libkokoro_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

#include "../../include/kokoro_b200.h"

namespace kk {

// thread-local last-error string (C ABI: kk_last_error)
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);

#define KK_CHECK_ARG(cond, ...)                 \
  do {                                          \
    if (!(cond)) {                              \
      kk::set_error(__VA_ARGS__);               \
      return KK_ERR_INVALID_ARG;                \
    }                                           \
  } while (0)

#define KK_CUDA(expr)                                        \
  do {                                                       \
    cudaError_t _e = (expr);                                 \
    if (_e != cudaSuccess) return kk::cuda_fail(_e, #expr);  \
  } while (0)

#define KK_LAUNCH_CHECK(name)                                     \
  do {                                                            \
    cudaError_t _e = cudaPeekAtLastError();                       \
    if (_e != cudaSuccess) return kk::cuda_fail(_e, name);        \
  } while (0)

static inline int cdiv(long long a, long long b) { return (int)((a + b - 1) / b); }

__device__ __forceinline__ float leaky(float v, float slope) { return v > 0.f ? v : v * slope; }

// Ragged batches (kokoro_b200.h): rows of utterance b at a level with mul*T + add rows per utterance, capped by the
// padded length L.  lens == NULL: L.
__device__ __forceinline__ int ragged_len(const int* __restrict__ lens, int mul, int add, int b, int L) {
#ifdef KK_NO_RAGGED   // lab builds: compile the ragged paths out (A/B of their cost)
  return L;
#endif
  if (!lens) return L;
  const int v = __ldg(lens + b) * mul + add;
  return v < L ? (v < 0 ? 0 : v) : L;
}

}  // namespace kk
