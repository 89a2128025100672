// Shared host/device helpers for the sm_100a kernels behind include/nerf_b200.h.
#pragma once
#include <cuda_runtime.h>

#include <atomic>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/nerf_b200.h"

namespace nb {

// ---- error plumbing (thread-local message, never throws across the ABI)
void set_error(const char* fmt, ...);
extern std::atomic<int64_t> g_launches;  // process-wide: autograd runs backward on its own thread

#define NB_REQUIRE(cond, ...)                \
  do {                                       \
    if (!(cond)) {                           \
      nb::set_error(__VA_ARGS__);            \
      return NB_ERR_INVALID;                 \
    }                                        \
  } while (0)

#define NB_CUDA(call)                                                                  \
  do {                                                                                 \
    cudaError_t e_ = (call);                                                           \
    if (e_ != cudaSuccess) {                                                           \
      nb::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
      return NB_ERR_CUDA;                                                              \
    }                                                                                  \
  } while (0)

// after a <<<>>> launch
#define NB_LAUNCHED()                                                                  \
  do {                                                                                 \
    nb::g_launches++;                                                                  \
    cudaError_t e_ = cudaGetLastError();                                               \
    if (e_ != cudaSuccess) {                                                           \
      nb::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(e_), __FILE__, __LINE__); \
      return NB_ERR_CUDA;                                                              \
    }                                                                                  \
  } while (0)

static inline cudaStream_t as_stream(nb_stream_t s) { return reinterpret_cast<cudaStream_t>(s); }

static inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline int64_t round_up(int64_t a, int64_t b) { return ceil_div(a, b) * b; }

int sm_count();  // SMs of the current device (cached)

// ---- device helpers
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// streaming 128-bit load / store (data touched exactly once: keep it out of L1)
__device__ __forceinline__ float4 ld_stream4(const float4* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ float ld_stream(const float* p) {
  float r;
  asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(r) : "l"(p));
  return r;
}

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// torch.linspace(0, 1, n)[i] in fp32 (CPU kernel): step = fp32(1/(n-1)); the lower half counts up
// from 0, the upper half counts down from 1 with a single rounding (fma).  Verified against
// torch.linspace for n = 2..512 (tests/test_host_formulas.py restates it in numpy).
__device__ __forceinline__ float linspace01(int i, int n) {
  if (n <= 1) return 0.0f;
  const float step = __fdiv_rn(1.0f, (float)(n - 1));
  if (i < n / 2) return __fmul_rn(step, (float)i);
  return __fmaf_rn(-step, (float)(n - 1 - i), 1.0f);
}

}  // namespace nb
