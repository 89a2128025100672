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
int set_sample_fine2(int v);  // volume.cu: 0 = one-ray-per-warp sample_fine kernel only; returns the previous value
int set_composite_lc(int v);  // volume.cu: 0 = strided compositing kernels only; returns the previous value

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
// torch.linspace bit for bit on the GPU for S in {2, 33, 64, 128} (tests/test_gpu_stages.py, sample_coarse).
__device__ __forceinline__ float linspace01(int i, int n) {
  if (n <= 1) return 0.0f;
  const float step = __fdiv_rn(1.0f, (float)(n - 1));
  if (i < n / 2) return __fmul_rn(step, (float)i);
  return __fmaf_rn(-step, (float)(n - 1 - i), 1.0f);
}

// ------------------------------------------------------------------------------------ in-kernel random numbers
// SURVEY 8(f) f4.  The reference draws the stratified jitter (run_nerf.py:192-199 / train_utils.py:58-65), the
// inverse-CDF uniforms (run_nerf_helpers.py:192-197) and the sigma noise (run_nerf.py:130-132 /
// volume_rendering_utils.py:27-38) from the framework generator; here they come from Philox4x32-10 evaluated inside
// the sampling / compositing kernels, so a training step reads no random arrays and launches no generator kernels.
// Counter-based: element (ray, s) of stream `stream` at step state[1] under seed state[0] is a pure function of those
// five numbers -- the compositing backward regenerates the noise its forward used, and nb_rng_fill (stages.cu)
// materialises exactly the arrays the kernels see (tests feed them back through the explicit-array path).
//   counter = { ray[31:0], ray[47:32] | (s & 31) << 16 | (s >> 7) << 21 | stream << 30, step[31:0], step[63:32] }
//   key     = seed;  the four outputs serve s, s+32, s+64, s+96 (one lane's consecutive sample groups)
struct Rng {
  const unsigned long long* state;  // device: {seed, step}; null = off
  unsigned stream;                  // 0 coarse jitter, 1 fine uniforms, 2 coarse noise, 3 fine noise
  float scale;                      // normal deviates are multiplied by this (raw_noise_std)
};
enum { RNG_JITTER = 0, RNG_FINE_U = 1, RNG_NOISE_C = 2, RNG_NOISE_F = 3 };

__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    const unsigned hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
    const unsigned hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += 0x9E3779B9u;
    k.y += 0xBB67AE85u;
  }
  return c;
}
__device__ __forceinline__ uint4 rng_block(const Rng& r, long long ray, int s) {
  const unsigned long long seed = r.state[0], step = r.state[1];
  const uint4 c = make_uint4((unsigned)ray,
                             ((unsigned)((unsigned long long)ray >> 32) & 0xFFFFu) | ((unsigned)(s & 31) << 16) |
                                 (((unsigned)(s >> 7) & 0x1FFu) << 21) | (r.stream << 30),
                             (unsigned)step, (unsigned)(step >> 32));
  return philox4x32_10(c, make_uint2((unsigned)seed, (unsigned)(seed >> 32)));
}
__device__ __forceinline__ float rng_u01(unsigned x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }  // [0, 1)
// uniform in [0, 1) for element (ray, s)
__device__ __forceinline__ float rng_uniform(const Rng& r, long long ray, int s) {
  const uint4 b = rng_block(r, ray, s);
  const int comp = (s >> 5) & 3;
  return rng_u01(comp == 0 ? b.x : comp == 1 ? b.y : comp == 2 ? b.z : b.w);
}
// scale * N(0, 1) for element (ray, s): Box-Muller on the output pair the element's component belongs to
__device__ __forceinline__ float rng_normal(const Rng& r, long long ray, int s) {
  const uint4 b = rng_block(r, ray, s);
  const int comp = (s >> 5) & 3;
  const float ua = rng_u01(comp < 2 ? b.x : b.z), ub = rng_u01(comp < 2 ? b.y : b.w);
  const float rad = sqrtf(-2.0f * logf(1.0f - ua));  // 1 - ua in (0, 1]
  float sn, cs;
  sincosf(6.283185307179586f * ub, &sn, &cs);
  return r.scale * rad * ((comp & 1) ? sn : cs);
}

// internal entry points with the generator argument (the extern "C" ones of include/nerf_b200.h pass it switched off)
int sample_coarse_impl(const float* rays, int stride, int64_t n, int S, int lindisp, const float* t_rand, Rng rng, float* z,
                       nb_stream_t stream);
int composite_impl(const float* raw, const float* z, const float* rays, int stride, int64_t n, int S, const float* noise,
                   Rng rng, int white, float* rgb, float* disp, float* acc, float* depth, float* weights, nb_stream_t stream);
int composite_backward_impl(const float* raw, const float* z, const float* rays, int stride, int64_t n, int S,
                            const float* noise, Rng rng, int white, const float* d_rgb, const float* d_disp,
                            const float* d_acc, const float* d_depth, float* d_raw, nb_stream_t stream);
int sample_fine_impl(const float* z, const float* weights, int64_t n, int S, const float* u, Rng rng, int ns,
                     float* z_samples, float* z_fine, float* z_std, nb_stream_t stream);

}  // namespace nb
