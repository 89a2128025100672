// Fused Adam step over a flat fp32 parameter buffer: what optimizer.step(); optimizer.zero_grad()
// (nerf-pytorch/train_nerf.py:288-289) or Adam.apply_gradients (nerf/run_nerf.py:827) do with ~7 multi-tensor launches,
// as one HBM-bound pass: 16 B read + 12 B written per parameter (+4 B to clear the gradient).
// Compiled --fmad=false so every product/sum rounds separately like the framework's element-wise kernels.
#include <algorithm>

#include "common.cuh"

namespace nb {

__global__ void adam_kernel(float4* __restrict__ p, float4* __restrict__ g, float4* __restrict__ m, float4* __restrict__ v,
                            long long n4, long long n, const float* __restrict__ hyper, int tf_style, int zero_grads) {
  // 1-beta are passed separately: the frameworks form them in double before rounding (1.0f - 0.999f is off by 1e-5)
  const float lr = hyper[0], b2 = hyper[2], eps = hyper[3], bc1 = hyper[4], bc2 = hyper[5], omb1 = hyper[6], omb2 = hyper[7];
  const float step_size = tf_style ? lr * sqrtf(bc2) / bc1 : lr / bc1;
  const float bc2_sqrt = sqrtf(bc2);
  auto upd = [&](float& pp, float gg, float& mm, float& vv) {
    mm = mm + (gg - mm) * omb1;                    // torch: exp_avg.lerp_(grad, 1 - beta1)
    vv = vv * b2 + (gg * gg) * omb2;               // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, 1 - beta2)
    const float denom = tf_style ? sqrtf(vv) + eps : sqrtf(vv) / bc2_sqrt + eps;
    pp = pp - step_size * (mm / denom);
  };
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    float4 pp = p[i], gg = g[i], mm = m[i], vv = v[i];
    upd(pp.x, gg.x, mm.x, vv.x);
    upd(pp.y, gg.y, mm.y, vv.y);
    upd(pp.z, gg.z, mm.z, vv.z);
    upd(pp.w, gg.w, mm.w, vv.w);
    p[i] = pp; m[i] = mm; v[i] = vv;
    if (zero_grads) g[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  // tail (n not a multiple of 4)
  if (blockIdx.x == 0 && threadIdx.x < (int)(n - 4 * n4)) {
    const long long i = 4 * n4 + threadIdx.x;
    float* ps = reinterpret_cast<float*>(p);
    float* gs = reinterpret_cast<float*>(g);
    float* ms = reinterpret_cast<float*>(m);
    float* vs = reinterpret_cast<float*>(v);
    float pp = ps[i], mm = ms[i], vv = vs[i];
    upd(pp, gs[i], mm, vv);
    ps[i] = pp; ms[i] = mm; vs[i] = vv;
    if (zero_grads) gs[i] = 0.0f;
  }
}

}  // namespace nb

using namespace nb;

extern "C" int nb_adam_step(float* params, float* grads, float* exp_avg, float* exp_avg_sq, int64_t n, const float* hyper,
                            int tf_style, int zero_grads, nb_stream_t stream) {
  NB_REQUIRE(params && grads && exp_avg && exp_avg_sq && hyper, "nb_adam_step: null pointer");
  NB_REQUIRE(n >= 0, "nb_adam_step: bad size");
  NB_REQUIRE(((reinterpret_cast<uintptr_t>(params) | reinterpret_cast<uintptr_t>(grads) | reinterpret_cast<uintptr_t>(exp_avg) |
               reinterpret_cast<uintptr_t>(exp_avg_sq)) & 15) == 0, "nb_adam_step: buffers must be 16-byte aligned");
  if (n == 0) return NB_OK;
  const long long n4 = n / 4;
  const int grid = (int)std::max<int64_t>(1, std::min<int64_t>(ceil_div(std::max<long long>(n4, 1), 256), 4 * sm_count()));
  adam_kernel<<<grid, 256, 0, as_stream(stream)>>>(reinterpret_cast<float4*>(params), reinterpret_cast<float4*>(grads),
                                                   reinterpret_cast<float4*>(exp_avg), reinterpret_cast<float4*>(exp_avg_sq),
                                                   n4, (long long)n, hyper, tf_style, zero_grads);
  NB_LAUNCHED();
  return NB_OK;
}
