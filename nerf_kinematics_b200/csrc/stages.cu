// HBM-bound stage kernels of the classic-NeRF hot path.
// Compiled with --fmad=false: the reference evaluates these expressions as separate fp32
// multiplies and adds, and the positional encoding amplifies a 1-ulp difference in a point
// coordinate by up to 2^9, so contraction is disabled for the whole file (all kernels here are
// bandwidth-bound, FMA would buy nothing).
#include <math.h>

#include <algorithm>
#include <stdarg.h>

#include "common.cuh"

namespace nb {

thread_local char g_err[512] = "";
std::atomic<int64_t> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int sm_count() {
  static int cached = 0;
  if (cached == 0) {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) == cudaSuccess &&
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0)
      cached = n;
    else
      return 148;
  }
  return cached;
}

struct Cam {
  float r[12];  // 3x4 row-major
};

// ------------------------------------------------------------------------------------ a1
// get_rays (nerf/run_nerf_helpers.py:123-130 == get_ray_bundle, nerf-pytorch/nerf/nerf_helpers.py:67-110):
// dir_cam = [(col - cx)/fx, -(row - cy)/fy, -1]; rays_d[k] = sum_j dir_cam[j] * R[k][j] (multiply, then sum over the
// last axis in this association); rays_o = c2w[:3, 3].  True IEEE divisions, no +0.5 pixel centre.
__device__ __forceinline__ void pixel_ray(int col, int row, float fx, float fy, float cx, float cy, const Cam& cam,
                                          float d[3]) {
  const float dx = ((float)col - cx) / fx;
  const float dy = -(((float)row - cy) / fy);
  const float dz = -1.0f;
#pragma unroll
  for (int k = 0; k < 3; ++k) d[k] = (dx * cam.r[4 * k + 0] + dy * cam.r[4 * k + 1]) + dz * cam.r[4 * k + 2];
}

// grid = (ceil(W / 256), ceil(rows / kRaysRows)): a block owns up to 256 consecutive pixels of kRaysRows rows; per row it
// stages the 768 direction floats in shared memory and writes them back as coalesced 16-byte stores (scalar stores when
// the row segment is not 16-byte aligned: W * 3 not a multiple of 4); the origins are the same 12 bytes for every ray.
constexpr int kRaysRows = 8;
__global__ void __launch_bounds__(256)
get_rays_kernel(int W, float fx, float fy, float cx, float cy, Cam cam, const float* __restrict__ cam_dev, int row0,
                int n_rows, float* __restrict__ ro, float* __restrict__ rd) {
  __shared__ __align__(16) float sd[768];
  if (cam_dev != nullptr) {  // pose in device memory (a captured frame graph reads the pose of the replay, not of the capture)
#pragma unroll
    for (int k = 0; k < 12; ++k) cam.r[k] = __ldg(cam_dev + k);
  }
  const int col0 = blockIdx.x * 256, col = col0 + threadIdx.x;
  const int npix = min(256, W - col0);
  const int nflt = 3 * npix;
  // origin pattern of this thread's float4 slot: element i of a row segment holds cam.r[4 * (i % 3) + 3]
  float4 o4;
  {
    const int i0 = (4 * threadIdx.x) % 3;
    o4.x = cam.r[4 * i0 + 3];
    o4.y = cam.r[4 * ((i0 + 1) % 3) + 3];
    o4.z = cam.r[4 * ((i0 + 2) % 3) + 3];
    o4.w = cam.r[4 * i0 + 3];
  }
  const int r_begin = blockIdx.y * kRaysRows, r_end = min(n_rows, r_begin + kRaysRows);
  for (int r = r_begin; r < r_end; ++r) {
    if (col < W) {
      float d[3];
      pixel_ray(col, row0 + r, fx, fy, cx, cy, cam, d);
      sd[3 * threadIdx.x + 0] = d[0];
      sd[3 * threadIdx.x + 1] = d[1];
      sd[3 * threadIdx.x + 2] = d[2];
    }
    __syncthreads();
    const int64_t base = ((int64_t)r * W + col0) * 3;
    if ((base & 3) == 0 && (nflt & 3) == 0) {
      float4* rd4 = reinterpret_cast<float4*>(rd + base);
      float4* ro4 = reinterpret_cast<float4*>(ro + base);
      if ((int)threadIdx.x < (nflt >> 2)) {
        rd4[threadIdx.x] = reinterpret_cast<const float4*>(sd)[threadIdx.x];
        ro4[threadIdx.x] = o4;
      }
    } else {
      for (int i = threadIdx.x; i < nflt; i += 256) {
        rd[base + i] = sd[i];
        ro[base + i] = cam.r[4 * (i % 3) + 3];
      }
    }
    __syncthreads();
  }
}

// rays of selected pixels only (the training loop draws N_rand pixel indices per step and the
// reference generates the whole image's rays to keep 1 024 of them, nerf-pytorch/train_nerf.py:240-253).
__global__ void get_rays_at_kernel(int W, float fx, float fy, float cx, float cy, Cam cam, const float* __restrict__ cam_dev,
                                   const int64_t* __restrict__ pix, int64_t n, float* __restrict__ ro,
                                   float* __restrict__ rd) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (cam_dev != nullptr) {
#pragma unroll
    for (int k = 0; k < 12; ++k) cam.r[k] = __ldg(cam_dev + k);
  }
  const int64_t p = pix[i];
  const int row = (int)(p / W), col = (int)(p - (int64_t)row * W);
  float d[3];
  pixel_ray(col, row, fx, fy, cx, cy, cam, d);
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    rd[i * 3 + k] = d[k];
    ro[i * 3 + k] = cam.r[4 * k + 3];
  }
}

// ------------------------------------------------------------------------------------ f3: to8b + squared error
__global__ void to8b_kernel(const float* __restrict__ x, const float* __restrict__ gt, int64_t n,
                            unsigned char* __restrict__ out8, double* __restrict__ mse_accum) {
  double local = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float v = x[i];
    if (out8 != nullptr) out8[i] = (unsigned char)(255.0f * fminf(fmaxf(v, 0.0f), 1.0f));  // numpy astype(uint8): truncation
    if (gt != nullptr) {
      const float d = v - gt[i];
      local += (double)(d * d);
    }
  }
  if (gt != nullptr && mse_accum != nullptr) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
    if ((threadIdx.x & 31) == 0) atomicAdd(mse_accum, local);
  }
}

// ------------------------------------------------------------------------------------ a3
struct NdcScale {
  float sx, sy, near, two_near, neg_two_near;
};

__device__ __forceinline__ void ndc_one(const NdcScale& s, float o[3], float d[3]) {
  float t = (-(s.near + o[2])) / d[2];
  float ox = o[0] + t * d[0], oy = o[1] + t * d[1], oz = o[2] + t * d[2];
  float o0 = (s.sx * ox) / oz;
  float o1 = (s.sy * oy) / oz;
  float o2 = 1.0f + s.two_near / oz;
  float d0 = s.sx * (d[0] / d[2] - ox / oz);
  float d1 = s.sy * (d[1] / d[2] - oy / oz);
  float d2 = s.neg_two_near / oz;
  o[0] = o0; o[1] = o1; o[2] = o2;
  d[0] = d0; d[1] = d1; d[2] = d2;
}

__global__ void ndc_kernel(NdcScale s, const float* __restrict__ ro, const float* __restrict__ rd, int64_t n,
                           float* __restrict__ oo, float* __restrict__ od) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float o[3] = {ro[3 * i], ro[3 * i + 1], ro[3 * i + 2]};
  float d[3] = {rd[3 * i], rd[3 * i + 1], rd[3 * i + 2]};
  ndc_one(s, o, d);
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    oo[3 * i + k] = o[k];
    od[3 * i + k] = d[k];
  }
}

// ------------------------------------------------------------------------------------ a2+a3+a4
__global__ void ray_batch_kernel(const float* __restrict__ ro, const float* __restrict__ rd,
                                 const float* __restrict__ vsrc, int64_t n, int use_viewdirs, int ndc, NdcScale s,
                                 float near, float far, const float* __restrict__ near_arr,
                                 const float* __restrict__ far_arr, float* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int stride = use_viewdirs ? 11 : 8;
  float o[3] = {ro[3 * i], ro[3 * i + 1], ro[3 * i + 2]};
  float d[3] = {rd[3 * i], rd[3 * i + 1], rd[3 * i + 2]};
  float v[3] = {0.f, 0.f, 0.f};
  if (use_viewdirs) {
    const float* src = vsrc ? vsrc : rd;
    float a = src[3 * i], b = src[3 * i + 1], c = src[3 * i + 2];
    float nrm = sqrtf((a * a + b * b) + c * c);
    v[0] = a / nrm; v[1] = b / nrm; v[2] = c / nrm;
  }
  if (ndc) ndc_one(s, o, d);
  float* q = out + i * stride;
  q[0] = o[0]; q[1] = o[1]; q[2] = o[2];
  q[3] = d[0]; q[4] = d[1]; q[5] = d[2];
  q[6] = near_arr ? near_arr[i] : near;
  q[7] = far_arr ? far_arr[i] : far;
  if (use_viewdirs) { q[8] = v[0]; q[9] = v[1]; q[10] = v[2]; }
}

// ------------------------------------------------------------------------------------ a5
__device__ __forceinline__ float coarse_z(float near, float far, int s, int S, int lindisp) {
  float t = linspace01(s, S);
  if (!lindisp) return near * (1.0f - t) + far * t;
  return 1.0f / ((1.0f / near) * (1.0f - t) + (1.0f / far) * t);
}

__global__ void sample_coarse_kernel(const float* __restrict__ rays, int stride, int64_t n, int S, int lindisp,
                                     const float* __restrict__ t_rand, Rng rng, float* __restrict__ z) {
  int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n * S) return;
  int64_t r = idx / S;
  int s = (int)(idx - r * S);
  float near = rays[r * stride + 6], far = rays[r * stride + 7];
  float zc = coarse_z(near, far, s, S, lindisp);
  if (t_rand != nullptr || rng.state != nullptr) {
    float lower = (s == 0) ? zc : 0.5f * (zc + coarse_z(near, far, s - 1, S, lindisp));
    float upper = (s == S - 1) ? zc : 0.5f * (coarse_z(near, far, s + 1, S, lindisp) + zc);
    zc = lower + (upper - lower) * (t_rand != nullptr ? t_rand[idx] : rng_uniform(rng, r, s));
  }
  z[idx] = zc;
}

// f4: the generator's step counter lives on the device so that a captured train step advances it on every replay
__global__ void rng_advance_kernel(unsigned long long* state) { state[1] += 1ull; }

// the numbers the kernels draw for rows of length row_len, as arrays (tests; hosts that want to log them)
__global__ void rng_fill_kernel(Rng rng, int normal, int64_t n, int row_len, float* __restrict__ out) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n * row_len) return;
  const int64_t r = idx / row_len;
  const int s = (int)(idx - r * row_len);
  out[idx] = normal ? rng_normal(rng, r, s) : rng_uniform(rng, r, s);
}

// ------------------------------------------------------------------------------------ a6
__global__ void posenc_kernel(const float* __restrict__ x, int64_t n, int L, int include_input, int C,
                              float* __restrict__ out, int out_stride) {
  int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n * C) return;
  int64_t p = idx / C;
  int c = (int)(idx - p * C);
  float v;
  if (include_input && c < 3) {
    v = x[3 * p + c];
  } else {
    int cc = c - (include_input ? 3 : 0);
    int f = cc / 6, r = cc - 6 * f;
    int comp = r % 3;
    float a = x[3 * p + comp] * exp2f((float)f);
    v = (r < 3) ? sinf(a) : cosf(a);
  }
  out[p * out_stride + c] = v;
}

// ------------------------------------------------------------------------------------ a13
__global__ void mse_kernel(const float* __restrict__ pred, const float* __restrict__ tgt, int64_t count,
                           float grad_scale, float* __restrict__ loss, float* __restrict__ d_pred) {
  float acc = 0.f;
  const float inv = 1.0f / (float)count;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x) {
    float df = pred[i] - tgt[i];
    acc += df * df;
    if (d_pred) d_pred[i] = grad_scale * (2.0f * df * inv);
  }
  acc = warp_sum(acc);
  __shared__ float part[32];
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) part[w] = acc;
  __syncthreads();
  if (w == 0) {
    float v = (lane < (int)(blockDim.x >> 5)) ? part[lane] : 0.f;
    v = warp_sum(v);
    if (lane == 0) atomicAdd(loss, v * inv);
  }
}

int sample_coarse_impl(const float* rays, int stride, int64_t n, int S, int lindisp, const float* t_rand, Rng rng, float* z,
                       nb_stream_t stream) {
  NB_REQUIRE(n >= 0 && S > 0 && (stride == 8 || stride == 11) && (n == 0 || (rays && z)), "nb_sample_coarse: bad arguments");
  if (n == 0) return NB_OK;
  sample_coarse_kernel<<<(unsigned)ceil_div(n * S, 256), 256, 0, as_stream(stream)>>>(rays, stride, n, S, lindisp, t_rand, rng, z);
  NB_LAUNCHED();
  return NB_OK;
}

}  // namespace nb

using namespace nb;

// ====================================================================================== C ABI
extern "C" {

int nb_abi_version(void) { return NB_ABI_VERSION; }
const char* nb_last_error(void) { return nb::g_err; }
int64_t nb_launch_count(void) { return nb::g_launches.load(); }
void nb_launch_count_reset(void) { nb::g_launches.store(0); }

int nb_device_count(void) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) {
    set_error("cudaGetDeviceCount: %s", cudaGetErrorString(e));
    cudaGetLastError();
    return NB_ERR_CUDA;
  }
  return n;
}

static int get_rays_common(int H, int W, float fx, float fy, float cx, float cy, const float* c2w_host,
                           const float* c2w_dev, int row0, int row1, float* rays_o, float* rays_d, nb_stream_t stream) {
  NB_REQUIRE(H > 0 && W > 0 && (c2w_host || c2w_dev), "nb_get_rays: bad arguments");
  NB_REQUIRE(0 <= row0 && row0 <= row1 && row1 <= H, "nb_get_rays: rows [%d,%d) outside image height %d", row0, row1, H);
  int64_t n = (int64_t)(row1 - row0) * W;
  if (n == 0) return NB_OK;
  NB_REQUIRE(rays_o && rays_d, "nb_get_rays: null output");
  Cam cam{};
  if (c2w_host) memcpy(cam.r, c2w_host, sizeof(cam.r));
  NB_REQUIRE(row1 - row0 <= 65535 * kRaysRows, "nb_get_rays: too many rows in one call");
  dim3 grid((unsigned)ceil_div(W, 256), (unsigned)ceil_div(row1 - row0, kRaysRows));
  get_rays_kernel<<<grid, 256, 0, as_stream(stream)>>>(W, fx, fy, cx, cy, cam, c2w_dev, row0, row1 - row0, rays_o, rays_d);
  NB_LAUNCHED();
  return NB_OK;
}

int nb_get_rays(int H, int W, float fx, float fy, float cx, float cy, const float* c2w, int row0, int row1,
                float* rays_o, float* rays_d, nb_stream_t stream) {
  NB_REQUIRE(c2w != nullptr, "nb_get_rays: null pose");
  return get_rays_common(H, W, fx, fy, cx, cy, c2w, nullptr, row0, row1, rays_o, rays_d, stream);
}

int nb_get_rays_dev(int H, int W, float fx, float fy, float cx, float cy, const float* c2w34_dev, int row0, int row1,
                    float* rays_o, float* rays_d, nb_stream_t stream) {
  NB_REQUIRE(c2w34_dev != nullptr, "nb_get_rays_dev: null pose");
  return get_rays_common(H, W, fx, fy, cx, cy, nullptr, c2w34_dev, row0, row1, rays_o, rays_d, stream);
}

// TF check_numerics (nerf/run_nerf.py:240-241) without a host round trip per tensor: flag |= 1 when any element is NaN/Inf
__global__ void check_finite_kernel(const float* __restrict__ x, int64_t n, int* __restrict__ flag, int bit) {
  bool bad = false;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    bad |= !isfinite(x[i]);
  if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) atomicOr(flag, bit);
}

int nb_check_finite(const float* x, int64_t n, int* flag, int bit, nb_stream_t stream) {
  NB_REQUIRE(n >= 0 && flag != nullptr && (n == 0 || x != nullptr), "nb_check_finite: bad arguments");
  if (n == 0) return NB_OK;
  const unsigned grid = (unsigned)std::min<int64_t>(ceil_div(n, 256), (int64_t)sm_count() * 8);
  check_finite_kernel<<<grid, 256, 0, as_stream(stream)>>>(x, n, flag, bit);
  NB_LAUNCHED();
  return NB_OK;
}

int nb_to8b(const float* x, const float* gt, int64_t n, unsigned char* out8, double* mse_accum, nb_stream_t stream) {
  NB_REQUIRE(n >= 0, "nb_to8b: negative size");
  if (n == 0) return NB_OK;
  NB_REQUIRE(x && (out8 || gt), "nb_to8b: null pointer");
  NB_REQUIRE(gt == nullptr || mse_accum != nullptr, "nb_to8b: gt given without an accumulator");
  const unsigned grid = (unsigned)std::min<int64_t>(ceil_div(n, 256), (int64_t)sm_count() * 8);
  to8b_kernel<<<grid, 256, 0, as_stream(stream)>>>(x, gt, n, out8, mse_accum);
  NB_LAUNCHED();
  return NB_OK;
}

static int get_rays_at_common(int H, int W, float fx, float fy, float cx, float cy, const float* c2w_host,
                              const float* c2w_dev, const int64_t* pixel_index, int64_t n, float* rays_o, float* rays_d,
                              nb_stream_t stream) {
  NB_REQUIRE(H > 0 && W > 0 && (c2w_host || c2w_dev) && n >= 0, "nb_get_rays_at: bad arguments");
  if (n == 0) return NB_OK;
  NB_REQUIRE(pixel_index && rays_o && rays_d, "nb_get_rays_at: null pointer");
  Cam cam{};
  if (c2w_host) memcpy(cam.r, c2w_host, sizeof(cam.r));
  get_rays_at_kernel<<<(unsigned)ceil_div(n, 256), 256, 0, as_stream(stream)>>>(W, fx, fy, cx, cy, cam, c2w_dev, pixel_index, n,
                                                                           rays_o, rays_d);
  NB_LAUNCHED();
  return NB_OK;
}

int nb_get_rays_at(int H, int W, float fx, float fy, float cx, float cy, const float* c2w, const int64_t* pixel_index,
                   int64_t n, float* rays_o, float* rays_d, nb_stream_t stream) {
  NB_REQUIRE(c2w != nullptr, "nb_get_rays_at: null pose");
  return get_rays_at_common(H, W, fx, fy, cx, cy, c2w, nullptr, pixel_index, n, rays_o, rays_d, stream);
}

int nb_get_rays_at_dev(int H, int W, float fx, float fy, float cx, float cy, const float* c2w34_dev,
                       const int64_t* pixel_index, int64_t n, float* rays_o, float* rays_d, nb_stream_t stream) {
  NB_REQUIRE(c2w34_dev != nullptr, "nb_get_rays_at_dev: null pose");
  return get_rays_at_common(H, W, fx, fy, cx, cy, nullptr, c2w34_dev, pixel_index, n, rays_o, rays_d, stream);
}

static NdcScale make_ndc_scale(int H, int W, double focal, double near) {
  // evaluated in double like the Python scalars of the reference, then rounded once to fp32
  NdcScale s;
  s.sx = (float)(-1.0 / (W / (2.0 * focal)));
  s.sy = (float)(-1.0 / (H / (2.0 * focal)));
  s.near = (float)near;
  s.two_near = (float)(2.0 * near);
  s.neg_two_near = (float)(-2.0 * near);
  return s;
}

int nb_ndc_rays(int H, int W, double focal, double near, const float* ro, const float* rd, int64_t n, float* oo,
                float* od, nb_stream_t stream) {
  NB_REQUIRE(H > 0 && W > 0 && n >= 0 && (n == 0 || (ro && rd && oo && od)), "nb_ndc_rays: bad arguments");
  if (n == 0) return NB_OK;
  ndc_kernel<<<(unsigned)ceil_div(n, 256), 256, 0, as_stream(stream)>>>(make_ndc_scale(H, W, focal, near), ro, rd, n, oo, od);
  NB_LAUNCHED();
  return NB_OK;
}

int nb_make_ray_batch(const float* ro, const float* rd, const float* vsrc, int64_t n, int use_viewdirs, int ndc, int H,
                      int W, double focal, double near, double far, const float* near_arr, const float* far_arr,
                      float* out, nb_stream_t stream) {
  NB_REQUIRE(n >= 0 && (n == 0 || (ro && rd && out)), "nb_make_ray_batch: bad arguments");
  NB_REQUIRE(!ndc || (H > 0 && W > 0), "nb_make_ray_batch: ndc needs H, W");
  if (n == 0) return NB_OK;
  NdcScale s = ndc ? make_ndc_scale(H, W, focal, 1.0) : NdcScale{0, 0, 0, 0, 0};
  ray_batch_kernel<<<(unsigned)ceil_div(n, 256), 256, 0, as_stream(stream)>>>(ro, rd, vsrc, n, use_viewdirs, ndc, s, (float)near, (float)far,
                                                                             near_arr, far_arr, out);
  NB_LAUNCHED();
  return NB_OK;
}

int nb_sample_coarse(const float* rays, int stride, int64_t n, int S, int lindisp, const float* t_rand, float* z,
                     nb_stream_t stream) {
  return nb::sample_coarse_impl(rays, stride, n, S, lindisp, t_rand, Rng{nullptr, 0u, 0.0f}, z, stream);
}

int nb_rng_advance(uint64_t* state, nb_stream_t stream) {
  NB_REQUIRE(state != nullptr, "nb_rng_advance: null state");
  rng_advance_kernel<<<1, 1, 0, as_stream(stream)>>>(reinterpret_cast<unsigned long long*>(state));
  NB_LAUNCHED();
  return NB_OK;
}

int nb_rng_fill(const uint64_t* state, int stream_id, int normal, float scale, int64_t n, int row_len, float* out,
                nb_stream_t stream) {
  NB_REQUIRE(state != nullptr && stream_id >= 0 && stream_id < 4 && n >= 0 && row_len > 0 && row_len <= 65536 &&
                 (n == 0 || out != nullptr),
             "nb_rng_fill: bad arguments");
  if (n == 0) return NB_OK;
  const Rng rng{reinterpret_cast<const unsigned long long*>(state), (unsigned)stream_id, scale};
  rng_fill_kernel<<<(unsigned)ceil_div(n * row_len, 256), 256, 0, as_stream(stream)>>>(rng, normal, n, row_len, out);
  NB_LAUNCHED();
  return NB_OK;
}

int nb_posenc(const float* x, int64_t n, int L, int include_input, float* out, int out_stride, nb_stream_t stream) {
  int C = (include_input ? 3 : 0) + 6 * L;
  NB_REQUIRE(n >= 0 && L >= 0 && C > 0 && out_stride >= C && (n == 0 || (x && out)), "nb_posenc: bad arguments");
  if (n == 0) return NB_OK;
  posenc_kernel<<<(unsigned)ceil_div(n * C, 256), 256, 0, as_stream(stream)>>>(x, n, L, include_input, C, out, out_stride);
  NB_LAUNCHED();
  return NB_OK;
}

int nb_mse_loss(const float* pred, const float* tgt, int64_t count, float grad_scale, float* loss, float* d_pred,
                nb_stream_t stream) {
  NB_REQUIRE(count > 0 && pred && tgt && loss, "nb_mse_loss: bad arguments");
  if (count <= (1 << 16)) {
    // a training batch (N_rand <= 21 845 rays): one block, so the sum has one order and the loss is the same bits on
    // every run; larger inputs (whole images) spread over the machine and combine their block sums atomically
    mse_kernel<<<1, 1024, 0, as_stream(stream)>>>(pred, tgt, count, grad_scale, loss, d_pred);
    NB_LAUNCHED();
    return NB_OK;
  }
  int blocks = (int)std::min<int64_t>(ceil_div(count, 256), 4 * sm_count());
  mse_kernel<<<blocks, 256, 0, as_stream(stream)>>>(pred, tgt, count, grad_scale, loss, d_pred);
  NB_LAUNCHED();
  return NB_OK;
}

}  // extern "C"
