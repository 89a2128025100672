// Per-chunk renderer: the body of render_rays (nerf/run_nerf.py:48-243) == predict_and_render_radiance
// (nerf-pytorch/nerf/train_utils.py:28-127) as one C-ABI call, and its backward for the train step.
// It sequences the stage kernels of stages.cu / volume.cu and the MLP passes on the caller's stream;
// all intermediates (depths, raw network output, weights) live in the caller-provided workspace.
#include <algorithm>

#include "model.h"

// Scratch of nb_render_rays_static (end of this file): lives in the library image, so the entry point allocates nothing.
// Every GPU that loads the library holds these bytes until the process exits.  In HOST code the name below is a host-side
// stand-in, not a device address: the host takes its address once per device with cudaGetSymbolAddress.
#define NB_STATIC_SCRATCH_BYTES (16u << 20)
static __device__ __align__(256) unsigned char g_static_scratch[NB_STATIC_SCRATCH_BYTES];

namespace nb {

struct RenderLayout {
  size_t off_zc = 0, off_rawc = 0, off_wc = 0, off_zf = 0, off_rawf = 0, off_scratch = 0, off_draw = 0, off_draw_c = 0, off_mlp_c = 0,
         off_mlp_f = 0, mlp_c_bytes = 0, mlp_f_bytes = 0, total = 0;
};

static RenderLayout render_layout(const nb_model* coarse, const nb_model* fine, int64_t n, const nb_render_cfg& c) {
  RenderLayout L;
  const int S = c.n_samples, I = c.n_importance;
  const int64_t nn = n > 0 ? n : 1;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    size_t o = off;
    off += (size_t)round_up((int64_t)bytes, 256);
    return o;
  };
  L.off_zc = take(sizeof(float) * nn * S);
  L.off_rawc = take(sizeof(float) * nn * S * 4);
  L.off_wc = take(sizeof(float) * nn * S);
  L.off_scratch = take(sizeof(float) * nn * 3);
  if (I > 0) {
    L.off_zf = take(sizeof(float) * nn * (S + I));
    L.off_rawf = take(sizeof(float) * nn * (S + I) * 4);
  }
  const nb_model* mf = fine ? fine : coarse;
  L.mlp_c_bytes = nb_mlp_workspace_bytes(coarse, c.precision, nn * S, c.save_for_backward);
  L.mlp_f_bytes = I > 0 ? nb_mlp_workspace_bytes(mf, c.precision, nn * (S + I), c.save_for_backward) : 0;
  if (c.save_for_backward) {
    L.off_draw = take(sizeof(float) * nn * (S + I) * 4);
    // the coarse pass's d_raw has its own buffer: the two networks' backward passes share nothing, so a caller may run
    // them as two calls on two streams (trainer.py does, inside its captured step)
    L.off_draw_c = I > 0 ? take(sizeof(float) * nn * S * 4) : L.off_draw;
    L.off_mlp_c = take(L.mlp_c_bytes);
    L.off_mlp_f = take(L.mlp_f_bytes);
  } else {
    L.off_mlp_c = L.off_mlp_f = take(std::max(L.mlp_c_bytes, L.mlp_f_bytes));
  }
  L.total = off;
  return L;
}

static inline float* at(void* ws, size_t off) { return reinterpret_cast<float*>(static_cast<char*>(ws) + off); }

// which of the four random inputs come from the in-kernel generator (f4): those the caller asked for (perturb /
// raw_noise_std) without supplying an array
struct RenderRng {
  Rng jitter, fine_u, noise_c, noise_f;
  explicit RenderRng(const nb_render_rand& r) {
    const unsigned long long* st = reinterpret_cast<const unsigned long long*>(r.philox_state);
    const bool jit = st != nullptr && r.perturb != 0, noi = st != nullptr && r.raw_noise_std > 0.0f;
    jitter = Rng{jit && r.t_rand == nullptr ? st : nullptr, RNG_JITTER, 0.0f};
    fine_u = Rng{jit && r.u_fine == nullptr ? st : nullptr, RNG_FINE_U, 0.0f};
    noise_c = Rng{noi && r.noise_c == nullptr ? st : nullptr, RNG_NOISE_C, r.raw_noise_std};
    noise_f = Rng{noi && r.noise_f == nullptr ? st : nullptr, RNG_NOISE_F, r.raw_noise_std};
  }
};

}  // namespace nb

using namespace nb;

#define NB_TRY(x)          \
  do {                     \
    int rc_ = (x);         \
    if (rc_ != NB_OK) return rc_; \
  } while (0)

static int check_cfg(const nb_model_t* coarse, int ray_stride, int64_t n, const nb_render_cfg* cfg) {
  NB_REQUIRE(coarse != nullptr && cfg != nullptr, "nb_render_rays: null model or cfg");
  NB_REQUIRE(n >= 0 && cfg->n_samples >= 2 && cfg->n_importance >= 0, "nb_render_rays: bad sizes");
  NB_REQUIRE(cfg->n_importance == 0 || cfg->n_samples >= 4, "nb_render_rays: hierarchical sampling needs N_samples >= 4");
  NB_REQUIRE(ray_stride == 8 || ray_stride == 11, "nb_render_rays: ray_stride must be 8 or 11");
  return NB_OK;
}

extern "C" {

size_t nb_render_workspace_bytes(const nb_model_t* coarse, const nb_model_t* fine, int ray_stride, int64_t n,
                                 const nb_render_cfg* cfg) {
  if (check_cfg(coarse, ray_stride, n, cfg) != NB_OK) return 0;
  return render_layout(coarse, fine, n, *cfg).total;
}

int nb_render_rays(const nb_model_t* coarse, const nb_model_t* fine, const float* rays, int ray_stride, int64_t n,
                   const nb_render_cfg* cfg, const nb_render_rand* rnd, const nb_render_out* out, void* ws,
                   size_t ws_bytes, nb_stream_t stream) {
  NB_TRY(check_cfg(coarse, ray_stride, n, cfg));
  NB_REQUIRE(out != nullptr, "nb_render_rays: null out");
  if (n == 0) return NB_OK;
  NB_REQUIRE(rays != nullptr, "nb_render_rays: null rays");
  const RenderLayout L = render_layout(coarse, fine, n, *cfg);
  if (ws == nullptr || ws_bytes < L.total) {
    set_error("nb_render_rays: workspace of %zu bytes given, %zu needed", ws_bytes, L.total);
    return NB_ERR_WORKSPACE;
  }
  const nb_render_rand none{nullptr, nullptr, nullptr, nullptr, nullptr, 0, 0.0f};
  const nb_render_rand& r = rnd ? *rnd : none;
  const RenderRng rr(r);
  const int S = cfg->n_samples, I = cfg->n_importance;
  const bool hier = I > 0;
  float* zc = at(ws, L.off_zc);
  float* rawc = at(ws, L.off_rawc);
  float* wc = at(ws, L.off_wc);
  float* scratch = at(ws, L.off_scratch);
  void* mlp_c = static_cast<char*>(ws) + L.off_mlp_c;
  void* mlp_f = static_cast<char*>(ws) + L.off_mlp_f;

  NB_TRY(sample_coarse_impl(rays, ray_stride, n, S, cfg->lindisp, r.t_rand, rr.jitter, zc, stream));
  NB_TRY(nb_mlp_forward(coarse, cfg->precision, rays, ray_stride, zc, n, S, rawc, mlp_c, L.mlp_c_bytes,
                        cfg->save_for_backward, stream));
  // coarse compositing; its weights feed sample_pdf
  float* c_rgb = hier ? (out->rgb0 ? out->rgb0 : scratch) : (out->rgb ? out->rgb : scratch);
  NB_TRY(composite_impl(rawc, zc, rays, ray_stride, n, S, r.noise_c, rr.noise_c, cfg->white_bkgd, c_rgb,
                        hier ? out->disp0 : out->disp, hier ? out->acc0 : out->acc, hier ? out->depth0 : out->depth, wc, stream));
  const float* last_raw = rawc;
  const float* last_z = zc;
  const float* last_w = wc;
  int last_S = S;
  if (hier) {
    float* zf = at(ws, L.off_zf);
    float* rawf = at(ws, L.off_rawf);
    NB_TRY(sample_fine_impl(zc, wc, n, S, r.u_fine, rr.fine_u, I, nullptr, zf, out->z_std, stream));
    const nb_model_t* mf = fine ? fine : coarse;
    NB_TRY(nb_mlp_forward(mf, cfg->precision, rays, ray_stride, zf, n, S + I, rawf, mlp_f, L.mlp_f_bytes,
                          cfg->save_for_backward, stream));
    NB_TRY(composite_impl(rawf, zf, rays, ray_stride, n, S + I, r.noise_f, rr.noise_f, cfg->white_bkgd,
                          out->rgb ? out->rgb : scratch, out->disp, out->acc, out->depth, out->weights, stream));
    last_raw = rawf;
    last_z = zf;
    last_w = nullptr;  // written directly above
    last_S = S + I;
  }
  const cudaStream_t s = as_stream(stream);
  if (out->raw) NB_CUDA(cudaMemcpyAsync(out->raw, last_raw, sizeof(float) * n * last_S * 4, cudaMemcpyDeviceToDevice, s));
  if (out->z_vals) NB_CUDA(cudaMemcpyAsync(out->z_vals, last_z, sizeof(float) * n * last_S, cudaMemcpyDeviceToDevice, s));
  if (out->weights && last_w)
    NB_CUDA(cudaMemcpyAsync(out->weights, last_w, sizeof(float) * n * last_S, cudaMemcpyDeviceToDevice, s));
  return NB_OK;
}

int nb_render_rays_backward(const nb_model_t* coarse, const nb_model_t* fine, const float* rays, int ray_stride,
                            int64_t n, const nb_render_cfg* cfg, const nb_render_rand* rnd,
                            const nb_render_grads* g, void* ws, size_t ws_bytes, float* grad_coarse,
                            float* grad_fine, nb_stream_t stream) {
  NB_TRY(check_cfg(coarse, ray_stride, n, cfg));
  NB_REQUIRE(cfg->save_for_backward, "nb_render_rays_backward: the forward call was not made with save_for_backward");
  NB_REQUIRE(g != nullptr && grad_coarse != nullptr, "nb_render_rays_backward: null grads");
  if (n == 0) return NB_OK;
  const RenderLayout L = render_layout(coarse, fine, n, *cfg);
  if (ws == nullptr || ws_bytes < L.total) {
    set_error("nb_render_rays_backward: workspace of %zu bytes given, %zu needed", ws_bytes, L.total);
    return NB_ERR_WORKSPACE;
  }
  const nb_render_rand none{nullptr, nullptr, nullptr, nullptr, nullptr, 0, 0.0f};
  const nb_render_rand& r = rnd ? *rnd : none;
  const RenderRng rr(r);
  const int S = cfg->n_samples, I = cfg->n_importance;
  const bool hier = I > 0;
  float* draw = at(ws, L.off_draw);
  float* draw_c = at(ws, L.off_draw_c);
  void* mlp_c = static_cast<char*>(ws) + L.off_mlp_c;
  void* mlp_f = static_cast<char*>(ws) + L.off_mlp_f;
  if (hier) {
    const nb_model_t* mf = fine ? fine : coarse;
    float* gf = fine ? grad_fine : grad_coarse;
    NB_REQUIRE(gf != nullptr, "nb_render_rays_backward: null grad_fine");
    if (g->d_rgb || g->d_disp || g->d_acc || g->d_depth) {
      NB_TRY(composite_backward_impl(at(ws, L.off_rawf), at(ws, L.off_zf), rays, ray_stride, n, S + I, r.noise_f,
                                     rr.noise_f, cfg->white_bkgd, g->d_rgb, g->d_disp, g->d_acc, g->d_depth, draw, stream));
      NB_TRY(nb_mlp_backward(mf, cfg->precision, n * (S + I), draw, mlp_f, L.mlp_f_bytes, gf, stream));
    }
    if (g->d_rgb0 || g->d_disp0 || g->d_acc0 || g->d_depth0) {
      NB_TRY(composite_backward_impl(at(ws, L.off_rawc), at(ws, L.off_zc), rays, ray_stride, n, S, r.noise_c,
                                   rr.noise_c, cfg->white_bkgd, g->d_rgb0, g->d_disp0, g->d_acc0, g->d_depth0, draw_c, stream));
      NB_TRY(nb_mlp_backward(coarse, cfg->precision, n * S, draw_c, mlp_c, L.mlp_c_bytes, grad_coarse, stream));
    }
  } else if (g->d_rgb || g->d_disp || g->d_acc || g->d_depth) {
    NB_TRY(composite_backward_impl(at(ws, L.off_rawc), at(ws, L.off_zc), rays, ray_stride, n, S, r.noise_c,
                                 rr.noise_c, cfg->white_bkgd, g->d_rgb, g->d_disp, g->d_acc, g->d_depth, draw, stream));
    NB_TRY(nb_mlp_backward(coarse, cfg->precision, n * S, draw, mlp_c, L.mlp_c_bytes, grad_coarse, stream));
  }
  return NB_OK;
}

size_t nb_render_static_scratch_bytes(void) { return (size_t)NB_STATIC_SCRATCH_BYTES; }

// nb_render_rays for callers that cannot allocate (CUDA graph capture, serving loops): the chunk's intermediates go to
// g_static_scratch instead of a caller workspace.  Inference only.  The call itself makes no cudaMalloc / cudaFree and
// uses no allocating library routine; the models must have been packed beforehand (nb_model_pack, set-up time).
// Every region of the scratch is written by this call before it is read (depths, raw network output, weights), so nothing
// has to start at zero and the previous call's contents do not matter.  All calls on a device share the one array: two
// calls that overlap on the GPU (different streams) race.
int nb_render_rays_static(const nb_model_t* coarse, const nb_model_t* fine, const float* rays, int ray_stride, int64_t n,
                          const nb_render_cfg* cfg, const nb_render_rand* rnd, const nb_render_out* out,
                          nb_stream_t stream) {
  NB_TRY(check_cfg(coarse, ray_stride, n, cfg));
  NB_REQUIRE(!cfg->save_for_backward, "nb_render_rays_static: inference only (save_for_backward must be 0)");
  const size_t need = render_layout(coarse, fine, n, *cfg).total;
  if (need > (size_t)NB_STATIC_SCRATCH_BYTES) {
    set_error("nb_render_rays_static: %lld rays need %zu bytes of scratch, the static array holds %zu (render fewer rays per "
              "call or use nb_render_rays with a workspace)", (long long)n, need, (size_t)NB_STATIC_SCRATCH_BYTES);
    return NB_ERR_WORKSPACE;
  }
  // device address of the array, once per device, with that device current
  constexpr int kMaxDev = 64;
  static void* addr[kMaxDev] = {};
  int dev = -1;
  NB_CUDA(cudaGetDevice(&dev));
  NB_REQUIRE(dev >= 0 && dev < kMaxDev, "nb_render_rays_static: device ordinal %d out of range", dev);
  if (addr[dev] == nullptr) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess || p == nullptr) {
      set_error("nb_render_rays_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
      return NB_ERR_CUDA;
    }
    addr[dev] = p;
  }
  return nb_render_rays(coarse, fine, rays, ray_stride, n, cfg, rnd, out, addr[dev], (size_t)NB_STATIC_SCRATCH_BYTES, stream);
}

}  // extern "C"
