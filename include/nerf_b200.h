/*
 * nerf_b200.h -- C-ABI of the B200-native classic-NeRF hot path.
 *
 * This is the drop-in boundary.  The reference has no native boundary of
 * its own except the un-vendored `torchsearchsorted` extension; every entry point below therefore
 * cites the *Python* reference function whose arithmetic it replaces (paths relative to the
 * reference checkout: "TF" = nerf/..., "PT" = nerf-pytorch/nerf/...).
 *
 * Conventions
 *   - plain C types only: raw DEVICE pointers (unless a parameter is documented as HOST), sizes,
 *     scalars, an opaque model handle and a cudaStream_t passed as void*;
 *   - the caller owns all device memory.  Inputs are borrowed for the duration of the call,
 *     outputs and workspaces are pre-allocated by the caller.  The library owns only the packed
 *     weight copies behind an nb_model_t;
 *   - every call is asynchronous on the given stream and never synchronises the device;
 *   - return value 0 on success, negative nb_status on failure; nb_last_error() returns a
 *     thread-local message.  Nothing throws across the ABI;
 *   - all tensors are fp32, row-major, ray-major (ray = row*W + col) unless stated otherwise.
 *   - there is no CPU fallback: without a CUDA device every compute call fails with NB_ERR_CUDA.
 */
#ifndef NERF_B200_H_
#define NERF_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NB_ABI_VERSION 3 /* 2: nb_render_rand gained philox_state, perturb, raw_noise_std; 3: *_dev pose variants, nb_check_finite, NB_PREC_F16 */

typedef enum nb_status {
  NB_OK = 0,
  NB_ERR_INVALID = -1,      /* bad argument */
  NB_ERR_CUDA = -2,         /* CUDA runtime error (message holds cudaGetErrorString) */
  NB_ERR_UNSUPPORTED = -3,  /* shape / option not supported by this kernel family */
  NB_ERR_WORKSPACE = -4     /* workspace too small */
} nb_status;

/* MLP layer graphs. */
typedef enum nb_graph {
  NB_GRAPH_NERF = 0,     /* TF init_nerf_model: nerf/run_nerf_helpers.py:80-118 (D x W, skip after layer 4) */
  NB_GRAPH_FLEXIBLE = 1  /* PT FlexibleNeRFModel at its 4x128 defaults: nerf-pytorch/nerf/models.py:185-256 */
} nb_graph;

typedef enum nb_precision {
  NB_PREC_FP32 = 0, /* FFMA path: fp32 operands and accumulation, layer by layer */
  NB_PREC_F16 = 1,  /* fused tcgen05 path: IEEE fp16 operands (kind::f16, saturating conversion at +-65504), fp32
                       accumulation in TMEM.  north_star calls this the "bf16 tensor-core path"; fp16 runs at the same
                       tcgen05 rate and is ~5x closer to the fp32 render */
  NB_PREC_BF16 = 1  /* historical alias of NB_PREC_F16 (same value, same kernels) */
} nb_precision;

typedef struct nb_model nb_model_t;
typedef void* nb_stream_t; /* cudaStream_t */

/* ---------------------------------------------------------------- misc */
int nb_abi_version(void);
const char* nb_last_error(void);
/* Number of CUDA devices visible, or a negative nb_status.  Never touches the oracle or a CPU path. */
int nb_device_count(void);
/* Kernels launched by this library (all threads of the process) since the last reset (bench.py: gpu_launches). */
int64_t nb_launch_count(void);
void nb_launch_count_reset(void);

/* Process-wide kernel selection knobs (no effect on results beyond rounding order).  Returns the previous value (>= 0) or
 * NB_ERR_INVALID for an unknown name.
 *   "fwd_2cta": 1 = inference launches of the tensor-core MLP use the CTA-pair kernel (tcgen05 cta_group::2: two SMs share
 *               every weight slot), 0 = the single-CTA kernel.  Initial value: environment NB_FWD_2CTA, else 0.
 *   "fwd_4tile": 1 (default) = inference launches of graphs whose stages are all N <= 128 (FlexibleNeRFModel 4x128) use the
 *               four-tiles-per-CTA kernel, 0 = the two-tile kernel.
 *   "composite_lc": 1 (default) = compositing of 64 / 128 / 192 samples per ray uses the lane-contiguous bulk-copy kernels,
 *               0 = the strided kernels for every sample count.
 *   "sample_fine2": 1 (default) = nb_sample_fine / the chunk renderer with 64 depths, 128 or 64 deterministic samples and
 *               16-byte aligned rows use the two-rays-per-warp kernel, 0 = the one-ray-per-warp kernel (bit-identical).
 *   "bwd_sm_reserve": number of SMs (0..64) the persistent backward kernels of the tensor-core train path leave free, so
 *               that a gradient all-reduce enqueued on another stream can run under them (data-parallel step). */
int nb_set_option(const char* name, int value);

/* ---------------------------------------------------------------- model handle
 * Replaces: TF create_nerf/init_nerf_model (nerf/run_nerf.py:378-404, run_nerf_helpers.py:80-118) and
 * PT FlexibleNeRFModel.__init__ (models.py:186-231) as far as the *weights layout* is concerned.
 * Parameters live in ONE caller-owned flat fp32 buffer: for every layer, in the canonical order
 * reported by nb_model_tensor_info, weight [out, in] row-major (nn.Linear convention) followed by
 * bias [out].  Names equal the reference state_dict keys for FLEXIBLE and
 * pts_linears.i / alpha_linear / bottleneck_linear / views_linears.0 / rgb_linear / output_linear
 * for NERF (Keras layer order of get_weights(), run_nerf.py:835-843).
 */
int nb_model_create(int graph, int depth, int width, int n_freq_xyz, int n_freq_dir, int use_viewdirs,
                    int skip_layer /* NERF: index of the layer after which [pts,h] is concatenated, -1 = none */,
                    nb_model_t** out);
int nb_model_destroy(nb_model_t* m);
int64_t nb_model_param_count(const nb_model_t* m);
int nb_model_num_tensors(const nb_model_t* m);
/* name_cap bytes are written to `name` (NUL terminated).  cols == 0 marks a bias vector of `rows`. */
int nb_model_tensor_info(const nb_model_t* m, int index, char* name, int name_cap, int64_t* rows, int64_t* cols,
                         int64_t* offset_floats);
/* Borrow the flat parameter buffer (device).  Must stay valid until replaced or the model is destroyed. */
int nb_model_set_params(nb_model_t* m, const float* flat_params, int64_t n_floats);
/* (Re)build the kernel-side copies of the weights for `precision` from the flat buffer. */
int nb_model_pack(nb_model_t* m, int precision, nb_stream_t stream);
/* FLOPs (2*MAC, unpadded) of one forward evaluation of one point: SURVEY 8d "algorithmic flops". */
int64_t nb_model_flops_per_point(const nb_model_t* m);

/* output stage of render_path (TF run_nerf.py:358-373, to8b run_nerf_helpers.py:17 =
 * (255 * clip(x, 0, 1)).astype(uint8), i.e. truncation) and the PSNR it prints against gt_imgs
 * (-10 * log10(mean((rgb - gt)^2))): `mse_accum` (device, 1 double, caller-zeroed) receives sum((x - gt)^2) when gt
 * is given; out8 may be null.  n = number of scalars. */
int nb_to8b(const float* x, const float* gt, int64_t n, unsigned char* out8, double* mse_accum, nb_stream_t stream);

/* TF check_numerics (nerf/run_nerf.py:240-241: every output of render_rays is checked for NaN/Inf) as a device-side
 * flag instead of a host synchronisation per tensor: *flag |= bit when x[0..n) holds a non-finite value.  `flag` is a
 * DEVICE int the caller zeroes and reads back once per frame / step. */
int nb_check_finite(const float* x, int64_t n, int* flag, int bit, nb_stream_t stream);

/* ---------------------------------------------------------------- a1 ray generation
 * Replaces get_rays (TF run_nerf_helpers.py:123-130) / get_ray_bundle (PT nerf_helpers.py:67-110).
 * dir_cam = [(col-cx)/fx, -(row-cy)/fy, -1]; rays_d = sum_j dir_cam[j]*c2w[k][j]; rays_o = c2w[:,3].
 * c2w34 is a HOST pointer to 12 floats (3x4 row-major).  Writes rows [row0,row1) of the image into
 * rays_o / rays_d, each [(row1-row0)*W, 3].
 */
int nb_get_rays(int H, int W, float fx, float fy, float cx, float cy, const float* c2w34_host, int row0, int row1,
                float* rays_o, float* rays_d, nb_stream_t stream);
/* Same, with the 3x4 pose read from DEVICE memory when the kernel runs: a CUDA graph captured around the per-frame
 * chain of render_path (TF run_nerf.py:339-375) then renders whatever pose was copied there before the replay. */
int nb_get_rays_dev(int H, int W, float fx, float fy, float cx, float cy, const float* c2w34_dev, int row0, int row1,
                    float* rays_o, float* rays_d, nb_stream_t stream);

/* rays of selected pixels only.  The training loops draw N_rand pixel indices per step and build
 * the whole image's bundle to keep them (TF run_nerf.py:762-774 `coords[select_inds]`, PT train_nerf.py:240-253);
 * this generates just those rays.  pixel_index: DEVICE int64 [n], row-major index row*W + col (the order of
 * `reshape(-1, 3)` on the [H, W, 3] bundle).  Bit-identical to gathering nb_get_rays' output. */
int nb_get_rays_at(int H, int W, float fx, float fy, float cx, float cy, const float* c2w34_host,
                   const int64_t* pixel_index, int64_t n, float* rays_o, float* rays_d, nb_stream_t stream);
int nb_get_rays_at_dev(int H, int W, float fx, float fy, float cx, float cy, const float* c2w34_dev,
                       const int64_t* pixel_index, int64_t n, float* rays_o, float* rays_d, nb_stream_t stream);

/* a3: ndc_rays (TF run_nerf_helpers.py:143-178 == PT nerf_helpers.py:170-197). In-place allowed.
 * focal / near are doubles because the reference evaluates -1/(W/(2*focal)) with Python (double) scalars
 * before the single rounding to fp32. */
int nb_ndc_rays(int H, int W, double focal, double near, const float* rays_o, const float* rays_d, int64_t n,
                float* ndc_o, float* ndc_d, nb_stream_t stream);

/* a2+a3+a4: unit view directions from the PRE-NDC directions, optional NDC, pack [o,d,near,far,(viewdir)]
 * (TF render run_nerf.py:297-325 == PT run_one_iter_of_nerf train_utils.py:141-168).
 * viewdir_src may be NULL (then rays_d is used) -- it differs only for c2w_staticcam.
 * near_arr/far_arr (device, [n]) override the scalars when non-NULL.  rays_out is [n, 8] or [n, 11]. */
int nb_make_ray_batch(const float* rays_o, const float* rays_d, const float* viewdir_src, int64_t n, int use_viewdirs,
                      int ndc, int H, int W, double focal, double near, double far, const float* near_arr,
                      const float* far_arr, float* rays_out, nb_stream_t stream);

/* ---------------------------------------------------------------- a5 stratified depths
 * TF render_rays run_nerf.py:181-199 == PT predict_and_render_radiance train_utils.py:45-65.
 * t_rand ([n,S] uniforms in [0,1)) may be NULL (perturb off). z_vals is [n,S]. */
int nb_sample_coarse(const float* rays, int ray_stride, int64_t n, int n_samples, int lindisp, const float* t_rand,
                     float* z_vals, nb_stream_t stream);

/* ---------------------------------------------------------------- a6 positional encoding
 * Embedder.embed (TF run_nerf_helpers.py:22-56) == positional_encoding (PT nerf_helpers.py:113-157).
 * x [n, 3] -> out [n, out_stride] (first 3+6L columns written; [x, sin(2^0 x), cos(2^0 x), ...]). */
int nb_posenc(const float* x, int64_t n, int n_freqs, int include_input, float* out, int out_stride,
              nb_stream_t stream);

/* ---------------------------------------------------------------- a7/a8 MLP
 * Replaces run_network + the network forward (TF run_nerf.py:30-45 + run_nerf_helpers.py:80-118;
 * PT train_utils.py:8-25 + models.py:233-256).  Points are o + d*z (rays/z form) or explicit.
 * raw is [n_rays*n_samples, 4] = [rgb, sigma] pre-activation.
 * With save_for_backward the workspace keeps what nb_mlp_backward needs (same workspace must be passed). */
size_t nb_mlp_workspace_bytes(const nb_model_t* m, int precision, int64_t n_points, int save_for_backward);
int nb_mlp_forward(const nb_model_t* m, int precision, const float* rays, int ray_stride, const float* z_vals,
                   int64_t n_rays, int n_samples, float* raw, void* workspace, size_t workspace_bytes,
                   int save_for_backward, nb_stream_t stream);
/* Explicit points [n_points,3]; viewdirs [n_points/points_per_dir, 3] or NULL (network_query_fn form). */
int nb_mlp_forward_points(const nb_model_t* m, int precision, const float* pts, const float* viewdirs,
                          int64_t n_points, int points_per_dir, float* raw, void* workspace, size_t workspace_bytes,
                          nb_stream_t stream);
/* Autograd of the network w.r.t. its parameters (TF train():807-826, PT train_nerf.py:286).
 * d_raw [n_points,4] is consumed (overwritten).  flat_grad (same layout as the flat parameters) is
 * ACCUMULATED into.  No gradient is produced for points / rays (nothing trainable upstream). */
int nb_mlp_backward(const nb_model_t* m, int precision, int64_t n_points, float* d_raw, void* workspace,
                    size_t workspace_bytes, float* flat_grad, nb_stream_t stream);
/* Introspection of the tensor-core train workspace (NB_PREC_BF16, save_for_backward): byte offsets, from
 * `workspace`, of the per-tile X record (forward operand tiles: [encx, encd, activations of every stage]) and dZ record
 * (gradient tiles in dgrad-chain order), and their sizes in 16 KB blocks (128 points x 64 fp16, SWIZZLE_128B image)
 * per 128-point tile.  Used by the parity tests to check dgrad / wgrad exactly against the saved activations. */
/* The constant the NB_PREC_F16 path scales its hidden activations by (a power of two; see csrc/tc_common.cuh kActScale):
 * activation blocks of the X record hold scale * h, the encoding blocks (0, 1) true values. */
float nb_mlp_activation_scale(void);
int nb_mlp_record_layout(const nb_model_t* m, int64_t n_points, const void* workspace, int64_t* xrec_offset,
                         int* xrec_blocks, int64_t* drec_offset, int* drec_blocks);

/* The 1-bit ReLU mask record of the same workspace: per 128-point tile and per ACTIVATION block b (X-record block b + 2),
 * 128 rows x 8 bytes; element i of the row's 64 pre-activations is bit 31 - (i & 31) of 32-bit word i >> 5, set where the
 * fp32 pre-activation is positive (its sign bit clear) -- what the dgrad chain multiplies by; a pre-activation too small
 * for the 16-bit X record (< 1e-6) still passes its gradient, as it does in the fp32 reference. */
int nb_mlp_mask_record_layout(const nb_model_t* m, int64_t n_points, const void* workspace, int64_t* mrec_offset,
                              int* mrec_blocks);

/* ---------------------------------------------------------------- a9 compositing
 * raw2outputs (TF run_nerf.py:93-163) == volume_render_radiance_field (PT volume_rendering_utils.py:6-53).
 * noise ([n,S], already multiplied by raw_noise_std) may be NULL.  depth / weights may be NULL. */
int nb_composite(const float* raw, const float* z_vals, const float* rays, int ray_stride, int64_t n, int n_samples,
                 const float* noise, int white_bkgd, float* rgb, float* disp, float* acc, float* depth,
                 float* weights, nb_stream_t stream);
/* Backward of the above w.r.t. raw.  Any of d_rgb [n,3], d_disp, d_acc, d_depth [n] may be NULL (= zero). */
int nb_composite_backward(const float* raw, const float* z_vals, const float* rays, int ray_stride, int64_t n,
                          int n_samples, const float* noise, int white_bkgd, const float* d_rgb,
                          const float* d_disp, const float* d_acc, const float* d_depth, float* d_raw,
                          nb_stream_t stream);

/* ---------------------------------------------------------------- a10/a11 hierarchical sampling
 * The reference's only native boundary: torchsearchsorted.searchsorted(a[B,n], v[B,m], side) -> int64[B,m]
 * (PT nerf_helpers.py:241-243,286-288; un-vendored github.com/aliutkus/torchsearchsorted, unpinned).
 * side_right != 0 counts entries <= v (numpy side='right').  a_rows == 1 broadcasts a over the rows of v. */
int nb_searchsorted(const float* a, int64_t a_rows, int n, const float* v, int64_t v_rows, int m, int side_right,
                    int64_t* out, nb_stream_t stream);
/* sample_pdf (TF run_nerf_helpers.py:183-211) == sample_pdf_2 (PT nerf_helpers.py:260-302).
 * bins [n, nb], weights [n, nb-1] with row strides; u ([n, ns]) NULL => deterministic linspace(0,1,ns).
 * samples [n, ns].  cdf_out ([n, nb], optional) receives the fp32 CDF actually used. */
int nb_sample_pdf(const float* bins, int bins_stride, const float* weights, int weights_stride, int64_t n, int nb,
                  const float* u, int ns, float* samples, float* cdf_out, nb_stream_t stream);
/* The fused form used by the renderer: from the coarse z_vals [n,S] and coarse weights [n,S] produce
 * z_samples [n,ns] (may be NULL) and the merged, sorted z_fine [n,S+ns]
 * (TF run_nerf.py:213-221 == PT train_utils.py:93-105: mids, weights[1:-1], sample_pdf, sort(cat)).
 * z_std ([n], optional) = population std of z_samples (TF run_nerf.py:238). */
int nb_sample_fine(const float* z_vals, const float* weights, int64_t n, int n_samples, const float* u, int ns,
                   float* z_samples, float* z_fine, float* z_std, nb_stream_t stream);
/* Ascending merge of two row-wise sorted arrays (torch.sort(cat) / tf.sort(concat), values only). */
int nb_merge_sorted(const float* a, int na, const float* b, int nb, int64_t n, float* out, nb_stream_t stream);

/* ---------------------------------------------------------------- a5..a12 per-chunk renderer
 * The whole of render_rays (TF nerf/run_nerf.py:48-243) == predict_and_render_radiance
 * (PT nerf-pytorch/nerf/train_utils.py:28-127) for one chunk of packed rays, as ONE call:
 * stratified depths -> coarse MLP -> compositing -> sample_pdf + sorted merge -> fine MLP -> compositing.
 * This is the entry point the reference-facing Python API (render / run_one_iter_of_nerf) sits on. */
typedef struct nb_render_cfg {
  int n_samples;         /* N_samples / num_coarse */
  int n_importance;      /* N_importance / num_fine; 0 = coarse pass only */
  int lindisp;
  int white_bkgd;
  int precision;         /* nb_precision of both MLP passes */
  int save_for_backward; /* keep what nb_render_rays_backward needs in the workspace */
} nb_render_cfg;

/* Random inputs (device, all optional): the reference draws them from the framework RNG, so parity is
 * defined only when they are NULL (perturb = 0, raw_noise_std = 0) or supplied by the test. */
typedef struct nb_render_rand {
  const float* t_rand;  /* [n, n_samples] uniforms for the stratified jitter (run_nerf.py:192-199) */
  const float* u_fine;  /* [n, n_importance] uniforms for sample_pdf (run_nerf_helpers.py:192-197) */
  const float* noise_c; /* [n, n_samples] sigma noise, already scaled by raw_noise_std (run_nerf.py:130-132) */
  const float* noise_f; /* [n, n_samples + n_importance] */
  /* in-kernel generator.  With philox_state != NULL (device, {seed, step}), every input above that is
   * NULL but asked for -- the jitter and the inverse-CDF uniforms when perturb != 0, the sigma noise when
   * raw_noise_std > 0 -- is drawn inside the sampling / compositing kernels (Philox4x32-10, counter = ray, sample,
   * input, step), statistically (not bitwise) matched to the reference's framework generator.  The backward call must
   * see the same state as its forward; nb_rng_advance moves to the next step. */
  const uint64_t* philox_state;
  int perturb;
  float raw_noise_std;
} nb_render_rand;

/* f4: state[1] += 1 on the stream (graph-capturable: a captured train step draws fresh numbers on every replay). */
int nb_rng_advance(uint64_t* state, nb_stream_t stream);
/* f4: the numbers the kernels draw, as an array: out[r, s] for r < n, s < row_len; input = 0 jitter, 1 inverse-CDF
 * uniforms, 2 coarse noise, 3 fine noise; normal != 0 -> scale * N(0,1), else U[0,1).  Feeding these arrays back
 * through t_rand / u_fine / noise_* reproduces the in-kernel result bit for bit (tests/test_gpu_rng.py). */
int nb_rng_fill(const uint64_t* state, int input, int normal, float scale, int64_t n, int row_len, float* out,
                nb_stream_t stream);

/* Outputs (device, any may be NULL).  With n_importance == 0 the coarse result is written to
 * rgb/disp/acc/depth and the *0 pointers are ignored. */
typedef struct nb_render_out {
  float *rgb, *disp, *acc, *depth;     /* [n,3] [n] [n] [n]   fine pass  (rgb_map, disp_map, acc_map, depth_map) */
  float *rgb0, *disp0, *acc0, *depth0; /* coarse pass (extras rgb0 / disp0 / acc0) */
  float* z_std;                        /* [n] std of the importance samples (run_nerf.py:238) */
  float* raw;                          /* [n, S(+I), 4] last-pass raw network output (retraw) */
  float* z_vals;                       /* [n, S(+I)] last-pass depths */
  float* weights;                      /* [n, S(+I)] last-pass compositing weights */
} nb_render_out;

/* Upstream gradients for nb_render_rays_backward (device, any may be NULL = zero). */
typedef struct nb_render_grads {
  const float *d_rgb, *d_disp, *d_acc, *d_depth;
  const float *d_rgb0, *d_disp0, *d_acc0, *d_depth0;
} nb_render_grads;

size_t nb_render_workspace_bytes(const nb_model_t* coarse, const nb_model_t* fine, int ray_stride, int64_t n,
                                 const nb_render_cfg* cfg);
/* `fine` may be NULL (the coarse network is then reused for the importance pass, run_nerf.py:226). */
int nb_render_rays(const nb_model_t* coarse, const nb_model_t* fine, const float* rays, int ray_stride, int64_t n,
                   const nb_render_cfg* cfg, const nb_render_rand* rnd, const nb_render_out* out, void* workspace,
                   size_t workspace_bytes, nb_stream_t stream);
/* nb_render_rays without a caller workspace and without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  The chunk's intermediates live in a file-scope `static __device__` array of
 * nb_render_static_scratch_bytes() = 16 MiB inside the library (held by every GPU that loads it until the process exits);
 * about 3 000 rays at 64 + 128 samples on the 16-bit path.  Inference only (cfg->save_for_backward must be 0); a call
 * whose intermediates exceed the array fails with NB_ERR_WORKSPACE and a message stating both sizes.  The call makes
 * no cudaMalloc / cudaFree and uses no allocating library routine; models must have been packed before (nb_model_pack).
 * All calls on a device share the array, which keeps the last call's contents (every region is written before it is
 * read, nothing needs clearing): calls that overlap on the GPU (different streams) RACE -- serialise them.
 * Results equal nb_render_rays on the same inputs bit for bit (same kernels, same order). */
size_t nb_render_static_scratch_bytes(void);
int nb_render_rays_static(const nb_model_t* coarse, const nb_model_t* fine, const float* rays, int ray_stride, int64_t n,
                          const nb_render_cfg* cfg, const nb_render_rand* rnd, const nb_render_out* out,
                          nb_stream_t stream);
/* Autograd of the call above w.r.t. the network parameters (TF train():807-826 / PT train_nerf.py:286):
 * compositing backward and MLP backward for the fine and the coarse pass; no gradient flows through
 * sample_pdf (stop_gradient, run_nerf.py:218 / detach, train_utils.py:103).  Same rays / cfg / rnd / workspace
 * as the forward call made with save_for_backward.  grad_coarse / grad_fine are flat buffers in the
 * parameter layout and are ACCUMULATED into (grad_fine may alias grad_coarse when fine == NULL). */
int nb_render_rays_backward(const nb_model_t* coarse, const nb_model_t* fine, const float* rays, int ray_stride,
                            int64_t n, const nb_render_cfg* cfg, const nb_render_rand* rnd,
                            const nb_render_grads* grads, void* workspace, size_t workspace_bytes,
                            float* grad_coarse, float* grad_fine, nb_stream_t stream);

/* ---------------------------------------------------------------- a13 loss
 * img2mse (TF run_nerf_helpers.py:11, PT nerf_helpers.py:9): loss_out[0] += mean((pred-target)^2) and
 * d_pred = grad_scale * 2*(pred-target)/count.  loss_out must be zeroed by the caller. */
int nb_mse_loss(const float* pred, const float* target, int64_t count, float grad_scale, float* loss_out,
                float* d_pred, nb_stream_t stream);

/* ---------------------------------------------------------------- a13 optimizer step
 * Adam over a flat fp32 parameter buffer, one launch, gradients consumed and zeroed for the next step.
 * Replaces optimizer.step() + zero_grad() (PT train_nerf.py:288-289, torch.optim.Adam: denom = sqrt(v)/sqrt(bc2)+eps,
 * p -= lr/bc1 * m/denom) or, with tf_style != 0, tf.train.AdamOptimizer.apply_gradients (TF run_nerf.py:827:
 * lr_t = lr*sqrt(bc2)/bc1, p -= lr_t*m/(sqrt(v)+eps)).
 * hyper is a DEVICE array of 8 floats {lr, beta1, beta2, eps, bc1 = 1-beta1^t, bc2 = 1-beta2^t, 1-beta1, 1-beta2}
 * (each formed in double precision on the host, like the frameworks do), so that a captured
 * CUDA graph of the whole train step sees the new learning rate / step count on every replay. */
int nb_adam_step(float* params, float* grads, float* exp_avg, float* exp_avg_sq, int64_t n, const float* hyper,
                 int tf_style, int zero_grads, nb_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* NERF_B200_H_ */
