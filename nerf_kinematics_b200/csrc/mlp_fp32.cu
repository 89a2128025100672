// fp32 MLP path (NB_PREC_FP32): positional encoding + layer-by-layer FFMA GEMMs, forward and backward.
//
// This is the accuracy path of the boundary (rendered rgb/depth/acc within 1e-3 of the reference,
// SURVEY.md 8c): fp32 operands, fp32 accumulation, activations kept in a caller-provided workspace.
// The throughput path is the fused tcgen05 kernel in mlp_tc.cu.  What is computed follows
// run_network + the network forward (nerf/run_nerf.py:30-45, nerf/run_nerf_helpers.py:22-56,80-118;
// nerf-pytorch/nerf/train_utils.py:8-25, nerf_helpers.py:113-157, models.py:233-256) and its autograd.
#include <math.h>

#include <algorithm>

#include "model.h"

namespace nb {

// implemented in mlp_tc.cu
int tc_forward(const nb_model* m, const float* rays, int ray_stride, const float* z_vals, const float* pts,
               const float* viewdirs, int points_per_dir, int64_t n_rays, int n_samples, int64_t n_points, float* raw,
               void* workspace, size_t workspace_bytes, int save_for_backward, cudaStream_t stream);
int tc_backward(const nb_model* m, int64_t n_points, float* d_raw, void* workspace, size_t workspace_bytes,
                float* flat_grad, cudaStream_t stream);
size_t tc_workspace_bytes(const nb_model* m, int64_t n_points, int save_for_backward);

// ------------------------------------------------------------------------------------ encoding
// One thread per (point, channel): coalesced stores.  Points come either from rays + depths
// (pts = o + d*z, evaluated as a separate multiply and add like the reference's broadcast expression
// train_utils.py:67 / run_nerf.py:202) or from an explicit [n,3] array.
struct EncodeSrc {
  const float* rays;  // [n_rays, stride] or null
  int stride;
  const float* z;     // [n_rays, S]
  int S;
  const float* pts;       // [n_points, 3] when rays == null
  const float* viewdirs;  // [n_points / per_dir, 3] when rays == null (may be null)
  int per_dir;
};

__device__ __forceinline__ float enc_channel(float x0, float x1, float x2, int c) {
  // channel order [x(3), sin(2^0 x)(3), cos(2^0 x)(3), sin(2^1 x), ...]  (nerf_helpers.py:149-157)
  if (c < 3) return c == 0 ? x0 : (c == 1 ? x1 : x2);
  const int cc = c - 3;
  const int f = cc / 6, r = cc - 6 * f;
  const int comp = r >= 3 ? r - 3 : r;
  const float x = comp == 0 ? x0 : (comp == 1 ? x1 : x2);
  const float a = __fmul_rn(x, exp2f((float)f));
  return r < 3 ? sinf(a) : cosf(a);
}

__global__ void encode_kernel(EncodeSrc src, int64_t p0, int64_t np, int Cx, int Cd, float* __restrict__ EX, int ldx,
                              float* __restrict__ ED, int ldd) {
  const int C = Cx + Cd;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= np * C) return;
  const int64_t pl = idx / C;  // point within this chunk
  const int c = (int)(idx - pl * C);
  const int64_t p = p0 + pl;
  float x0, x1, x2;
  if (c < Cx) {
    if (src.rays != nullptr) {
      const int64_t r = p / src.S;
      const float* ry = src.rays + r * src.stride;
      const float zz = src.z[p];
      x0 = __fadd_rn(ry[0], __fmul_rn(ry[3], zz));
      x1 = __fadd_rn(ry[1], __fmul_rn(ry[4], zz));
      x2 = __fadd_rn(ry[2], __fmul_rn(ry[5], zz));
    } else {
      x0 = src.pts[3 * p]; x1 = src.pts[3 * p + 1]; x2 = src.pts[3 * p + 2];
    }
    EX[pl * ldx + c] = enc_channel(x0, x1, x2, c);
  } else {
    const float* v;
    if (src.rays != nullptr) v = src.rays + (p / src.S) * src.stride + 8;
    else v = src.viewdirs + (p / src.per_dir) * 3;
    ED[pl * ldd + (c - Cx)] = enc_channel(v[0], v[1], v[2], c - Cx);
  }
}

// ------------------------------------------------------------------------------------ SGEMM
// C[i][j] (op)= epilogue( sum_c A(i,c) * B(j,c) ).  AMODE/BMODE 0: the contraction index is the
// contiguous one (A(i,c) = A[i*lda + c]); 1: the row index is contiguous (A(i,c) = A[c*lda + i]).
struct Gemm {
  const float* A; int64_t lda;
  const float* B; int64_t ldb;
  float* C; int64_t ldc;
  int64_t M; int N; int64_t K;
  const float* bias;          // [N] or null
  const float* mask; int64_t ldm;  // multiply by (mask[i][j] > 0) when non-null
  int relu, accumulate;
  int64_t k_split;            // contraction elements per blockIdx.z
  int64_t c_split;            // C elements between the partial results of consecutive blockIdx.z (split-K), else 0
};

constexpr int BM = 128, BN = 128, BK = 16, PAD = 4;

template <int AMODE, int BMODE>
__global__ void __launch_bounds__(256) sgemm_kernel(Gemm g) {
  __shared__ __align__(16) float As[BK][BM + PAD];
  __shared__ __align__(16) float Bs[BK][BN + PAD];
  const int t = threadIdx.x;
  const int tx = t & 15, ty = t >> 4;
  const int64_t i0 = (int64_t)blockIdx.x * BM;
  const int j0 = blockIdx.y * BN;
  const int64_t kbeg = (int64_t)blockIdx.z * g.k_split;
  const int64_t kend = kbeg + g.k_split < g.K ? kbeg + g.k_split : g.K;
  float acc[8][8];
#pragma unroll
  for (int a = 0; a < 8; ++a)
#pragma unroll
    for (int b = 0; b < 8; ++b) acc[a][b] = 0.f;

  for (int64_t k0 = kbeg; k0 < kend; k0 += BK) {
    // ---- stage A tile
    if (AMODE == 0) {
      const int c = t & 15;
      const int64_t kc = k0 + c;
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        const int i = (t >> 4) + 16 * r;
        const int64_t gi = i0 + i;
        As[c][i] = (gi < g.M && kc < kend) ? g.A[gi * g.lda + kc] : 0.f;
      }
    } else {
      const int i = t & 127;
      const int64_t gi = i0 + i;
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        const int c = (t >> 7) + 2 * r;
        const int64_t kc = k0 + c;
        As[c][i] = (gi < g.M && kc < kend) ? g.A[kc * g.lda + gi] : 0.f;
      }
    }
    // ---- stage B tile
    if (BMODE == 0) {
      const int c = t & 15;
      const int64_t kc = k0 + c;
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        const int j = (t >> 4) + 16 * r;
        const int gj = j0 + j;
        Bs[c][j] = (gj < g.N && kc < kend) ? g.B[(int64_t)gj * g.ldb + kc] : 0.f;
      }
    } else {
      const int j = t & 127;
      const int gj = j0 + j;
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        const int c = (t >> 7) + 2 * r;
        const int64_t kc = k0 + c;
        Bs[c][j] = (gj < g.N && kc < kend) ? g.B[kc * g.ldb + gj] : 0.f;
      }
    }
    __syncthreads();
#pragma unroll
    for (int c = 0; c < BK; ++c) {
      const float4 a0 = *reinterpret_cast<const float4*>(&As[c][ty * 4]);
      const float4 a1 = *reinterpret_cast<const float4*>(&As[c][64 + ty * 4]);
      const float4 b0 = *reinterpret_cast<const float4*>(&Bs[c][tx * 4]);
      const float4 b1 = *reinterpret_cast<const float4*>(&Bs[c][64 + tx * 4]);
      const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      const float bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
      for (int a = 0; a < 8; ++a)
#pragma unroll
        for (int b = 0; b < 8; ++b) acc[a][b] = fmaf(av[a], bv[b], acc[a][b]);
    }
    __syncthreads();
  }
  // ---- epilogue
#pragma unroll
  for (int a = 0; a < 8; ++a) {
    const int64_t gi = i0 + (a < 4 ? ty * 4 + a : 64 + ty * 4 + (a - 4));
    if (gi >= g.M) continue;
#pragma unroll
    for (int b = 0; b < 8; ++b) {
      const int gj = j0 + (b < 4 ? tx * 4 + b : 64 + tx * 4 + (b - 4));
      if (gj >= g.N) continue;
      float v = acc[a][b];
      float* cp = g.C + (int64_t)blockIdx.z * g.c_split + gi * g.ldc + gj;
      if (g.accumulate) v += *cp;
      if (g.bias != nullptr) v += g.bias[gj];
      if (g.relu) v = fmaxf(v, 0.f);
      if (g.mask != nullptr && !(g.mask[gi * g.ldm + gj] > 0.f)) v = 0.f;
      *cp = v;
    }
  }
}

static int launch_gemm(int amode, int bmode, Gemm g, int splits, cudaStream_t s) {
  if (g.M <= 0 || g.N <= 0) return NB_OK;
  if (splits < 1) splits = 1;
  g.k_split = round_up(ceil_div(g.K, splits), BK);
  const int z = (int)ceil_div(g.K, g.k_split);
  if (z > 1 && g.c_split == 0) {
    set_error("launch_gemm: split-K needs a partial-result buffer");
    return NB_ERR_INVALID;
  }
  dim3 grid((unsigned)ceil_div(g.M, BM), (unsigned)ceil_div(g.N, BN), (unsigned)z);
  if (amode == 0 && bmode == 0) sgemm_kernel<0, 0><<<grid, 256, 0, s>>>(g);
  else if (amode == 0 && bmode == 1) sgemm_kernel<0, 1><<<grid, 256, 0, s>>>(g);
  else if (amode == 1 && bmode == 1) sgemm_kernel<1, 1><<<grid, 256, 0, s>>>(g);
  else sgemm_kernel<1, 0><<<grid, 256, 0, s>>>(g);
  NB_LAUNCHED();
  return NB_OK;
}

// dst[i][j] += part[0][i][j] + part[1][i][j] + ... in that order: the split-K partial results of a weight gradient (or
// the per-block partial column sums of a bias gradient) are summed by one thread per element, so a gradient is the same
// bits on every run (atomics would make the order a race)
__global__ void ordered_sum_kernel(const float* __restrict__ part, int n_parts, int64_t part_stride, int rows, int cols,
                                   float* __restrict__ dst, int64_t ld) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= (int64_t)rows * cols) return;
  float v = part[e];
  for (int q = 1; q < n_parts; ++q) v += part[q * part_stride + e];
  dst[(e / cols) * ld + (e % cols)] += v;
}

static int launch_ordered_sum(const float* part, int n_parts, int64_t part_stride, int rows, int cols, float* dst,
                              int64_t ld, cudaStream_t s) {
  const int64_t n = (int64_t)rows * cols;
  ordered_sum_kernel<<<(unsigned)ceil_div(n, 256), 256, 0, s>>>(part, n_parts, part_stride, rows, cols, dst, ld);
  NB_LAUNCHED();
  return NB_OK;
}

constexpr int64_t kColsumRows = 2048;  // rows per block of colsum_kernel

// part[blockIdx.x][j] = sum of dY[i][j] over the block's rows; db[j] += the partials in block order (ordered_sum_kernel)
__global__ void colsum_kernel(const float* __restrict__ dY, int64_t ld, int64_t M, int N, int64_t rows_per_block,
                              float* __restrict__ part) {
  const int j = blockIdx.y * 32 + (threadIdx.x & 31);
  const int sub = threadIdx.x >> 5;  // 8 row lanes
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
  const int64_t r1 = r0 + rows_per_block < M ? r0 + rows_per_block : M;
  float acc = 0.f;
  if (j < N)
    for (int64_t i = r0 + sub; i < r1; i += 8) acc += dY[i * ld + j];
  __shared__ float lanes[8][33];
  lanes[sub][threadIdx.x & 31] = acc;
  __syncthreads();
  if (sub == 0 && j < N) {
    float v = 0.f;
#pragma unroll
    for (int q = 0; q < 8; ++q) v += lanes[q][threadIdx.x & 31];
    part[(int64_t)blockIdx.x * N + j] = v;
  }
}

static int launch_colsum(const float* dY, int64_t ld, int64_t M, int N, float* db, float* part, cudaStream_t s) {
  if (M <= 0) return NB_OK;
  const int blocks = (int)ceil_div(M, kColsumRows);
  dim3 grid((unsigned)blocks, (unsigned)ceil_div(N, 32));
  colsum_kernel<<<grid, 256, 0, s>>>(dY, ld, M, N, kColsumRows, part);
  NB_LAUNCHED();
  return launch_ordered_sum(part, blocks, N, 1, N, db, N, s);
}

__global__ void copy_cols_kernel(const float* __restrict__ src, int64_t lds, float* __restrict__ dst, int64_t ldd,
                                 int64_t M, int N) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= M * N) return;
  const int64_t i = idx / N;
  const int j = (int)(idx - i * N);
  dst[i * ldd + j] = src[i * lds + j];
}

// ------------------------------------------------------------------------------------ workspace
struct Fp32Layout {
  int64_t P = 0;        // points the activation buffers are sized for
  bool save = false;
  int ldx = 0, ldd = 0;
  size_t off_ex = 0, off_ed = 0, off_h = 0, off_feat = 0, off_dirh = 0, off_g0 = 0, off_g1 = 0, off_gfeat = 0,
         off_gdir = 0, off_part = 0, total = 0;
  int n_h = 0;
};

constexpr int64_t kFwdChunk = 1 << 18;  // points per internal pass when nothing is saved

static Fp32Layout fp32_layout(const Graph& g, int64_t n_points, bool save) {
  Fp32Layout L;
  L.save = save;
  L.P = save ? n_points : std::min<int64_t>(n_points, kFwdChunk);
  if (L.P < 1) L.P = 1;
  L.ldx = g.Cx;
  L.ldd = g.Cd > 0 ? g.Cd : 1;
  L.n_h = save ? g.depth : 2;
  size_t off = 0;
  auto take = [&](size_t floats) {
    size_t o = off;
    off += round_up((int64_t)(floats * sizeof(float)), 256);
    return o;
  };
  L.off_ex = take((size_t)L.P * L.ldx);
  L.off_ed = take((size_t)L.P * L.ldd);
  L.off_h = take((size_t)L.P * g.width * L.n_h);
  L.off_feat = take((size_t)L.P * std::max(g.feat_width, 1));
  L.off_dirh = take((size_t)L.P * std::max(g.dir_width, 1));
  if (save) {
    L.off_g0 = take((size_t)L.P * g.width);
    L.off_g1 = take((size_t)L.P * g.width);
    L.off_gfeat = take((size_t)L.P * std::max(g.feat_width, 1));
    L.off_gdir = take((size_t)L.P * std::max(g.dir_width, 1));
    // partial results of the gradient reductions (split-K weight gradients: splits x tiles <= 2 x SMs tiles of BM x BN;
    // bias gradients: one row per 2048 points), summed in a fixed order by ordered_sum_kernel
    L.off_part = take(std::max<size_t>((size_t)2 * sm_count() * BM * BN,
                                       (size_t)ceil_div(L.P, kColsumRows) *
                                           std::max(std::max(g.width, g.feat_width), std::max(g.dir_width, 4))));
  }
  L.total = off;
  return L;
}

static inline float* wsf(void* ws, size_t off) { return reinterpret_cast<float*>(static_cast<char*>(ws) + off); }

static bool layer_has_relu(const Graph& g, int i) { return !(i == 0 && !g.act0); }

// ------------------------------------------------------------------------------------ forward
static int fp32_forward_chunk(const nb_model* m, const Fp32Layout& L, void* ws, EncodeSrc src, int64_t p0, int64_t np,
                              float* raw, cudaStream_t s) {
  const Graph& g = m->g;
  const float* W = m->params;
  float* EX = wsf(ws, L.off_ex);
  float* ED = wsf(ws, L.off_ed);
  float* H = wsf(ws, L.off_h);
  float* FEAT = wsf(ws, L.off_feat);
  float* DIRH = wsf(ws, L.off_dirh);
  {
    const int64_t total = np * (g.Cx + g.Cd);
    encode_kernel<<<(unsigned)ceil_div(total, 256), 256, 0, s>>>(src, p0, np, g.Cx, g.Cd, EX, L.ldx, ED, L.ldd);
    NB_LAUNCHED();
  }
  auto hbuf = [&](int i) { return H + (size_t)(L.save ? i : (i & 1)) * L.P * g.width; };
  int rc;
  for (int i = 0; i < g.depth; ++i) {
    const Lin& l = g.lins[g.trunk[i]];
    Gemm q{};
    q.C = hbuf(i); q.ldc = g.width; q.M = np; q.N = l.out;
    q.bias = W + l.b_off;
    q.relu = layer_has_relu(g, i);
    if (i == 0) {
      q.A = EX; q.lda = L.ldx; q.B = W + l.w_off; q.ldb = l.in; q.K = g.Cx;
      if ((rc = launch_gemm(0, 0, q, 1, s)) != NB_OK) return rc;
    } else if (i - 1 == g.skip) {
      // input = concat([pts encoding, h]) (TF order, run_nerf_helpers.py:99-100): two passes over K
      Gemm q1 = q;
      q1.A = EX; q1.lda = L.ldx; q1.B = W + l.w_off; q1.ldb = l.in; q1.K = g.Cx; q1.bias = nullptr; q1.relu = 0;
      if ((rc = launch_gemm(0, 0, q1, 1, s)) != NB_OK) return rc;
      q.A = hbuf(i - 1); q.lda = g.width; q.B = W + l.w_off + g.Cx; q.ldb = l.in; q.K = g.width; q.accumulate = 1;
      if ((rc = launch_gemm(0, 0, q, 1, s)) != NB_OK) return rc;
    } else {
      q.A = hbuf(i - 1); q.lda = g.width; q.B = W + l.w_off; q.ldb = l.in; q.K = g.width;
      if ((rc = launch_gemm(0, 0, q, 1, s)) != NB_OK) return rc;
    }
  }
  const float* HL = hbuf(g.depth - 1);
  float* rawc = raw + p0 * 4;
  if (g.use_viewdirs) {
    const Lin& la = g.lins[g.alpha];
    const Lin& lf = g.lins[g.feat];
    const Lin& ld = g.lins[g.dir];
    const Lin& lr = g.lins[g.rgb];
    Gemm q{};
    // alpha -> raw[:, 3]
    q.A = HL; q.lda = g.width; q.B = W + la.w_off; q.ldb = la.in; q.K = la.in; q.C = rawc + 3; q.ldc = 4; q.M = np; q.N = 1;
    q.bias = W + la.b_off;
    if ((rc = launch_gemm(0, 0, q, 1, s)) != NB_OK) return rc;
    // feat / bottleneck
    q = Gemm{};
    q.A = HL; q.lda = g.width; q.B = W + lf.w_off; q.ldb = lf.in; q.K = lf.in; q.C = FEAT; q.ldc = g.feat_width; q.M = np;
    q.N = lf.out; q.bias = W + lf.b_off; q.relu = g.feat_act;
    if ((rc = launch_gemm(0, 0, q, 1, s)) != NB_OK) return rc;
    // dir layer on concat([feat, dir encoding])
    q = Gemm{};
    q.A = FEAT; q.lda = g.feat_width; q.B = W + ld.w_off; q.ldb = ld.in; q.K = g.feat_width; q.C = DIRH; q.ldc = g.dir_width;
    q.M = np; q.N = ld.out;
    if ((rc = launch_gemm(0, 0, q, 1, s)) != NB_OK) return rc;
    q.A = ED; q.lda = L.ldd; q.B = W + ld.w_off + g.feat_width; q.K = g.Cd; q.accumulate = 1; q.bias = W + ld.b_off; q.relu = 1;
    if ((rc = launch_gemm(0, 0, q, 1, s)) != NB_OK) return rc;
    // rgb -> raw[:, 0:3]
    q = Gemm{};
    q.A = DIRH; q.lda = g.dir_width; q.B = W + lr.w_off; q.ldb = lr.in; q.K = lr.in; q.C = rawc; q.ldc = 4; q.M = np; q.N = 3;
    q.bias = W + lr.b_off;
    if ((rc = launch_gemm(0, 0, q, 1, s)) != NB_OK) return rc;
  } else {
    const Lin& lo = g.lins[g.out4];
    Gemm q{};
    q.A = HL; q.lda = g.width; q.B = W + lo.w_off; q.ldb = lo.in; q.K = lo.in; q.C = rawc; q.ldc = 4; q.M = np; q.N = 4;
    q.bias = W + lo.b_off;
    if ((rc = launch_gemm(0, 0, q, 1, s)) != NB_OK) return rc;
  }
  return NB_OK;
}

static int fp32_forward(const nb_model* m, EncodeSrc src, int64_t n_points, float* raw, void* ws, size_t ws_bytes,
                        int save, cudaStream_t s) {
  const Graph& g = m->g;
  if (g.skip >= g.depth - 1 && g.skip >= 0) {
    set_error("nb_mlp_forward: a skip connection after the last trunk layer is not supported");
    return NB_ERR_UNSUPPORTED;
  }
  const Fp32Layout L = fp32_layout(g, n_points, save != 0);
  if (ws == nullptr || ws_bytes < L.total) {
    set_error("nb_mlp_forward: workspace of %zu bytes given, %zu needed", ws_bytes, L.total);
    return NB_ERR_WORKSPACE;
  }
  for (int64_t p0 = 0; p0 < n_points; p0 += L.P) {
    const int64_t np = std::min<int64_t>(L.P, n_points - p0);
    int rc = fp32_forward_chunk(m, L, ws, src, p0, np, raw, s);
    if (rc != NB_OK) return rc;
  }
  return NB_OK;
}

// ------------------------------------------------------------------------------------ backward
// wgrad: dW[out, in_off : in_off+Kin] += dY[P, out]^T . X[P, Kin]   (contraction over the points)
static int wgrad(const float* dY, int64_t ldy, int out, const float* X, int64_t ldx, int Kin, float* dW, int64_t ldw,
                 int64_t P, float* part, cudaStream_t s) {
  Gemm q{};
  q.A = dY; q.lda = ldy; q.B = X; q.ldb = ldx; q.C = part; q.ldc = Kin; q.M = out; q.N = Kin; q.K = P;
  q.c_split = (int64_t)out * Kin;
  const int tiles = (int)(ceil_div(out, BM) * ceil_div(Kin, BN));
  int splits = (int)std::min<int64_t>(std::max<int64_t>(1, (2 * sm_count()) / tiles), ceil_div(P, 512));
  const int64_t k_split = round_up(ceil_div(P, splits), BK);
  const int z = (int)ceil_div(P, k_split);
  int rc = launch_gemm(1, 1, q, splits, s);
  if (rc != NB_OK) return rc;
  return launch_ordered_sum(part, z, q.c_split, out, Kin, dW, ldw, s);
}

// dgrad: dX[P, Kin] (=|+=) (dY[P, out] . W[out, in_off : in_off+Kin]) * (mask > 0)
static int dgrad(const float* dY, int64_t ldy, int out, const float* Wp, int64_t ldw, int Kin, float* dX, int64_t ldxo,
                 int64_t P, const float* mask, int64_t ldm, int accumulate, cudaStream_t s) {
  Gemm q{};
  q.A = dY; q.lda = ldy; q.B = Wp; q.ldb = ldw; q.C = dX; q.ldc = ldxo; q.M = P; q.N = Kin; q.K = out;
  q.mask = mask; q.ldm = ldm; q.accumulate = accumulate;
  return launch_gemm(0, 1, q, 1, s);
}

static int fp32_backward(const nb_model* m, int64_t P, float* d_raw, void* ws, size_t ws_bytes, float* grad,
                         cudaStream_t s) {
  const Graph& g = m->g;
  const float* W = m->params;
  const Fp32Layout L = fp32_layout(g, P, true);
  if (ws == nullptr || ws_bytes < L.total) {
    set_error("nb_mlp_backward: workspace of %zu bytes given, %zu needed", ws_bytes, L.total);
    return NB_ERR_WORKSPACE;
  }
  float* EX = wsf(ws, L.off_ex);
  float* ED = wsf(ws, L.off_ed);
  float* H = wsf(ws, L.off_h);
  float* FEAT = wsf(ws, L.off_feat);
  float* DIRH = wsf(ws, L.off_dirh);
  float* G[2] = {wsf(ws, L.off_g0), wsf(ws, L.off_g1)};
  float* GFEAT = wsf(ws, L.off_gfeat);
  float* GDIR = wsf(ws, L.off_gdir);
  float* PART = wsf(ws, L.off_part);
  auto hbuf = [&](int i) { return H + (size_t)i * L.P * g.width; };
  const int D = g.depth;
  const float* HL = hbuf(D - 1);
  const float* maskL = layer_has_relu(g, D - 1) ? HL : nullptr;
  int rc;
#define NB_TRY(x) do { if ((rc = (x)) != NB_OK) return rc; } while (0)
  int cur = 0;  // G[cur] holds dZ of the trunk layer being processed
  if (g.use_viewdirs) {
    const Lin& la = g.lins[g.alpha];
    const Lin& lf = g.lins[g.feat];
    const Lin& ld = g.lins[g.dir];
    const Lin& lr = g.lins[g.rgb];
    // rgb head
    NB_TRY(wgrad(d_raw, 4, 3, DIRH, g.dir_width, lr.in, grad + lr.w_off, lr.in, P, PART, s));
    NB_TRY(launch_colsum(d_raw, 4, P, 3, grad + lr.b_off, PART, s));
    NB_TRY(dgrad(d_raw, 4, 3, W + lr.w_off, lr.in, lr.in, GDIR, g.dir_width, P, DIRH, g.dir_width, 0, s));
    // dir layer
    NB_TRY(wgrad(GDIR, g.dir_width, ld.out, FEAT, g.feat_width, g.feat_width, grad + ld.w_off, ld.in, P, PART, s));
    if (g.Cd > 0)
      NB_TRY(wgrad(GDIR, g.dir_width, ld.out, ED, L.ldd, g.Cd, grad + ld.w_off + g.feat_width, ld.in, P, PART, s));
    NB_TRY(launch_colsum(GDIR, g.dir_width, P, ld.out, grad + ld.b_off, PART, s));
    NB_TRY(dgrad(GDIR, g.dir_width, ld.out, W + ld.w_off, ld.in, g.feat_width, GFEAT, g.feat_width, P,
                 g.feat_act ? FEAT : nullptr, g.feat_width, 0, s));
    // feat / bottleneck
    NB_TRY(wgrad(GFEAT, g.feat_width, lf.out, HL, g.width, lf.in, grad + lf.w_off, lf.in, P, PART, s));
    NB_TRY(launch_colsum(GFEAT, g.feat_width, P, lf.out, grad + lf.b_off, PART, s));
    NB_TRY(dgrad(GFEAT, g.feat_width, lf.out, W + lf.w_off, lf.in, lf.in, G[cur], g.width, P, nullptr, 0, 0, s));
    // alpha head (sums into the same dH, then the ReLU mask of the last trunk layer)
    NB_TRY(wgrad(d_raw + 3, 4, 1, HL, g.width, la.in, grad + la.w_off, la.in, P, PART, s));
    NB_TRY(launch_colsum(d_raw + 3, 4, P, 1, grad + la.b_off, PART, s));
    NB_TRY(dgrad(d_raw + 3, 4, 1, W + la.w_off, la.in, la.in, G[cur], g.width, P, maskL, g.width, 1, s));
  } else {
    const Lin& lo = g.lins[g.out4];
    NB_TRY(wgrad(d_raw, 4, 4, HL, g.width, lo.in, grad + lo.w_off, lo.in, P, PART, s));
    NB_TRY(launch_colsum(d_raw, 4, P, 4, grad + lo.b_off, PART, s));
    NB_TRY(dgrad(d_raw, 4, 4, W + lo.w_off, lo.in, lo.in, G[cur], g.width, P, maskL, g.width, 0, s));
  }
  for (int i = D - 1; i >= 0; --i) {
    const Lin& l = g.lins[g.trunk[i]];
    const float* dZ = G[cur];
    NB_TRY(launch_colsum(dZ, g.width, P, l.out, grad + l.b_off, PART, s));
    if (i == 0) {
      NB_TRY(wgrad(dZ, g.width, l.out, EX, L.ldx, g.Cx, grad + l.w_off, l.in, P, PART, s));
      break;
    }
    const bool cat = (i - 1 == g.skip);
    const int off = cat ? g.Cx : 0;
    if (cat) NB_TRY(wgrad(dZ, g.width, l.out, EX, L.ldx, g.Cx, grad + l.w_off, l.in, P, PART, s));
    NB_TRY(wgrad(dZ, g.width, l.out, hbuf(i - 1), g.width, g.width, grad + l.w_off + off, l.in, P, PART, s));
    const float* mk = layer_has_relu(g, i - 1) ? hbuf(i - 1) : nullptr;
    NB_TRY(dgrad(dZ, g.width, l.out, W + l.w_off + off, l.in, g.width, G[cur ^ 1], g.width, P, mk, g.width, 0, s));
    cur ^= 1;
  }
#undef NB_TRY
  return NB_OK;
}

}  // namespace nb

using namespace nb;

extern "C" {

size_t nb_mlp_workspace_bytes(const nb_model_t* m, int precision, int64_t n_points, int save_for_backward) {
  if (m == nullptr || n_points < 0) return 0;
  if (precision == NB_PREC_BF16) return tc_workspace_bytes(m, n_points, save_for_backward);
  return fp32_layout(m->g, n_points, save_for_backward != 0).total;
}

int nb_mlp_forward(const nb_model_t* m, int precision, const float* rays, int ray_stride, const float* z_vals,
                   int64_t n_rays, int n_samples, float* raw, void* workspace, size_t workspace_bytes,
                   int save_for_backward, nb_stream_t stream) {
  NB_REQUIRE(m != nullptr && m->params != nullptr, "nb_mlp_forward: model has no parameters (nb_model_set_params)");
  NB_REQUIRE(n_rays >= 0 && n_samples > 0, "nb_mlp_forward: bad sizes");
  NB_REQUIRE(ray_stride == 8 || ray_stride == 11, "nb_mlp_forward: ray_stride must be 8 or 11");
  NB_REQUIRE(!m->g.use_viewdirs || ray_stride == 11, "nb_mlp_forward: the model uses view directions but rays are [n, 8]");
  if (n_rays == 0) return NB_OK;
  NB_REQUIRE(rays && z_vals && raw, "nb_mlp_forward: null pointer");
  const int64_t P = n_rays * n_samples;
  if (precision == NB_PREC_BF16)
    return tc_forward(m, rays, ray_stride, z_vals, nullptr, nullptr, 1, n_rays, n_samples, P, raw, workspace,
                      workspace_bytes, save_for_backward, as_stream(stream));
  NB_REQUIRE(precision == NB_PREC_FP32, "nb_mlp_forward: unknown precision %d", precision);
  EncodeSrc src{rays, ray_stride, z_vals, n_samples, nullptr, nullptr, 1};
  return fp32_forward(m, src, P, raw, workspace, workspace_bytes, save_for_backward, as_stream(stream));
}

int nb_mlp_forward_points(const nb_model_t* m, int precision, const float* pts, const float* viewdirs, int64_t n_points,
                          int points_per_dir, float* raw, void* workspace, size_t workspace_bytes, nb_stream_t stream) {
  NB_REQUIRE(m != nullptr && m->params != nullptr, "nb_mlp_forward_points: model has no parameters");
  NB_REQUIRE(n_points >= 0, "nb_mlp_forward_points: bad sizes");
  if (n_points == 0) return NB_OK;
  NB_REQUIRE(pts && raw, "nb_mlp_forward_points: null pointer");
  NB_REQUIRE(!m->g.use_viewdirs || (viewdirs != nullptr && points_per_dir > 0),
             "nb_mlp_forward_points: the model uses view directions but none were given");
  if (precision == NB_PREC_BF16)
    return tc_forward(m, nullptr, 0, nullptr, pts, viewdirs, points_per_dir, 0, 0, n_points, raw, workspace,
                      workspace_bytes, 0, as_stream(stream));
  NB_REQUIRE(precision == NB_PREC_FP32, "nb_mlp_forward_points: unknown precision %d", precision);
  EncodeSrc src{nullptr, 0, nullptr, 1, pts, viewdirs, points_per_dir > 0 ? points_per_dir : 1};
  return fp32_forward(m, src, n_points, raw, workspace, workspace_bytes, 0, as_stream(stream));
}

int nb_mlp_backward(const nb_model_t* m, int precision, int64_t n_points, float* d_raw, void* workspace,
                    size_t workspace_bytes, float* flat_grad, nb_stream_t stream) {
  NB_REQUIRE(m != nullptr && m->params != nullptr, "nb_mlp_backward: model has no parameters");
  NB_REQUIRE(n_points >= 0, "nb_mlp_backward: bad sizes");
  if (n_points == 0) return NB_OK;
  NB_REQUIRE(d_raw && flat_grad, "nb_mlp_backward: null pointer");
  if (precision == NB_PREC_BF16)
    return tc_backward(m, n_points, d_raw, workspace, workspace_bytes, flat_grad, as_stream(stream));
  NB_REQUIRE(precision == NB_PREC_FP32, "nb_mlp_backward: unknown precision %d", precision);
  return fp32_backward(m, n_points, d_raw, workspace, workspace_bytes, flat_grad, as_stream(stream));
}

}  // extern "C"
