// Fused bf16 tensor-core MLP (NB_PREC_BF16) for sm_100a: positional encoding + every layer of the
// network in ONE persistent kernel; activations never leave the SM between layers.
//
// What it computes: run_network + the network forward (nerf/run_nerf.py:30-45 with the Embedder of
// nerf/run_nerf_helpers.py:22-56 and the graph of :80-118; nerf-pytorch/nerf/train_utils.py:8-25 with
// models.py:233-256) for a tile of 128 points at a time.
//
// How (Blackwell execution model):
//   * one CTA per SM, persistent over 128-point tiles; 6 warps with fixed roles:
//       warp 0    weight producer  -- streams pre-swizzled bf16 weight chunks (N x 64 K-elements, the exact
//                                     UMMA SWIZZLE_128B K-major image) from L2 into a shared-memory ring with
//                                     bulk async copies (cp.async.bulk -> UBLKCP) completing on mbarriers;
//       warp 1    MMA issuer       -- one thread issues tcgen05.mma (kind::f16, bf16 x bf16 -> fp32, M=128,
//                                     N<=256, K=16) with A = activations in shared memory, B = the ring,
//                                     D = fp32 accumulators in TMEM (two 256-column buffers, alternating by
//                                     layer); tcgen05.commit releases ring slots / publishes accumulators;
//       warps 2-5 epilogue         -- thread == row: encode the tile's points (sin/cos by double-angle
//                                     recurrence) straight into the swizzled A-operand layout, then per layer
//                                     tcgen05.ld the accumulator 64 columns at a time, add bias, ReLU, pack to
//                                     bf16 and store it as the next layer's A operand.  Each finished
//                                     64-column block is published on its own mbarrier, so the next layer's
//                                     MMAs start on block 0 while blocks 1..3 are still being converted.
//   * the skip concat ([pts, h], K = 64 + 256) and the view-direction concat ([feat, dirs]) are extra
//     K-blocks fed from the resident encoding tiles, not copies; alpha (N=1) rides along the
//     bottleneck / fc_feat GEMM as extra output columns; rgb is an N=16 MMA.
#include <algorithm>
#include <vector>

#include "tc_common.cuh"

namespace nb {

// ------------------------------------------------------------------------------------ weight packing
__global__ void pack_weights_kernel(const float* __restrict__ params, const PackSeg* __restrict__ segs,
                                    unsigned char* __restrict__ out) {
  const PackSeg sg = segs[blockIdx.y];
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= sg.n_rows * 8) return;
  const int r = idx >> 3, c = idx & 7;
  const int n = sg.row0 + r;
  const float* src = params + sg.src_off + (long long)r * sg.ld + sg.col0;
  float v[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const int k = c * 8 + e;
    v[e] = k < sg.ncols ? src[k] : 0.0f;
  }
  uint4 pk;
  pk.x = pack_h2(v[0], v[1], false);
  pk.y = pack_h2(v[2], v[3], false);
  pk.z = pack_h2(v[4], v[5], false);
  pk.w = pack_h2(v[6], v[7], false);
  *reinterpret_cast<uint4*>(out + sg.dst_off + (long long)n * 128 + ((c ^ (n & 7)) << 4)) = pk;
}

__global__ void pack_bias_kernel(const float* __restrict__ params, const BiasSeg* __restrict__ segs,
                                 float* __restrict__ out) {
  const BiasSeg sg = segs[blockIdx.x];
  for (int i = threadIdx.x; i < sg.count; i += blockDim.x) out[sg.dst + i] = params[sg.src_off + i];
}



// ------------------------------------------------------------------------------------ point source
struct TcSrc {
  const float* rays;  // [n_rays, stride] or null
  int stride;
  const float* z;  // [n_rays, S]
  int S;
  const float* pts;       // [n_points, 3] when rays == null
  const float* viewdirs;  // [n_points / per_dir, 3]
  int per_dir;
};

// Encode one 3-vector into 64 fp16 channels [v, sin(2^0 v), cos(2^0 v), ...] and store them as row `row` of a
// SWIZZLE_128B K-block.  Higher octaves by the double-angle recurrence from one accurate sincosf per component:
// the absolute error doubles per octave (<= 2^9 * 6e-8), far below the fp16 quantum (4.9e-4) of the stored value.
__device__ __forceinline__ void encode_row(unsigned char* block, int row, float x0, float x1, float x2, int L) {
  float s[3], c[3];
  sincosf(x0, &s[0], &c[0]);
  sincosf(x1, &s[1], &c[1]);
  sincosf(x2, &s[2], &c[2]);
  float ch[64];
  ch[0] = x0; ch[1] = x1; ch[2] = x2;
#pragma unroll
  for (int f = 0; f < 10; ++f) {
    const bool on = f < L;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      ch[3 + 6 * f + k] = on ? s[k] : 0.0f;
      ch[6 + 6 * f + k] = on ? c[k] : 0.0f;
      const float s2 = 2.0f * s[k] * c[k];
      const float c2 = 1.0f - 2.0f * s[k] * s[k];
      s[k] = s2; c[k] = c2;
    }
  }
  ch[63] = 0.0f;
  unsigned char* rp = block + row * 128;
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    uint4 pk;
    pk.x = pack_h2(ch[8 * q + 0], ch[8 * q + 1], false);
    pk.y = pack_h2(ch[8 * q + 2], ch[8 * q + 3], false);
    pk.z = pack_h2(ch[8 * q + 4], ch[8 * q + 5], false);
    pk.w = pack_h2(ch[8 * q + 6], ch[8 * q + 7], false);
    *reinterpret_cast<uint4*>(rp + ((q ^ (row & 7)) << 4)) = pk;
  }
}

// ------------------------------------------------------------------------------------ the kernel
__global__ void __launch_bounds__(kThreadsTc, 1)
mlp_tc_kernel(const __grid_constant__ TcPlan plan, const unsigned char* __restrict__ wpk, const float* __restrict__ bias_g,
              TcSrc src, long long n_points, float* __restrict__ raw, unsigned char* __restrict__ rec) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  // carve (all tile regions 1024-byte aligned for SWIZZLE_128B)
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  unsigned char* act = smem;
  unsigned char* encx = act + plan.n_act_blocks * kBlockBytes;
  unsigned char* encd = encx + kBlockBytes;
  unsigned char* ring = encd + kBlockBytes;
  float* bias_s = reinterpret_cast<float*>(ring + kRing * plan.ring_stage_bytes);
  uint64_t* bars = reinterpret_cast<uint64_t*>(bias_s + ((plan.n_bias + 3) & ~3));
  // barrier indices
  //   [0,kRing) full, [kRing,2kRing) empty, then enc_ready, act_ready[4], acc_full[2]
  uint64_t* bar_full = bars;
  uint64_t* bar_empty = bars + kRing;
  uint64_t* bar_enc = bars + 2 * kRing;
  uint64_t* bar_act = bar_enc + 1;
  uint64_t* bar_acc = bar_act + 4;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_acc + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long n_tiles = (n_points + kTileM - 1) / kTileM;

  if (threadIdx.x == 0) {
    for (int i = 0; i < kRing; ++i) {
      mbar_init(smem_u32(bar_full + i), 1);
      mbar_init(smem_u32(bar_empty + i), 1);
    }
    mbar_init(smem_u32(bar_enc), 128);
    for (int i = 0; i < 4; ++i) mbar_init(smem_u32(bar_act + i), 128);
    mbar_init(smem_u32(bar_acc), 1);
    mbar_init(smem_u32(bar_acc + 1), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {  // TMEM: all 512 columns (one CTA per SM)
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  for (int i = threadIdx.x; i < plan.n_bias; i += blockDim.x) bias_s[i] = bias_g[i];
  // encoding tiles: padding columns are rewritten every tile by encode_row, nothing to clear
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ============================ weight producer ============================
    if (lane == 0) {
      uint32_t slot = 0, phase = 0;
      for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        for (int s = 0; s < plan.n_stages; ++s) {
          const TcStage& st = plan.st[s];
          for (int kb = 0; kb < st.n_kb; ++kb) {
            mbar_wait(smem_u32(bar_empty + slot), phase ^ 1);
            const uint32_t fb = smem_u32(bar_full + slot);
            mbar_expect_tx(fb, (uint32_t)st.chunk_bytes);
            bulk_g2s(smem_u32(ring + slot * plan.ring_stage_bytes), wpk + st.w_off + (size_t)kb * st.chunk_bytes,
                     (uint32_t)st.chunk_bytes, fb);
            if (++slot == kRing) { slot = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ============================ MMA issuer ============================
    if (lane == 0) {
      uint32_t slot = 0, phase = 0, enc_phase = 0;
      uint32_t act_phase[4] = {0, 0, 0, 0};
      for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        mbar_wait(smem_u32(bar_enc), enc_phase);
        enc_phase ^= 1;
        tc_fence_after();
        // train: every operand tile is also written to the per-tile X record, in the exact swizzled image the
        // backward kernels load (K-major for recompute-free dgrad masks, MN-major operand of the wgrad MMAs)
        unsigned char* rec_tile = rec ? rec + (size_t)tile * plan.rec_blocks * kBlockBytes : nullptr;
        if (rec_tile) {
          bulk_s2g(rec_tile, smem_u32(encx), kBlockBytes);
          if (plan.Cd > 0) bulk_s2g(rec_tile + kBlockBytes, smem_u32(encd), kBlockBytes);
          bulk_commit();
        }
        for (int s = 0; s < plan.n_stages; ++s) {
          const TcStage& st = plan.st[s];
          const int buf = s & 1;
          const uint32_t d_main = tmem_base + buf * 256;
          const uint32_t d_extra = tmem_base + (buf ^ 1) * 256;
          const uint32_t idesc_main = umma_idesc(st.n_main);
          const uint32_t idesc_extra = umma_idesc(16);
          for (int kb = 0; kb < st.n_kb; ++kb) {
            const int sidx = st.a_src[kb];
            uint32_t a_base;
            if (sidx < 4) {
              mbar_wait(smem_u32(bar_act + sidx), act_phase[sidx]);
              act_phase[sidx] ^= 1;
              tc_fence_after();
              a_base = smem_u32(act + sidx * kBlockBytes);
              if (rec_tile) {
                bulk_s2g(rec_tile + (size_t)(plan.st[s - 1].rec_blk + sidx) * kBlockBytes, a_base, kBlockBytes);
                bulk_commit();
              }
            } else {
              a_base = smem_u32(sidx == SRC_ENCX ? encx : encd);
            }
            mbar_wait(smem_u32(bar_full + slot), phase);
            tc_fence_after();
            const uint32_t b_base = smem_u32(ring + slot * plan.ring_stage_bytes);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const uint32_t accum = (kb | k) != 0 ? 1u : 0u;
              tc_mma(d_main, umma_desc(a_base + k * 32), umma_desc(b_base + k * 32), idesc_main, accum);
              if (st.n_extra)
                tc_mma(d_extra, umma_desc(a_base + k * 32), umma_desc(b_base + st.n_main * 128 + k * 32), idesc_extra,
                       accum);
            }
            tc_commit(smem_u32(bar_empty + slot));  // frees the ring slot when these MMAs have read it
            if (++slot == kRing) { slot = 0; phase ^= 1; }
          }
          if (rec_tile) bulk_wait_read0();     // record stores have read the tiles this stage's epilogue overwrites
          tc_commit(smem_u32(bar_acc + buf));  // accumulator of this stage complete
        }
      }
      if (rec) bulk_wait0();
    }
  } else {
    // ============================ epilogue warps (thread == row) ============================
    const int q = warp & 3;             // TMEM lane quarter this warp may access
    const int row = q * 32 + lane;      // row of the tile
    const uint32_t lane_addr = ((uint32_t)(q * 32)) << 16;
    uint32_t acc_phase[2] = {0, 0};
    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
      long long p = tile * kTileM + row;
      const bool in_range = p < n_points;
      if (!in_range) p = n_points - 1;
      // ---- positional encoding of this row's point and view direction
      {
        float x0, x1, x2, v0 = 0.f, v1 = 0.f, v2 = 0.f;
        if (src.rays != nullptr) {
          const long long r = p / src.S;
          const float* ry = src.rays + r * src.stride;
          const float zz = src.z[p];
          x0 = __fadd_rn(ry[0], __fmul_rn(ry[3], zz));
          x1 = __fadd_rn(ry[1], __fmul_rn(ry[4], zz));
          x2 = __fadd_rn(ry[2], __fmul_rn(ry[5], zz));
          if (plan.Cd > 0) { v0 = ry[8]; v1 = ry[9]; v2 = ry[10]; }
        } else {
          x0 = src.pts[3 * p]; x1 = src.pts[3 * p + 1]; x2 = src.pts[3 * p + 2];
          if (plan.Cd > 0) {
            const float* vv = src.viewdirs + (p / src.per_dir) * 3;
            v0 = vv[0]; v1 = vv[1]; v2 = vv[2];
          }
        }
        encode_row(encx, row, x0, x1, x2, plan.Lx);
        if (plan.Cd > 0) encode_row(encd, row, v0, v1, v2, plan.Ld);
      }
      tc_fence_before();  // our TMEM loads of the previous tile are ordered before the MMAs this arrive releases
      fence_async_smem();
      mbar_arrive(smem_u32(bar_enc));

      float alpha_raw = 0.0f;
      for (int s = 0; s < plan.n_stages; ++s) {
        const TcStage& st = plan.st[s];
        const int buf = s & 1;
        mbar_wait(smem_u32(bar_acc + buf), acc_phase[buf]);
        acc_phase[buf] ^= 1;
        tc_fence_after();
        const uint32_t t_main = tmem_base + lane_addr + buf * 256;
        const float* bs = bias_s + st.b_off;
        if (st.epi == EPI_HEAD) {
          const uint32_t ta = st.alpha_col >= 0 ? t_main + st.alpha_col : tmem_base + lane_addr + (buf ^ 1) * 256;
          const uint32_t a = tmem_ld1(ta);
          tc_wait_ld();
          alpha_raw = __uint_as_float(a) + bs[st.n_act];
        }
        if (st.epi == EPI_HIDDEN || st.epi == EPI_HEAD) {
          const bool relu = st.relu != 0;
          const int nblk = st.n_act >> 6;
          for (int j = 0; j < nblk; ++j) {
            uint32_t v[64];
            NB_TMEM_LD32(t_main + j * 64, v);
            {
              uint32_t* v2 = v + 32;
              NB_TMEM_LD32(t_main + j * 64 + 32, v2);
            }
            tc_wait_ld();
            unsigned char* rp = act + j * kBlockBytes + row * 128;
            const float* bj = bs + j * 64;
#pragma unroll
            for (int c = 0; c < 8; ++c) {
              const float4 b0 = *reinterpret_cast<const float4*>(bj + 8 * c);
              const float4 b1 = *reinterpret_cast<const float4*>(bj + 8 * c + 4);
              uint4 pk;
              pk.x = pack_h2(__uint_as_float(v[8 * c + 0]) + b0.x, __uint_as_float(v[8 * c + 1]) + b0.y, relu);
              pk.y = pack_h2(__uint_as_float(v[8 * c + 2]) + b0.z, __uint_as_float(v[8 * c + 3]) + b0.w, relu);
              pk.z = pack_h2(__uint_as_float(v[8 * c + 4]) + b1.x, __uint_as_float(v[8 * c + 5]) + b1.y, relu);
              pk.w = pack_h2(__uint_as_float(v[8 * c + 6]) + b1.z, __uint_as_float(v[8 * c + 7]) + b1.w, relu);
              *reinterpret_cast<uint4*>(rp + ((c ^ (row & 7)) << 4)) = pk;
            }
            tc_fence_before();
            fence_async_smem();
            mbar_arrive(smem_u32(bar_act + j));
          }
        } else {
          uint32_t v[4];
          tmem_ld4(t_main, v);
          tc_wait_ld();
          float4 o;
          o.x = __uint_as_float(v[0]) + bs[0];
          o.y = __uint_as_float(v[1]) + bs[1];
          o.z = __uint_as_float(v[2]) + bs[2];
          o.w = st.epi == EPI_RGB ? alpha_raw : __uint_as_float(v[3]) + bs[3];
          if (in_range) reinterpret_cast<float4*>(raw)[tile * kTileM + row] = o;
        }
      }
    }
  }
  // ---- teardown
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
  }
}

// ------------------------------------------------------------------------------------ host: plan + packing
void packed_free(PackedBf16* p) {
  if (p == nullptr) return;
  bw_free(p);
  cudaFree(p->d_w);
  cudaFree(p->d_bias);
  cudaFree(p->d_segs);
  cudaFree(p->d_bsegs);
  delete p;
}

static bool build_plan(const Graph& g, TcPlan& plan, std::vector<PackSeg>& segs, std::vector<BiasSeg>& bsegs,
                       size_t& w_bytes, const char** why) {
  memset(&plan, 0, sizeof(plan));
  *why = nullptr;
  if (g.width % 128 != 0 || g.width > 256) { *why = "width must be 128 or 256"; return false; }
  if (g.use_viewdirs && (g.feat_width % 64 != 0 || g.feat_width > 256)) { *why = "feature width must be <= 256"; return false; }
  if (g.Cx > 64 || g.Cd > 64) { *why = "more than 10 encoding frequencies"; return false; }
  if (g.skip >= g.depth - 1) { *why = "skip connection after the last trunk layer"; return false; }
  const int wb = g.width / 64;
  plan.n_act_blocks = std::max(wb, g.use_viewdirs ? g.feat_width / 64 : 0);
  plan.Cx = g.Cx; plan.Cd = g.Cd; plan.Lx = g.n_freq_xyz; plan.Ld = g.n_freq_dir;
  plan.rec_blocks = 2;  // [encx, encd]
  size_t woff = 0;
  int boff = 0, ring_max = 0;
  auto begin_stage = [&](int n_main, int n_extra, int epi, int relu, int n_act, int alpha_col) -> TcStage& {
    TcStage& st = plan.st[plan.n_stages++];
    st.n_kb = 0;
    st.n_main = n_main; st.n_extra = n_extra; st.chunk_bytes = (n_main + n_extra) * 128;
    st.w_off = (int)woff; st.b_off = boff; st.epi = epi; st.relu = relu; st.n_act = n_act; st.alpha_col = alpha_col;
    st.rec_blk = plan.rec_blocks;
    plan.rec_blocks += n_act / 64;
    boff += (n_main + n_extra + 3) & ~3;
    ring_max = std::max(ring_max, st.chunk_bytes);
    return st;
  };
  // adds one K-block to the stage: rows from up to two source matrices
  auto add_kb = [&](TcStage& st, int a_src, std::initializer_list<PackSeg> rows) {
    for (PackSeg sg : rows) {
      sg.dst_off = (long long)woff;
      segs.push_back(sg);
    }
    st.a_src[st.n_kb++] = a_src;
    woff += st.chunk_bytes;
  };
  auto seg = [](const Lin& l, int row0, int col0, int ncols) {
    PackSeg s{};
    s.src_off = l.w_off; s.row0 = row0; s.n_rows = l.out; s.ld = l.in; s.col0 = col0; s.ncols = ncols;
    return s;
  };
  if (g.depth + 3 > kMaxStages) { *why = "too many layers"; return false; }
  for (int i = 0; i < g.depth; ++i) {
    const Lin& l = g.lins[g.trunk[i]];
    TcStage& st = begin_stage(g.width, 0, EPI_HIDDEN, !(i == 0 && !g.act0), g.width, 0);
    bsegs.push_back(BiasSeg{st.b_off, l.out, l.b_off});
    if (i == 0) {
      add_kb(st, SRC_ENCX, {seg(l, 0, 0, g.Cx)});
    } else {
      int off = 0;
      if (i - 1 == g.skip) {
        add_kb(st, SRC_ENCX, {seg(l, 0, 0, g.Cx)});
        off = g.Cx;
      }
      for (int j = 0; j < wb; ++j) add_kb(st, j, {seg(l, 0, off + 64 * j, 64)});
    }
  }
  if (g.use_viewdirs) {
    const Lin& la = g.lins[g.alpha];
    const Lin& lf = g.lins[g.feat];
    const Lin& ld = g.lins[g.dir];
    const Lin& lr = g.lins[g.rgb];
    if (g.dir_width % 64 != 0) { *why = "direction layer width must be a multiple of 64"; return false; }
    const bool split = g.feat_width + 16 > 256;
    TcStage& sh = begin_stage(split ? g.feat_width : g.feat_width + 16, split ? 16 : 0, EPI_HEAD, g.feat_act ? 1 : 0,
                              g.feat_width, split ? -1 : g.feat_width);
    bsegs.push_back(BiasSeg{sh.b_off, lf.out, lf.b_off});
    bsegs.push_back(BiasSeg{sh.b_off + g.feat_width, 1, la.b_off});
    for (int j = 0; j < wb; ++j) add_kb(sh, j, {seg(lf, 0, 64 * j, 64), seg(la, g.feat_width, 64 * j, 64)});
    // direction layer: activation blocks first (its first wait also orders it after the alpha read), encoding last
    TcStage& sd = begin_stage(g.dir_width, 0, EPI_HIDDEN, 1, g.dir_width, 0);
    bsegs.push_back(BiasSeg{sd.b_off, ld.out, ld.b_off});
    for (int j = 0; j < g.feat_width / 64; ++j) add_kb(sd, j, {seg(ld, 0, 64 * j, 64)});
    add_kb(sd, SRC_ENCD, {seg(ld, 0, g.feat_width, g.Cd)});
    TcStage& sr = begin_stage(16, 0, EPI_RGB, 0, 0, 0);
    bsegs.push_back(BiasSeg{sr.b_off, 3, lr.b_off});
    for (int j = 0; j < g.dir_width / 64; ++j) add_kb(sr, j, {seg(lr, 0, 64 * j, 64)});
  } else {
    const Lin& lo = g.lins[g.out4];
    TcStage& so = begin_stage(16, 0, EPI_OUT4, 0, 0, 0);
    bsegs.push_back(BiasSeg{so.b_off, 4, lo.b_off});
    for (int j = 0; j < wb; ++j) add_kb(so, j, {seg(lo, 0, 64 * j, 64)});
  }
  for (int s = 0; s < plan.n_stages; ++s)
    if (plan.st[s].n_kb > kMaxKB) { *why = "too many K blocks"; return false; }
  plan.n_bias = boff;
  plan.ring_stage_bytes = (int)round_up(ring_max, 1024);
  w_bytes = woff;
  return true;
}

static size_t tc_smem_bytes(const TcPlan& plan) {
  size_t b = 1024;  // alignment slack
  b += (size_t)(plan.n_act_blocks + 2) * kBlockBytes;
  b += (size_t)kRing * plan.ring_stage_bytes;
  b += (size_t)((plan.n_bias + 3) & ~3) * sizeof(float);
  b += (2 * kRing + 1 + 4 + 2) * sizeof(uint64_t) + 16;
  return b;
}

int packed_build(nb_model* m, cudaStream_t stream) {
  if (m->packed == nullptr) {
    PackedBf16* p = new PackedBf16();
    std::vector<PackSeg> segs;
    std::vector<BiasSeg> bsegs;
    const char* why = nullptr;
    p->supported = build_plan(m->g, p->plan, segs, bsegs, p->w_bytes, &why);
    if (!p->supported) {
      delete p;
      set_error("bf16 tensor-core path does not support this graph: %s (use NB_PREC_FP32)", why ? why : "?");
      return NB_ERR_UNSUPPORTED;
    }
    p->smem_bytes = tc_smem_bytes(p->plan);
    if (p->smem_bytes > 227 * 1024) {
      delete p;
      set_error("bf16 tensor-core path: %zu bytes of shared memory needed", p->smem_bytes);
      return NB_ERR_UNSUPPORTED;
    }
    p->n_segs = (int)segs.size();
    p->n_bsegs = (int)bsegs.size();
    cudaError_t e = cudaMalloc(&p->d_w, p->w_bytes);
    if (e == cudaSuccess) e = cudaMalloc(&p->d_bias, sizeof(float) * std::max(p->plan.n_bias, 4));
    if (e == cudaSuccess) e = cudaMalloc(&p->d_segs, sizeof(PackSeg) * segs.size());
    if (e == cudaSuccess) e = cudaMalloc(&p->d_bsegs, sizeof(BiasSeg) * bsegs.size());
    // padding rows / bias slots never change: clear them once
    if (e == cudaSuccess) e = cudaMemsetAsync(p->d_w, 0, p->w_bytes, stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(p->d_bias, 0, sizeof(float) * std::max(p->plan.n_bias, 4), stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(p->d_segs, segs.data(), sizeof(PackSeg) * segs.size(), cudaMemcpyHostToDevice, stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(p->d_bsegs, bsegs.data(), sizeof(BiasSeg) * bsegs.size(), cudaMemcpyHostToDevice, stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(stream);  // the host vectors die with this scope
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(mlp_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) {
      packed_free(p);
      set_error("bf16 pack setup failed: %s", cudaGetErrorString(e));
      return NB_ERR_CUDA;
    }
    m->packed = p;
    int rc = bw_setup(m, stream);
    if (rc != NB_OK) return rc;
  }
  PackedBf16* p = m->packed;
  p->wt_fresh = false;  // the transposed images are rebuilt by the next backward call
  int max_rows = 0;
  // one launch over all segments; blockIdx.x covers the largest segment
  max_rows = 272;
  dim3 grid((unsigned)ceil_div(max_rows * 8, 256), (unsigned)p->n_segs);
  pack_weights_kernel<<<grid, 256, 0, stream>>>(m->params, p->d_segs, p->d_w);
  NB_LAUNCHED();
  pack_bias_kernel<<<p->n_bsegs, 128, 0, stream>>>(m->params, p->d_bsegs, p->d_bias);
  NB_LAUNCHED();
  return NB_OK;
}

size_t tc_workspace_bytes(const nb_model* m, int64_t n_points, int save_for_backward) {
  if (!save_for_backward || m->packed == nullptr || !m->packed->bw_supported) return 0;
  return tc_train_workspace_bytes(m->packed, n_points);
}

int tc_forward(const nb_model* m, const float* rays, int ray_stride, const float* z_vals, const float* pts,
               const float* viewdirs, int points_per_dir, int64_t n_rays, int n_samples, int64_t n_points, float* raw,
               void* ws, size_t ws_bytes, int save_for_backward, cudaStream_t stream) {
  (void)n_rays;
  if (m->packed == nullptr) {
    set_error("nb_mlp_forward: nb_model_pack(NB_PREC_BF16) has not been called");
    return NB_ERR_INVALID;
  }
  const PackedBf16* p = m->packed;
  unsigned char* rec = nullptr;
  if (save_for_backward) {
    if (!p->bw_supported) {
      set_error("tensor-core backward does not support this graph (train it with NB_PREC_FP32)");
      return NB_ERR_UNSUPPORTED;
    }
    const size_t need = tc_train_workspace_bytes(p, n_points);
    if (ws == nullptr || ws_bytes < need) {
      set_error("nb_mlp_forward: workspace of %zu bytes given, %zu needed", ws_bytes, need);
      return NB_ERR_WORKSPACE;
    }
    rec = tc_train_xrec(p, ws, n_points);
  }
  TcSrc src{rays, ray_stride, z_vals, n_samples > 0 ? n_samples : 1, pts, viewdirs, points_per_dir > 0 ? points_per_dir : 1};
  const int64_t n_tiles = ceil_div(n_points, kTileM);
  const int grid = (int)std::min<int64_t>(n_tiles, sm_count());
  mlp_tc_kernel<<<grid, kThreadsTc, p->smem_bytes, stream>>>(p->plan, p->d_w, p->d_bias, src, (long long)n_points, raw, rec);
  NB_LAUNCHED();
  return NB_OK;
}

}  // namespace nb
