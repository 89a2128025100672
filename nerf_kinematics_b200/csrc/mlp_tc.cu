// Fused 16-bit tensor-core MLP (NB_PREC_BF16) for sm_100a: positional encoding + every layer of the
// network in ONE persistent kernel; activations never leave the SM between layers.
//
// What it computes: run_network + the network forward (nerf/run_nerf.py:30-45 with the Embedder of
// nerf/run_nerf_helpers.py:22-56 and the graph of :80-118; nerf-pytorch/nerf/train_utils.py:8-25 with
// models.py:233-256) for PAIRS of 128-point tiles.
//
// How (Blackwell execution model):
//   * one CTA per SM, persistent over a contiguous range of tiles taken two at a time; 20 warps with fixed roles:
//       warp 0     weight producer -- streams pre-swizzled 16-bit weight chunks (<= 256 N-rows x 32 K-elements, the
//                                     exact UMMA SWIZZLE_64B K-major image) from L2 into a 3-slot shared-memory ring
//                                     with bulk async copies (cp.async.bulk -> UBLKCP) completing on mbarriers;
//       warps 1-2  MMA issuers     -- one warp per tile walks a step program in warp-uniform control flow and issues
//                                     tcgen05.mma (kind::f16, M=128, N<=256, K=16) from an elected lane: every ring
//                                     slot feeds BOTH tiles of the pair (A = that tile's activations in shared memory,
//                                     D = that tile's 256 fp32 TMEM columns), halving the weight traffic per point;
//                                     tcgen05.commit releases ring slots / publishes the accumulators; in training
//                                     the same warps bulk-store every operand block to the per-tile X record;
//       warp 3     builds the step program, then idles;
//       warps 4-19 epilogue        -- 8 warps per tile (TMEM lane quarter x column-block parity), thread == row:
//                                     encode the tile's points (sin/cos by double-angle recurrence) straight into the
//                                     swizzled A-operand layout, then per layer tcgen05.ld 64 accumulator columns,
//                                     ReLU, pack to 16 bit and store them as the next layer's A operand (the bias is
//                                     already in the accumulator: it rides in the MMA against the encoding block's
//                                     constant-one channel).
//                                     Each finished 64-column block is published on its own mbarrier and the
//                                     accumulator is handed back as soon as the last tcgen05.ld has landed, so the
//                                     next layer's MMAs start while the remaining blocks are still being converted.
//   * the skip concat ([pts, h], K = 64 + 256) and the view-direction concat ([feat, dirs]) are extra
//     K-blocks fed from the resident encoding block (position encoding first; re-filled with the direction
//     encoding once the skip layer has consumed it); alpha (N=1) is an fp32 dot product in the epilogue of the
//     last trunk layer; rgb is an N=16 MMA.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "tc_common.cuh"

namespace nb {

// ------------------------------------------------------------------------------------ weight packing
__device__ __forceinline__ void pack_weights_body(const float* __restrict__ params, const PackSeg& sg,
                                                  unsigned char* __restrict__ out, int idx) {
  if (idx >= sg.n_rows * 8) return;
  const int r = idx >> 3, c = idx & 7;
  const int n = sg.row0 + r;
  const float* src = params + sg.src_off + (long long)r * sg.ld + sg.col0;
  float v[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const int k = c * 8 + e;
    // element 63 multiplies the encoding block's constant-one channel: the layer's bias rides along in the MMA
    v[e] = (k < sg.ncols ? src[k] : (k == 63 && sg.bias_off >= 0 ? params[sg.bias_off + r] : 0.0f)) * sg.scale;
  }
  uint4 pk;
  pk.x = pack_h2(v[0], v[1], false);
  pk.y = pack_h2(v[2], v[3], false);
  pk.z = pack_h2(v[4], v[5], false);
  pk.w = pack_h2(v[6], v[7], false);
  // K-block = two half-blocks of 32 elements: [chunk_rows x 64 B] each, SWIZZLE_64B (16-byte chunk ^= (row >> 1) & 3)
  const int h = c >> 2, c4 = c & 3;
  *reinterpret_cast<uint4*>(out + sg.dst_off + (long long)h * sg.chunk_rows * 64 + (long long)n * 64 +
                            ((c4 ^ ((n >> 1) & 3)) << 4)) = pk;
}

__global__ void pack_weights_kernel(const float* __restrict__ params, const PackSeg* __restrict__ segs,
                                    unsigned char* __restrict__ out) {
  pack_weights_body(params, segs[blockIdx.y], out, blockIdx.x * blockDim.x + threadIdx.x);
}

__global__ void pack_bias_kernel(const float* __restrict__ params, const BiasSeg* __restrict__ segs,
                                 float* __restrict__ out) {
  const BiasSeg sg = segs[blockIdx.x];
  for (int i = threadIdx.x; i < sg.count; i += blockDim.x) out[sg.dst + i] = params[sg.src_off + i];
}



// One launch for everything a parameter update invalidates: forward image, biases and (train) the transposed image of
// the dgrad chain.  blockIdx.y walks [forward segments | bias segments | transposed segments].
__global__ void pack_all_kernel(const float* __restrict__ params, const PackSeg* __restrict__ segs, int n_segs,
                                unsigned char* __restrict__ out, const BiasSeg* __restrict__ bsegs, int n_bsegs,
                                float* __restrict__ bias_out, const PackSegT* __restrict__ segs_t, int n_segs_t,
                                unsigned char* __restrict__ out_t) {
  int y = blockIdx.y;
  if (y < n_segs) {
    pack_weights_body(params, segs[y], out, blockIdx.x * blockDim.x + threadIdx.x);
    return;
  }
  y -= n_segs;
  if (y < n_bsegs) {
    if (blockIdx.x == 0) {
      const BiasSeg sg = bsegs[y];
      for (int i = threadIdx.x; i < sg.count; i += blockDim.x) bias_out[sg.dst + i] = params[sg.src_off + i];
    }
    return;
  }
  y -= n_bsegs;
  if (y < n_segs_t) pack_weights_t_body(params, segs_t[y], out_t, blockIdx.x * blockDim.x + threadIdx.x);
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

// Encode one 3-vector into 64 fp16 channels [v, sin(2^0 v), cos(2^0 v), ...] and store 16-byte chunks
// [kChunk0, kChunk0 + kChunks) of them as row `row` of a SWIZZLE_128B K-block.  Higher octaves by the double-angle
// recurrence from one accurate sincosf per component: the absolute error doubles per octave (<= 2^9 * 6e-8), far
// below the fp16 quantum (4.9e-4) of the stored value.
template <int kChunk0, int kChunks>
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
      // explicit roundings: left to the compiler, the contraction into FMAs differs between instantiations (kernels), and a
      // last-bit difference at octave 0 is 2^9 ulp at octave 9 -- enough to flip fp16 roundings in ~10 % of the points
      const float s2 = __fmul_rn(__fmul_rn(2.0f, s[k]), c[k]);
      const float c2 = __fmaf_rn(__fmul_rn(-2.0f, s[k]), s[k], 1.0f);
      s[k] = s2; c[k] = c2;
    }
  }
  ch[63] = 1.0f;  // constant-one channel: element 63 of a layer's encoding weights holds its bias (pack_weights_body)
  unsigned char* rp = block + row * 128;
#pragma unroll
  for (int q = kChunk0; q < kChunk0 + kChunks; ++q) {
    uint4 pk;
    pk.x = pack_h2(ch[8 * q + 0], ch[8 * q + 1], false);
    pk.y = pack_h2(ch[8 * q + 2], ch[8 * q + 3], false);
    pk.z = pack_h2(ch[8 * q + 4], ch[8 * q + 5], false);
    pk.w = pack_h2(ch[8 * q + 6], ch[8 * q + 7], false);
    *reinterpret_cast<uint4*>(rp + ((q ^ (row & 7)) << 4)) = pk;
  }
}

// shared-memory accesses of the epilogue by shared-space address; volatile keeps the bias loads where they are
// written (after tcgen05.wait::ld) instead of hoisted above the accumulator load, which would spill
__device__ __forceinline__ float4 lds_f4(uint32_t saddr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(saddr));
  return v;
}
__device__ __forceinline__ void sts_u4(uint32_t saddr, const uint4& v) {
  asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(saddr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
// accumulator columns [c0, c0+32) of one row (+ bias, optional ReLU) -> 16-bit -> four swizzled 16-byte chunks of the
// row's 128-byte line at rp; kAlpha also accumulates <activation, alpha weights> in fp32
template <bool kAlpha, int kRelu = -1>  // kRelu: 0 / 1 = compile-time activation choice, -1 = the run-time flag
__device__ __forceinline__ void epi_convert32(const uint32_t (&v)[64], int c0, uint32_t rp, uint32_t wa, int sw,
                                              bool relu_rt, float& adot) {
  const bool relu = kRelu < 0 ? relu_rt : kRelu != 0;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    const int e0 = c0 + 8 * c;
    const float a0 = __uint_as_float(v[e0 + 0]), a1 = __uint_as_float(v[e0 + 1]);
    const float a2 = __uint_as_float(v[e0 + 2]), a3 = __uint_as_float(v[e0 + 3]);
    const float a4 = __uint_as_float(v[e0 + 4]), a5 = __uint_as_float(v[e0 + 5]);
    const float a6 = __uint_as_float(v[e0 + 6]), a7 = __uint_as_float(v[e0 + 7]);
    if (kAlpha) {
      const float4 w0 = lds_f4(wa + 4 * e0), w1 = lds_f4(wa + 4 * e0 + 16);
      const float lo = relu ? 0.0f : -3.0e38f;
      adot = fmaf(fmaxf(a0, lo), w0.x, adot); adot = fmaf(fmaxf(a1, lo), w0.y, adot);
      adot = fmaf(fmaxf(a2, lo), w0.z, adot); adot = fmaf(fmaxf(a3, lo), w0.w, adot);
      adot = fmaf(fmaxf(a4, lo), w1.x, adot); adot = fmaf(fmaxf(a5, lo), w1.y, adot);
      adot = fmaf(fmaxf(a6, lo), w1.z, adot); adot = fmaf(fmaxf(a7, lo), w1.w, adot);
    }
    uint4 pk;
    pk.x = pack_h2(a0, a1, relu);
    pk.y = pack_h2(a2, a3, relu);
    pk.z = pack_h2(a4, a5, relu);
    pk.w = pack_h2(a6, a7, relu);
    sts_u4(rp + ((((c0 >> 3) + c) ^ sw) << 4), pk);
  }
}


// flags of the MMA issuer's step program
enum : uint32_t { PF_WAIT_ACT = 4u, PF_WAIT_ENCD = 8u, PF_FIRST = 16u, PF_LAST = 32u, PF_ACC = 64u, PF_BIAS = 128u };
// constant upper halves of the shared-memory operand descriptors (see umma_desc / umma_desc_sw64)
constexpr uint64_t kDescHiSw128 = ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
constexpr uint64_t kDescHiSw64 = ((uint64_t)1 << 16) | ((uint64_t)(512 >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)4 << 61);

// The MMA issuers' step program, built in shared memory at kernel start: one entry per ring slot.  x = byte offset of the A
// operand half-block inside a tile's region, y = instruction descriptor, z = flags (PF_*; bits 0-1 = activation block to
// wait for), w = X-record block the awaited operand is saved to (train launches).  Lane s of ONE warp builds stage s from
// the plan's step0 / encd_wait (a single thread walking all ~90 entries sat on every launch's critical path before the
// first __syncthreads: ~10 us of the fixed cost per launch measured by tools/fixed_cost_sweep.py).
template <bool kTwoCta>
__device__ __forceinline__ void build_step_program(const TcPlan& plan, uint4* prog, int s) {
  if (s >= plan.n_stages) return;
  const TcStage& st = plan.st[s];
  int i = st.step0;
  bool encd_pending = st.encd_wait != 0;  // the encoding block was rewritten by an earlier stage: its next reader waits for bar_encd
  for (int kb = 0; kb < st.n_kb; ++kb) {
    const int sidx = st.a_src[kb];
    // SRC_BIAS: one K=16 MMA over channels 48..63 of the encoding block (channel 63 == 1) against a chunk that is
    // zero except for the biases in element 63 -- only the second half-block of that chunk is ever loaded
    for (int h = sidx == SRC_BIAS ? 1 : 0; h < 2; ++h) {
      uint4 e;
      e.x = (uint32_t)((sidx < 4 ? sidx : plan.n_act_blocks) * kBlockBytes + h * 64);
      e.y = kTwoCta ? umma_idesc2(st.n_main) : umma_idesc(st.n_main);
      e.z = sidx == SRC_BIAS ? PF_BIAS : 0u;
      e.w = 0u;
      if (h == 0 && sidx < 4) { e.z |= PF_WAIT_ACT | (uint32_t)sidx; e.w = (uint32_t)(plan.st[s - 1].rec_blk + sidx); }
      if ((sidx == SRC_ENCD || sidx == SRC_BIAS) && encd_pending) { e.z |= PF_WAIT_ENCD; e.w = 1u; encd_pending = false; }
      if (kb == 0 && h == 0) e.z |= PF_FIRST;
      if (kb == st.n_kb - 1 && h == 1) e.z |= PF_LAST;
      if ((kb | h) != 0) e.z |= PF_ACC;
      prog[i++] = e;
    }
  }
}

// ------------------------------------------------------------------------------------ the kernel
__global__ void __launch_bounds__(kThreadsTc, 1)
mlp_tc_kernel(const __grid_constant__ TcPlan plan, const unsigned char* __restrict__ wpk, const float* __restrict__ bias_g,
              TcSrc src, long long n_points, float* __restrict__ raw, unsigned char* __restrict__ rec,
              unsigned char* __restrict__ mrec, int dbg) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  // carve (all tile regions 1024-byte aligned for SWIZZLE_128B)
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int tile_blocks = plan.n_act_blocks + 1;             // activation blocks + the encoding block
  const int tile_bytes = tile_blocks * kBlockBytes;
  unsigned char* act = smem;                                 // [2 tiles][tile_blocks][16 KB]
  unsigned char* ring = act + 2 * tile_bytes;
  float* bias_s = reinterpret_cast<float*>(ring + kFwdRing * kFwdSlotBytes);
  float* alpha_part = bias_s + ((plan.n_bias + 3) & ~3);     // [2 tiles][2 parities][128 rows]
  uint64_t* bars = reinterpret_cast<uint64_t*>(alpha_part + 2 * 2 * kTileM);
  uint64_t* bar_full = bars;                   // [kFwdRing]
  uint64_t* bar_empty = bar_full + kFwdRing;   // [kFwdRing]
  uint64_t* bar_enc = bar_empty + kFwdRing;    // [2]  position encoding of tile t written (256 arrivals)
  uint64_t* bar_encd = bar_enc + 2;            // [2]  direction encoding of tile t written (128)
  uint64_t* bar_act = bar_encd + 2;            // [2][4] activation block j of tile t written (128)
  uint64_t* bar_drain = bar_act + 8;           // [2]  accumulator of tile t read out by all 8 warps (256)
  uint64_t* bar_acc = bar_drain + 2;           // [2]  accumulator of tile t's current stage complete
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_acc + 2);
  uint4* prog = reinterpret_cast<uint4*>(tmem_slot + 4);   // [n_steps] MMA issuer program, one entry per ring slot

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long n_tiles = (n_points + kTileM - 1) / kTileM;
  // contiguous, balanced tile range of this CTA; taken two tiles at a time, a trailing single tile runs alone
  const long long tile_begin = n_tiles * blockIdx.x / gridDim.x;
  const long long tile_end = n_tiles * (blockIdx.x + 1) / gridDim.x;

  if (threadIdx.x == 0) {
    for (int i = 0; i < kFwdRing; ++i) {
      mbar_init(smem_u32(bar_full + i), 1);
      mbar_init(smem_u32(bar_empty + i), 2);   // one commit per MMA issuer warp
    }
    for (int t = 0; t < 2; ++t) {
      mbar_init(smem_u32(bar_acc + t), 1);
      mbar_init(smem_u32(bar_enc + t), 256);
      mbar_init(smem_u32(bar_encd + t), 128);
      for (int i = 0; i < 4; ++i) mbar_init(smem_u32(bar_act + 4 * t + i), 128);
      mbar_init(smem_u32(bar_drain + t), 256);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 3) build_step_program<false>(plan, prog, lane);
  if (warp == 1) {  // TMEM: all 512 columns (one CTA per SM): tile t owns columns [256 t, 256 t + 256)
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  for (int i = threadIdx.x; i < plan.n_bias; i += blockDim.x) bias_s[i] = bias_g[i];
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ============================ weight producer ============================
    if (lane == 0) {
      uint32_t slot = 0, phase = 0;
      const uint64_t pol_w = l2_evict_last();
      for (long long tile = tile_begin; tile < tile_end; tile += 2) {
        for (int s = 0; s < plan.n_stages; ++s) {
          const TcStage& st = plan.st[s];
          for (int kb = 0; kb < st.n_kb; ++kb) {
            for (int h = st.a_src[kb] == SRC_BIAS ? 1 : 0; h < 2; ++h) {
              const uint32_t bytes = (uint32_t)st.n_main * 64u;
              mbar_wait(smem_u32(bar_empty + slot), phase ^ 1);
              const uint32_t fb = smem_u32(bar_full + slot);
              if (NB_DBG(dbg & 1)) { mbar_arrive(fb); if (++slot == kFwdRing) { slot = 0; phase ^= 1; } continue; }
              mbar_expect_tx(fb, bytes);
              bulk_g2s_hint(smem_u32(ring + slot * kFwdSlotBytes),
                            wpk + st.w_off + (size_t)kb * st.chunk_bytes + (size_t)h * bytes, bytes, fb, pol_w);
              if (++slot == kFwdRing) { slot = 0; phase ^= 1; }
            }
          }
        }
      }
    }
  } else if (warp == 1 || warp == 2) {
    // ============================ MMA issuers (one warp per tile of the pair) ============================
    // The whole warp walks the per-pair step program (one entry per ring slot, built in shared memory at kernel
    // start) in warp-uniform control flow, so descriptors and phases live in uniform registers; only the tcgen05 /
    // bulk-copy instructions are predicated on one elected lane.  The issue loop has to stay far below the 256
    // tensor cycles one slot feeds per tile, or the tensor pipe starves (measured: tools/umma_bench.cu).
    const int t = warp - 1;
    uint32_t slot = 0, phase = 0;
    uint32_t ph_act = 0;     // bit j: parity of bar_act[t][j]
    uint32_t ph_enc = 0, ph_encd = 0, ph_drain = 0;
    bool armed = false;      // an epilogue of this tile slot has run: its accumulator hand-back must be awaited
    const uint32_t act_s = smem_u32(act) + t * tile_bytes, ring_s = smem_u32(ring);
    const uint32_t bar_full_s = smem_u32(bar_full), bar_empty_s = smem_u32(bar_empty), bar_act_s = smem_u32(bar_act + 4 * t);
    const uint32_t bar_enc_s = smem_u32(bar_enc + t), bar_encd_s = smem_u32(bar_encd + t);
    const uint32_t bar_drain_s = smem_u32(bar_drain + t), bar_acc_s = smem_u32(bar_acc + t);
    const uint32_t enc_off = plan.n_act_blocks * kBlockBytes;
    const uint32_t d_tmem = tmem_base + t * 256;
    const int n_steps = plan.n_steps;
    const size_t rec_stride = (size_t)plan.rec_blocks * kBlockBytes;
    const uint64_t pol_rec = l2_evict_first();
    for (long long tile0 = tile_begin; tile0 < tile_end; tile0 += 2) {
      if (tile0 + t >= tile_end) {
        // trailing single tile: this issuer only keeps the ring's two-commit protocol going
        for (int i = 0; i < n_steps; ++i) {
          mbar_wait(bar_full_s + 8 * slot, phase);
          if (elect_one()) tc_commit(bar_empty_s + 8 * slot);
          __syncwarp();
          if (++slot == kFwdRing) { slot = 0; phase ^= 1; }
        }
        break;
      }
      // train: every operand block is also written to the per-tile X record, in the exact swizzled image the
      // backward kernels load (K-major for recompute-free dgrad masks, MN-major operand of the wgrad MMAs)
      unsigned char* rec_t = rec ? rec + (size_t)(tile0 + t) * rec_stride : nullptr;
      mbar_wait(bar_enc_s, ph_enc);
      ph_enc ^= 1;
      if (rec && !NB_DBG(dbg & 4) && elect_one()) {
        bulk_s2g_hint(rec_t, act_s + enc_off, kBlockBytes, pol_rec);
        bulk_commit();
      }
      __syncwarp();
      uint4 e = prog[0];
      for (int i = 0; i < n_steps; ++i) {
        const uint4 en = prog[i + 1 < n_steps ? i + 1 : i];  // prefetch
        const uint32_t fl = e.z;
        if (fl & (PF_WAIT_ACT | PF_WAIT_ENCD)) {
          if (fl & PF_WAIT_ACT) {
            const uint32_t j = fl & 3u;
            mbar_wait(bar_act_s + 8 * j, (ph_act >> j) & 1u);
            ph_act ^= 1u << j;
          } else {
            mbar_wait(bar_encd_s, ph_encd);
            ph_encd ^= 1;
          }
          if (rec && !NB_DBG(dbg & 4) && elect_one()) {
            bulk_s2g_hint(rec_t + (size_t)e.w * kBlockBytes, act_s + (e.x & ~127u), kBlockBytes, pol_rec);
            bulk_commit();
          }
          __syncwarp();
        }
        if ((fl & PF_FIRST) && armed) {  // the previous stage's epilogue has read this tile's accumulator out
          mbar_wait(bar_drain_s, ph_drain);
          ph_drain ^= 1;
        }
        armed = true;
        mbar_wait(bar_full_s + 8 * slot, phase);
        tc_fence_after();
        const uint64_t a_desc = kDescHiSw128 | (uint64_t)(((act_s + e.x) & 0x3FFFFu) >> 4);
        const uint64_t b_desc = kDescHiSw64 | (uint64_t)(((ring_s + slot * kFwdSlotBytes) & 0x3FFFFu) >> 4);
        if (elect_one()) {
          if (!(fl & PF_BIAS)) tc_mma(d_tmem, a_desc, b_desc, e.y, (fl & PF_ACC) ? 1u : 0u);
          tc_mma(d_tmem, a_desc + 2, b_desc + 2, e.y, 1u);
          tc_commit(bar_empty_s + 8 * slot);  // (1 of 2 arrivals) frees the ring slot when these MMAs have read it
          if (fl & PF_LAST) {
            if (rec && !NB_DBG(dbg & 8)) bulk_wait_read0();  // record stores have read the blocks this stage's epilogue overwrites
            tc_commit(bar_acc_s);        // accumulator of this stage complete
          }
        }
        __syncwarp();
        if (++slot == kFwdRing) { slot = 0; phase ^= 1; }
        e = en;
      }
    }
    if (rec && elect_one()) bulk_wait0();
    __syncwarp();
  } else if (warp >= 4) {
    // ============================ epilogue warps (thread == row) ============================
    const int e = warp - 4;
    const int t = e >> 3;               // tile of the pair
    const int par = (e >> 2) & 1;       // handles 64-column blocks j == par (mod 2)
    const int q = warp & 3;             // TMEM lane quarter this warp may access
    const int row = q * 32 + lane;      // row of the tile
    const int sw = row & 7;
    const uint32_t t_main = tmem_base + (((uint32_t)(q * 32)) << 16) + t * 256;
    unsigned char* act_t = act + t * tile_bytes;
    unsigned char* enc_t = act_t + plan.n_act_blocks * kBlockBytes;
    float* apart = alpha_part + (t * 2 + par) * kTileM;
    uint32_t acc_phase = 0;
    for (long long tile0 = tile_begin; tile0 < tile_end; tile0 += 2) {
      const long long tile = tile0 + t;
      if (tile >= tile_end) break;      // trailing single tile: the second half of the CTA sits it out
      long long p = tile * kTileM + row;
      const bool in_range = p < n_points;
      if (!in_range) p = n_points - 1;
      // ---- positional encoding of this row's point (the two parities write 4 chunks of the row each)
      float v0 = 0.f, v1 = 0.f, v2 = 0.f;
      {
        float x0, x1, x2;
        if (src.rays != nullptr) {
          const long long r = p / src.S;
          const float* ry = src.rays + r * src.stride;
          const float zz = src.z[p];
          x0 = __fadd_rn(ry[0], __fmul_rn(ry[3], zz));
          x1 = __fadd_rn(ry[1], __fmul_rn(ry[4], zz));
          x2 = __fadd_rn(ry[2], __fmul_rn(ry[5], zz));
          if (plan.Cd > 0 && par == 1) { v0 = ry[8]; v1 = ry[9]; v2 = ry[10]; }
        } else {
          x0 = src.pts[3 * p]; x1 = src.pts[3 * p + 1]; x2 = src.pts[3 * p + 2];
          if (plan.Cd > 0 && par == 1) {
            const float* vv = src.viewdirs + (p / src.per_dir) * 3;
            v0 = vv[0]; v1 = vv[1]; v2 = vv[2];
          }
        }
        if (par == 0) encode_row<0, 4>(enc_t, row, x0, x1, x2, plan.Lx);
        else encode_row<4, 4>(enc_t, row, x0, x1, x2, plan.Lx);
      }
      fence_async_smem();
      mbar_arrive(smem_u32(bar_enc + t));

      float alpha_raw = 0.0f;
      for (int s = 0; s < plan.n_stages; ++s) {
        const TcStage& st = plan.st[s];
        mbar_wait(smem_u32(bar_acc + t), acc_phase);
        acc_phase ^= 1;
        tc_fence_after();
        const float* bs = bias_s + st.b_off;
        if (st.epi == EPI_HIDDEN) {
          const bool relu = st.relu != 0;
          const int nblk = st.n_act >> 6;
          float adot = 0.0f;
          if (par >= nblk) {  // nothing to read in this stage
            tc_fence_before();
            mbar_arrive(smem_u32(bar_drain + t));
          }
          for (int j = par; j < nblk; j += 2) {
            uint32_t v[64];
            NB_TMEM_LD32(t_main + j * 64, v);
            {
              uint32_t* v2 = v + 32;
              NB_TMEM_LD32(t_main + j * 64 + 32, v2);
            }
            NB_TMEM_WAIT64(v);
            if (j + 2 >= nblk) {  // last accumulator read of this stage: hand the TMEM columns back
              tc_fence_before();
              mbar_arrive(smem_u32(bar_drain + t));
            }
            if (NB_DBG(dbg & 2)) { fence_async_smem(); mbar_arrive(smem_u32(bar_act + 4 * t + j)); continue; }
            if (mrec != nullptr && relu && !NB_DBG(dbg & 16)) {
              // train: 1-bit ReLU mask of this row's 64 pre-activations for the dgrad chain (8 B per row instead of the
              // 128-byte activation row it used to re-read): element i of the block -> bit 31 - (i & 31) of word i >> 5,
              // set where the accumulator's sign bit is clear (funnel shifts collect the sign bits, one per value)
              uint32_t n0 = 0u, n1 = 0u;
#pragma unroll
              for (int i = 0; i < 32; ++i) {
                n0 = __funnelshift_l(v[i], n0, 1);
                n1 = __funnelshift_l(v[32 + i], n1, 1);
              }
              *reinterpret_cast<uint2*>(mrec + (((size_t)tile * (plan.rec_blocks - 2) + (st.rec_blk - 2 + j)) * kTileM + row) * 8) =
                  make_uint2(~n0, ~n1);
            }
            const uint32_t rp = smem_u32(act_t) + j * kBlockBytes + row * 128;
            if (st.alpha_dot) {
              const uint32_t wa = smem_u32(bias_s) + (plan.alpha_w_off + j * 64) * 4;
              epi_convert32<true>(v, 0, rp, wa, sw, relu, adot);
              epi_convert32<true>(v, 32, rp, wa, sw, relu, adot);
            } else if (relu) {  // the common case without a branch per 16-byte chunk
              epi_convert32<false, 1>(v, 0, rp, 0u, sw, true, adot);
              epi_convert32<false, 1>(v, 32, rp, 0u, sw, true, adot);
            } else {
              epi_convert32<false, 0>(v, 0, rp, 0u, sw, false, adot);
              epi_convert32<false, 0>(v, 32, rp, 0u, sw, false, adot);
            }
            fence_async_smem();
            mbar_arrive(smem_u32(bar_act + 4 * t + j));
          }
          if (st.alpha_dot) apart[row] = adot;
          if (st.write_encd && par == 1) {
            // every MMA reading the position encoding has completed (bar_acc of this stage): the block now
            // becomes the direction encoding for the view-dependent layer
            encode_row<0, 8>(enc_t, row, v0, v1, v2, plan.Ld);
            fence_async_smem();
            mbar_arrive(smem_u32(bar_encd + t));
          }
        } else {
          uint32_t v[4] = {0u, 0u, 0u, 0u};
          if (par == 0) {
            tmem_ld4(t_main, v);
            tc_wait_ld();
          }
          tc_fence_before();
          mbar_arrive(smem_u32(bar_drain + t));
          if (par == 0) {
            float4 o;
            // the heads read kActScale * h: undo the scale in fp32, then add the (unscaled, fp32) bias
            o.x = __uint_as_float(v[0]) * kInvActScale + bs[0];
            o.y = __uint_as_float(v[1]) * kInvActScale + bs[1];
            o.z = __uint_as_float(v[2]) * kInvActScale + bs[2];
            if (st.epi == EPI_RGB) {
              const float* ap = alpha_part + (t * 2) * kTileM;
              alpha_raw = (ap[row] + ap[kTileM + row]) * kInvActScale + bias_s[plan.alpha_b_off];
              o.w = alpha_raw;
            } else {
              o.w = __uint_as_float(v[3]) * kInvActScale + bs[3];
            }
            if (in_range) reinterpret_cast<float4*>(raw)[tile * kTileM + row] = o;
          }
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

// ------------------------------------------------------------------------------------ four-tile kernel for narrow graphs
// Inference kernel for graphs whose stages are all N <= 128 (FlexibleNeRFModel 4x128: every checkpoint the reference
// ships).  A 128-wide layer is 9 MMAs of ~80 cycles against the same per-stage hand-off chain as a 256-wide one
// (commit -> epilogue wakes -> tcgen05.ld -> convert -> publish -> next MMA, ~2 000 cycles), so two tiles in flight leave
// the tensor pipe idle most of the time (0.24 of peak).  Narrow tiles are cheap, though: 3 operand blocks (48 KB) and 128
// TMEM columns each, so FOUR fit in one CTA.  Same roles and step program as mlp_tc_kernel with
//   warp 0        weight producer (ring of 8 KB slots: <= 128 rows x 32 K-elements)
//   warps 1-4     MMA issuers, one per tile slot
//   warps 5-20    epilogue, four warps per tile (one per TMEM lane quarter, thread == row, all column blocks)
constexpr int kThreadsTc4 = 672;
constexpr int kRing4 = 3;
constexpr int kSlot4Bytes = 128 * 64;

__global__ void __launch_bounds__(kThreadsTc4, 1)
mlp_tc4_kernel(const __grid_constant__ TcPlan plan, const unsigned char* __restrict__ wpk, const float* __restrict__ bias_g,
               TcSrc src, long long n_points, float* __restrict__ raw) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int tile_blocks = plan.n_act_blocks + 1;
  const int tile_bytes = tile_blocks * kBlockBytes;
  unsigned char* act = smem;                                 // [4 tiles][tile_blocks][16 KB]
  unsigned char* ring = act + 4 * tile_bytes;                // [kRing4][8 KB]
  float* bias_s = reinterpret_cast<float*>(ring + kRing4 * kSlot4Bytes);
  uint64_t* bars = reinterpret_cast<uint64_t*>(bias_s + ((plan.n_bias + 3) & ~3));
  uint64_t* bar_full = bars;                    // [kRing4]
  uint64_t* bar_empty = bar_full + kRing4;      // [kRing4]  four commits (one per issuer) free a slot
  uint64_t* bar_enc = bar_empty + kRing4;       // [4]  position encoding of tile t written (128 arrivals)
  uint64_t* bar_encd = bar_enc + 4;             // [4]  direction encoding (128)
  uint64_t* bar_act = bar_encd + 4;             // [4][2] activation block j of tile t written (128)
  uint64_t* bar_drain = bar_act + 8;            // [4]  accumulator of tile t read out (128)
  uint64_t* bar_acc = bar_drain + 4;            // [4]  accumulator of tile t's current stage complete
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_acc + 4);
  uint4* prog = reinterpret_cast<uint4*>(tmem_slot + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long n_tiles = (n_points + kTileM - 1) / kTileM;
  const long long tile_begin = n_tiles * blockIdx.x / gridDim.x;
  const long long tile_end = n_tiles * (blockIdx.x + 1) / gridDim.x;
  const long long n_rounds = (tile_end - tile_begin + 3) >> 2;

  if (threadIdx.x == 0) {
    for (int i = 0; i < kRing4; ++i) {
      mbar_init(smem_u32(bar_full + i), 1);
      mbar_init(smem_u32(bar_empty + i), 4);
    }
    for (int t = 0; t < 4; ++t) {
      mbar_init(smem_u32(bar_acc + t), 1);
      mbar_init(smem_u32(bar_enc + t), 128);
      mbar_init(smem_u32(bar_encd + t), 128);
      mbar_init(smem_u32(bar_act + 2 * t), 128);
      mbar_init(smem_u32(bar_act + 2 * t + 1), 128);
      mbar_init(smem_u32(bar_drain + t), 128);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) build_step_program<false>(plan, prog, lane);
  if (warp == 1) {  // TMEM: all 512 columns; tile slot t owns columns [128 t, 128 t + 128)
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  for (int i = threadIdx.x; i < plan.n_bias; i += blockDim.x) bias_s[i] = bias_g[i];
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int n_steps = plan.n_steps;

  if (warp == 0) {
    // ============================ weight producer ============================
    if (lane == 0) {
      uint32_t slot = 0, phase = 0;
      const uint64_t pol_w = l2_evict_last();
      for (long long rd = 0; rd < n_rounds; ++rd) {
        for (int s = 0; s < plan.n_stages; ++s) {
          const TcStage& st = plan.st[s];
          for (int kb = 0; kb < st.n_kb; ++kb) {
            for (int h = st.a_src[kb] == SRC_BIAS ? 1 : 0; h < 2; ++h) {
              const uint32_t bytes = (uint32_t)st.n_main * 64u;
              mbar_wait(smem_u32(bar_empty + slot), phase ^ 1);
              const uint32_t fb = smem_u32(bar_full + slot);
              mbar_expect_tx(fb, bytes);
              bulk_g2s_hint(smem_u32(ring + slot * kSlot4Bytes),
                            wpk + st.w_off + (size_t)kb * st.chunk_bytes + (size_t)h * bytes, bytes, fb, pol_w);
              if (++slot == kRing4) { slot = 0; phase ^= 1; }
            }
          }
        }
      }
    }
  } else if (warp <= 4) {
    // ============================ MMA issuers (one warp per tile slot) ============================
    const int t = warp - 1;
    uint32_t slot = 0, phase = 0;
    uint32_t ph_act = 0, ph_enc = 0, ph_encd = 0, ph_drain = 0;
    bool armed = false;
    const uint32_t act_s = smem_u32(act) + t * tile_bytes, ring_s = smem_u32(ring);
    const uint32_t bar_full_s = smem_u32(bar_full), bar_empty_s = smem_u32(bar_empty), bar_act_s = smem_u32(bar_act + 2 * t);
    const uint32_t bar_enc_s = smem_u32(bar_enc + t), bar_encd_s = smem_u32(bar_encd + t);
    const uint32_t bar_drain_s = smem_u32(bar_drain + t), bar_acc_s = smem_u32(bar_acc + t);
    const uint32_t d_tmem = tmem_base + t * 128;
    for (long long rd = 0; rd < n_rounds; ++rd) {
      if (tile_begin + 4 * rd + t >= tile_end) {
        // unpopulated slot of the last round: only keep the ring's four-commit protocol going
        for (int i = 0; i < n_steps; ++i) {
          mbar_wait(bar_full_s + 8 * slot, phase);
          if (elect_one()) tc_commit(bar_empty_s + 8 * slot);
          __syncwarp();
          if (++slot == kRing4) { slot = 0; phase ^= 1; }
        }
        continue;
      }
      mbar_wait(bar_enc_s, ph_enc);
      ph_enc ^= 1;
      uint4 e = prog[0];
      for (int i = 0; i < n_steps; ++i) {
        const uint4 en = prog[i + 1 < n_steps ? i + 1 : i];
        const uint32_t fl = e.z;
        if (fl & PF_WAIT_ACT) {
          const uint32_t j = fl & 3u;
          mbar_wait(bar_act_s + 8 * j, (ph_act >> j) & 1u);
          ph_act ^= 1u << j;
        } else if (fl & PF_WAIT_ENCD) {
          mbar_wait(bar_encd_s, ph_encd);
          ph_encd ^= 1;
        }
        if ((fl & PF_FIRST) && armed) {
          mbar_wait(bar_drain_s, ph_drain);
          ph_drain ^= 1;
        }
        armed = true;
        mbar_wait(bar_full_s + 8 * slot, phase);
        tc_fence_after();
        const uint64_t a_desc = kDescHiSw128 | (uint64_t)(((act_s + e.x) & 0x3FFFFu) >> 4);
        const uint64_t b_desc = kDescHiSw64 | (uint64_t)(((ring_s + slot * kSlot4Bytes) & 0x3FFFFu) >> 4);
        if (elect_one()) {
          if (!(fl & PF_BIAS)) tc_mma(d_tmem, a_desc, b_desc, e.y, (fl & PF_ACC) ? 1u : 0u);
          tc_mma(d_tmem, a_desc + 2, b_desc + 2, e.y, 1u);
          tc_commit(bar_empty_s + 8 * slot);
          if (fl & PF_LAST) tc_commit(bar_acc_s);
        }
        __syncwarp();
        if (++slot == kRing4) { slot = 0; phase ^= 1; }
        e = en;
      }
    }
  } else {
    // ============================ epilogue warps (thread == row), four per tile ============================
    const int ew = warp - 5;
    const int t = ew >> 2;
    const int q = warp & 3;             // TMEM lane quarter this warp may access
    const int row = q * 32 + lane;
    const int sw = row & 7;
    const uint32_t t_main = tmem_base + (((uint32_t)(q * 32)) << 16) + t * 128;
    unsigned char* act_t = act + t * tile_bytes;
    unsigned char* enc_t = act_t + plan.n_act_blocks * kBlockBytes;
    uint32_t acc_phase = 0;
    for (long long rd = 0; rd < n_rounds; ++rd) {
      const long long tile = tile_begin + 4 * rd + t;
      if (tile >= tile_end) break;
      long long p = tile * kTileM + row;
      const bool in_range = p < n_points;
      if (!in_range) p = n_points - 1;
      float v0 = 0.f, v1 = 0.f, v2 = 0.f;
      {
        float x0, x1, x2;
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
        encode_row<0, 8>(enc_t, row, x0, x1, x2, plan.Lx);
      }
      fence_async_smem();
      mbar_arrive(smem_u32(bar_enc + t));

      float alpha_dot = 0.0f;   // <last trunk activation row, alpha weights>, formed where that layer is converted
      for (int s = 0; s < plan.n_stages; ++s) {
        const TcStage& st = plan.st[s];
        mbar_wait(smem_u32(bar_acc + t), acc_phase);
        acc_phase ^= 1;
        tc_fence_after();
        const float* bs = bias_s + st.b_off;
        if (st.epi == EPI_HIDDEN) {
          const bool relu = st.relu != 0;
          const int nblk = st.n_act >> 6;
          // <activation, alpha weights>: one partial sum per column block, added at the end -- the summation order of the
          // two-tile kernel, where the two blocks belong to different warps
          float adot0 = 0.0f, adot = 0.0f;
          for (int j = 0; j < nblk; ++j) {
            if (j == 1) { adot0 = adot; adot = 0.0f; }
            uint32_t v[64];
            NB_TMEM_LD32(t_main + j * 64, v);
            {
              uint32_t* v2p = v + 32;
              NB_TMEM_LD32(t_main + j * 64 + 32, v2p);
            }
            NB_TMEM_WAIT64(v);
            if (j + 1 >= nblk) {  // last accumulator read of this stage: hand the TMEM columns back
              tc_fence_before();
              mbar_arrive(smem_u32(bar_drain + t));
            }
            const uint32_t rp = smem_u32(act_t) + j * kBlockBytes + row * 128;
            if (st.alpha_dot) {
              const uint32_t wa = smem_u32(bias_s) + (plan.alpha_w_off + j * 64) * 4;
              epi_convert32<true>(v, 0, rp, wa, sw, relu, adot);
              epi_convert32<true>(v, 32, rp, wa, sw, relu, adot);
            } else if (relu) {
              epi_convert32<false, 1>(v, 0, rp, 0u, sw, true, adot);
              epi_convert32<false, 1>(v, 32, rp, 0u, sw, true, adot);
            } else {
              epi_convert32<false, 0>(v, 0, rp, 0u, sw, false, adot);
              epi_convert32<false, 0>(v, 32, rp, 0u, sw, false, adot);
            }
            fence_async_smem();
            mbar_arrive(smem_u32(bar_act + 2 * t + j));
          }
          if (st.alpha_dot) alpha_dot = adot0 + adot;
          if (st.write_encd) {
            encode_row<0, 8>(enc_t, row, v0, v1, v2, plan.Ld);
            fence_async_smem();
            mbar_arrive(smem_u32(bar_encd + t));
          }
        } else {
          uint32_t v[4];
          tmem_ld4(t_main, v);
          tc_wait_ld();
          float4 o;
          o.x = __uint_as_float(v[0]) * kInvActScale + bs[0];
          o.y = __uint_as_float(v[1]) * kInvActScale + bs[1];
          o.z = __uint_as_float(v[2]) * kInvActScale + bs[2];
          if (st.epi == EPI_RGB) o.w = alpha_dot * kInvActScale + bias_s[plan.alpha_b_off];
          else o.w = __uint_as_float(v[3]) * kInvActScale + bs[3];
          tc_fence_before();
          mbar_arrive(smem_u32(bar_drain + t));
          if (in_range) reinterpret_cast<float4*>(raw)[tile * kTileM + row] = o;
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
  }
}

// ------------------------------------------------------------------------------------ CTA-pair forward kernel
// Same network, same roles, same step program as mlp_tc_kernel, for inference (no activation record), run by CLUSTERS OF
// TWO CTAs (two SMs of a TPC).  Each CTA still owns two 128-point tiles; MMA stream t joins tile t of both CTAs into
// one tcgen05.mma.cta_group::2 of M = 256: the leader CTA (cluster rank 0) issues it, every CTA supplies its own 128
// activation rows and HALF of the weight rows (N/2) from its own shared memory.  Per SM the B-operand reads and the
// weight ring fill are half of the single-CTA kernel's -- the shared-memory port is what that kernel saturates
//  -- and the ring holds twice as many slots.
//   * barriers the issuers wait on live in the leader; the peer CTA's epilogue warps arrive on them remotely
//     (mbarrier.arrive.release.cluster on the mapa address), one arrival per warp;
//   * a weight slot is ready when BOTH halves have landed: the peer's warp 1 relays its local "full" barrier to the
//     leader's peer_full barrier;
//   * tcgen05.commit multicasts its arrival to both CTAs (slot free, accumulator ready).
constexpr int kRing2 = 6;
constexpr int kSlot2Bytes = 128 * 64;   // half of a <= 256-row slot

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreadsTc, 1)
mlp_tc2_kernel(const __grid_constant__ TcPlan plan, const unsigned char* __restrict__ wpk, const float* __restrict__ bias_g,
               TcSrc src, long long n_points, float* __restrict__ raw) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int tile_blocks = plan.n_act_blocks + 1;
  const int tile_bytes = tile_blocks * kBlockBytes;
  unsigned char* act = smem;                                 // [2 tiles][tile_blocks][16 KB]
  unsigned char* ring = act + 2 * tile_bytes;                // [kRing2][8 KB]
  float* bias_s = reinterpret_cast<float*>(ring + kRing2 * kSlot2Bytes);
  float* alpha_part = bias_s + ((plan.n_bias + 3) & ~3);     // [2 tiles][2 parities][128 rows]
  uint64_t* bars = reinterpret_cast<uint64_t*>(alpha_part + 2 * 2 * kTileM);
  uint64_t* bar_full = bars;                    // [kRing2] this CTA's half of the slot has landed
  uint64_t* bar_pfull = bar_full + kRing2;      // [kRing2] (leader) the peer's half has landed
  uint64_t* bar_empty = bar_pfull + kRing2;     // [kRing2] both MMA streams have consumed the slot (multicast commits)
  uint64_t* bar_enc = bar_empty + kRing2;       // [2]  (leader) position encoding of stream t written in both CTAs (16 warps)
  uint64_t* bar_encd = bar_enc + 2;             // [2]  (leader) direction encoding (8 warps)
  uint64_t* bar_act = bar_encd + 2;             // [2][4] (leader) activation block j of stream t written (8 warps)
  uint64_t* bar_drain = bar_act + 8;            // [2]  (leader) accumulator of stream t read out (16 warps)
  uint64_t* bar_acc = bar_drain + 2;            // [2]  accumulator of stream t complete (multicast commit)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_acc + 2);
  uint4* prog = reinterpret_cast<uint4*>(tmem_slot + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const long long n_tiles = (n_points + kTileM - 1) / kTileM;
  const long long n_clusters = gridDim.x >> 1, cluster = blockIdx.x >> 1;
  // contiguous, balanced tile range of this cluster, taken four tiles at a time:
  // round tiles base + {0, 1, 2, 3} = (stream 0, CTA 0), (stream 0, CTA 1), (stream 1, CTA 0), (stream 1, CTA 1)
  const long long tile_begin = n_tiles * cluster / n_clusters;
  const long long tile_end = n_tiles * (cluster + 1) / n_clusters;

  if (threadIdx.x == 0) {
    for (int i = 0; i < kRing2; ++i) {
      mbar_init(smem_u32(bar_full + i), 1);
      mbar_init(smem_u32(bar_pfull + i), 1);
      mbar_init(smem_u32(bar_empty + i), 2);
    }
    for (int t = 0; t < 2; ++t) {
      mbar_init(smem_u32(bar_acc + t), 1);
      mbar_init(smem_u32(bar_enc + t), 16);
      mbar_init(smem_u32(bar_encd + t), 8);
      for (int i = 0; i < 4; ++i) mbar_init(smem_u32(bar_act + 4 * t + i), 8);
      mbar_init(smem_u32(bar_drain + t), 16);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 3) build_step_program<true>(plan, prog, lane);
  if (warp == 1) {  // TMEM: all 512 columns in both CTAs of the pair; stream t owns columns [256 t, 256 t + 256)
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
  }
  for (int i = threadIdx.x; i < plan.n_bias; i += blockDim.x) bias_s[i] = bias_g[i];
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();   // barriers of both CTAs are initialised before anybody arrives on a remote one
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int n_steps = plan.n_steps;
  // rounds of this cluster and which streams are populated in the last one
  const long long n_local = tile_end - tile_begin;
  const long long n_rounds = (n_local + 3) >> 2;

  if (warp == 0) {
    // ============================ weight producer: this CTA's half of every slot ============================
    if (lane == 0) {
      uint32_t slot = 0, phase = 0;
      const uint64_t pol_w = l2_evict_last();
      for (long long rd = 0; rd < n_rounds; ++rd) {
        for (int s = 0; s < plan.n_stages; ++s) {
          const TcStage& st = plan.st[s];
          const uint32_t half_rows_bytes = (uint32_t)(st.n_main >> 1) * 64u;   // N/2 rows x 32 K-elements
          for (int kb = 0; kb < st.n_kb; ++kb) {
            for (int h = st.a_src[kb] == SRC_BIAS ? 1 : 0; h < 2; ++h) {
              mbar_wait(smem_u32(bar_empty + slot), phase ^ 1);
              const uint32_t fb = smem_u32(bar_full + slot);
              mbar_expect_tx(fb, half_rows_bytes);
              bulk_g2s_hint(smem_u32(ring + slot * kSlot2Bytes),
                            wpk + st.w_off + (size_t)kb * st.chunk_bytes + (size_t)h * (st.n_main * 64u) + (size_t)rank * half_rows_bytes,
                            half_rows_bytes, fb, pol_w);
              if (++slot == kRing2) { slot = 0; phase ^= 1; }
            }
          }
        }
      }
    }
  } else if (warp == 1 && !leader) {
    // ============================ peer: relay "my half has landed" to the leader ============================
    if (lane == 0) {
      uint32_t slot = 0, phase = 0;
      const uint32_t pf0 = mapa_u32(smem_u32(bar_pfull), 0);
      for (long long it = 0; it < n_rounds * n_steps; ++it) {
        mbar_wait(smem_u32(bar_full + slot), phase);
        mbar_arrive_cluster(pf0 + 8 * slot);
        if (++slot == kRing2) { slot = 0; phase ^= 1; }
      }
    }
  } else if ((warp == 1 || warp == 2) && leader) {
    // ============================ MMA issuers of the pair (one warp per stream) ============================
    const int t = warp - 1;
    uint32_t slot = 0, phase = 0;
    uint32_t ph_act = 0, ph_enc = 0, ph_encd = 0, ph_drain = 0;
    bool armed = false;
    const uint32_t act_s = smem_u32(act) + t * tile_bytes, ring_s = smem_u32(ring);
    const uint32_t bar_full_s = smem_u32(bar_full), bar_pfull_s = smem_u32(bar_pfull), bar_empty_s = smem_u32(bar_empty);
    const uint32_t bar_act_s = smem_u32(bar_act + 4 * t);
    const uint32_t bar_enc_s = smem_u32(bar_enc + t), bar_encd_s = smem_u32(bar_encd + t);
    const uint32_t bar_drain_s = smem_u32(bar_drain + t), bar_acc_s = smem_u32(bar_acc + t);
    const uint32_t d_tmem = tmem_base + t * 256;
    for (long long rd = 0; rd < n_rounds; ++rd) {
      const bool active = tile_begin + 4 * rd + 2 * t < tile_end;   // the stream has at least its CTA-0 tile
      if (!active) {
        // unpopulated stream of the last round: only keep the ring's two-commit protocol going
        for (int i = 0; i < n_steps; ++i) {
          mbar_wait(bar_full_s + 8 * slot, phase);
          mbar_wait_cluster(bar_pfull_s + 8 * slot, phase);
          if (elect_one()) tc_commit2(bar_empty_s + 8 * slot);
          __syncwarp();
          if (++slot == kRing2) { slot = 0; phase ^= 1; }
        }
        continue;
      }
      mbar_wait_cluster(bar_enc_s, ph_enc);
      ph_enc ^= 1;
      uint4 e = prog[0];
      for (int i = 0; i < n_steps; ++i) {
        const uint4 en = prog[i + 1 < n_steps ? i + 1 : i];
        const uint32_t fl = e.z;
        if (fl & PF_WAIT_ACT) {
          const uint32_t j = fl & 3u;
          mbar_wait_cluster(bar_act_s + 8 * j, (ph_act >> j) & 1u);
          ph_act ^= 1u << j;
        } else if (fl & PF_WAIT_ENCD) {
          mbar_wait_cluster(bar_encd_s, ph_encd);
          ph_encd ^= 1;
        }
        if ((fl & PF_FIRST) && armed) {
          mbar_wait_cluster(bar_drain_s, ph_drain);
          ph_drain ^= 1;
        }
        armed = true;
        mbar_wait(bar_full_s + 8 * slot, phase);
        mbar_wait_cluster(bar_pfull_s + 8 * slot, phase);
        tc_fence_after();
        const uint64_t a_desc = kDescHiSw128 | (uint64_t)(((act_s + e.x) & 0x3FFFFu) >> 4);
        const uint64_t b_desc = kDescHiSw64 | (uint64_t)(((ring_s + slot * kSlot2Bytes) & 0x3FFFFu) >> 4);
        if (elect_one()) {
          if (!(fl & PF_BIAS)) tc_mma2(d_tmem, a_desc, b_desc, e.y, (fl & PF_ACC) ? 1u : 0u);
          tc_mma2(d_tmem, a_desc + 2, b_desc + 2, e.y, 1u);
          tc_commit2(bar_empty_s + 8 * slot);
          if (fl & PF_LAST) tc_commit2(bar_acc_s);
        }
        __syncwarp();
        if (++slot == kRing2) { slot = 0; phase ^= 1; }
        e = en;
      }
    }
  } else if (warp >= 4) {
    // ============================ epilogue warps (thread == row), both CTAs ============================
    const int ew = warp - 4;
    const int t = ew >> 3;
    const int par = (ew >> 2) & 1;
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const int sw = row & 7;
    const uint32_t t_main = tmem_base + (((uint32_t)(q * 32)) << 16) + t * 256;
    unsigned char* act_t = act + t * tile_bytes;
    unsigned char* enc_t = act_t + plan.n_act_blocks * kBlockBytes;
    float* apart = alpha_part + (t * 2 + par) * kTileM;
    // the issuers' barriers live in the leader CTA
    // (the leader's own warps use the plain shared::cta window of the same barriers: mapa of rank 0 from rank 0 is
    // the same object, but the shared::cluster form goes through the cluster path)
    const uint32_t l_enc = mapa_u32(smem_u32(bar_enc + t), 0), l_encd = mapa_u32(smem_u32(bar_encd + t), 0);
    const uint32_t l_act = mapa_u32(smem_u32(bar_act + 4 * t), 0), l_drain = mapa_u32(smem_u32(bar_drain + t), 0);
    const uint32_t c_enc = smem_u32(bar_enc + t), c_encd = smem_u32(bar_encd + t);
    const uint32_t c_act = smem_u32(bar_act + 4 * t), c_drain = smem_u32(bar_drain + t);
#define NB_ARRIVE2(local, remote) do { if (leader) mbar_arrive(local); else mbar_arrive_cluster(remote); } while (0)
    uint32_t acc_phase = 0;
    for (long long rd = 0; rd < n_rounds; ++rd) {
      const long long stream_tile0 = tile_begin + 4 * rd + 2 * t;
      if (stream_tile0 >= tile_end) break;          // this stream is not populated in the cluster's last round
      const long long tile = stream_tile0 + rank;   // may be one past the range: a dummy tile (clamped points, no store)
      const bool tile_real = tile < tile_end;
      long long p = tile * kTileM + row;
      const bool in_range = tile_real && p < n_points;
      if (p >= n_points) p = n_points - 1;
      float v0 = 0.f, v1 = 0.f, v2 = 0.f;
      {
        float x0, x1, x2;
        if (src.rays != nullptr) {
          const long long r = p / src.S;
          const float* ry = src.rays + r * src.stride;
          const float zz = src.z[p];
          x0 = __fadd_rn(ry[0], __fmul_rn(ry[3], zz));
          x1 = __fadd_rn(ry[1], __fmul_rn(ry[4], zz));
          x2 = __fadd_rn(ry[2], __fmul_rn(ry[5], zz));
          if (plan.Cd > 0 && par == 1) { v0 = ry[8]; v1 = ry[9]; v2 = ry[10]; }
        } else {
          x0 = src.pts[3 * p]; x1 = src.pts[3 * p + 1]; x2 = src.pts[3 * p + 2];
          if (plan.Cd > 0 && par == 1) {
            const float* vv = src.viewdirs + (p / src.per_dir) * 3;
            v0 = vv[0]; v1 = vv[1]; v2 = vv[2];
          }
        }
        if (par == 0) encode_row<0, 4>(enc_t, row, x0, x1, x2, plan.Lx);
        else encode_row<4, 4>(enc_t, row, x0, x1, x2, plan.Lx);
      }
      fence_async_smem();
      __syncwarp();
      if (lane == 0) NB_ARRIVE2(c_enc, l_enc);

      for (int s = 0; s < plan.n_stages; ++s) {
        const TcStage& st = plan.st[s];
        mbar_wait(smem_u32(bar_acc + t), acc_phase);
        acc_phase ^= 1;
        tc_fence_after();
        const float* bs = bias_s + st.b_off;
        if (st.epi == EPI_HIDDEN) {
          const bool relu = st.relu != 0;
          const int nblk = st.n_act >> 6;
          float adot = 0.0f;
          if (par >= nblk) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) NB_ARRIVE2(c_drain, l_drain);
          }
          for (int j = par; j < nblk; j += 2) {
            uint32_t v[64];
            NB_TMEM_LD32(t_main + j * 64, v);
            {
              uint32_t* v2p = v + 32;
              NB_TMEM_LD32(t_main + j * 64 + 32, v2p);
            }
            NB_TMEM_WAIT64(v);
            if (j + 2 >= nblk) {
              tc_fence_before();
              __syncwarp();
              if (lane == 0) NB_ARRIVE2(c_drain, l_drain);
            }
            const uint32_t rp = smem_u32(act_t) + j * kBlockBytes + row * 128;
            if (st.alpha_dot) {
              const uint32_t wa = smem_u32(bias_s) + (plan.alpha_w_off + j * 64) * 4;
              epi_convert32<true>(v, 0, rp, wa, sw, relu, adot);
              epi_convert32<true>(v, 32, rp, wa, sw, relu, adot);
            } else if (relu) {
              epi_convert32<false, 1>(v, 0, rp, 0u, sw, true, adot);
              epi_convert32<false, 1>(v, 32, rp, 0u, sw, true, adot);
            } else {
              epi_convert32<false, 0>(v, 0, rp, 0u, sw, false, adot);
              epi_convert32<false, 0>(v, 32, rp, 0u, sw, false, adot);
            }
            fence_async_smem();
            __syncwarp();
            if (lane == 0) NB_ARRIVE2(c_act + 8 * j, l_act + 8 * j);
          }
          if (st.alpha_dot) apart[row] = adot;
          if (st.write_encd && par == 1) {
            encode_row<0, 8>(enc_t, row, v0, v1, v2, plan.Ld);
            fence_async_smem();
            __syncwarp();
            if (lane == 0) NB_ARRIVE2(c_encd, l_encd);
          }
        } else {
          uint32_t v[4] = {0u, 0u, 0u, 0u};
          if (par == 0) {
            tmem_ld4(t_main, v);
            tc_wait_ld();
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) NB_ARRIVE2(c_drain, l_drain);
          if (par == 0) {
            float4 o;
            o.x = __uint_as_float(v[0]) * kInvActScale + bs[0];
            o.y = __uint_as_float(v[1]) * kInvActScale + bs[1];
            o.z = __uint_as_float(v[2]) * kInvActScale + bs[2];
            if (st.epi == EPI_RGB) {
              const float* ap = alpha_part + (t * 2) * kTileM;
              o.w = (ap[row] + ap[kTileM + row]) * kInvActScale + bias_s[plan.alpha_b_off];
            } else {
              o.w = __uint_as_float(v[3]) * kInvActScale + bs[3];
            }
            if (in_range) reinterpret_cast<float4*>(raw)[tile * kTileM + row] = o;
          }
        }
      }
    }
  }
#undef NB_ARRIVE2
  // ---- teardown: nobody leaves (or frees tensor memory) while the other CTA may still signal or be signalled
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 1) {
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
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
  if (g.Cx > 63 || g.Cd > 63) { *why = "more than 10 encoding frequencies"; return false; }
  if (g.skip >= g.depth - 1) { *why = "skip connection after the last trunk layer"; return false; }
  const int wb = g.width / 64;
  plan.n_act_blocks = std::max(wb, g.use_viewdirs ? g.feat_width / 64 : 0);
  plan.Cx = g.Cx; plan.Cd = g.Cd; plan.Lx = g.n_freq_xyz; plan.Ld = g.n_freq_dir;
  plan.alpha_w_off = plan.alpha_b_off = -1;
  plan.rec_blocks = 2;  // [encx, encd]
  size_t woff = 0;
  int boff = 0;
  auto begin_stage = [&](int n_main, int epi, int relu, int n_act) -> TcStage& {
    TcStage& st = plan.st[plan.n_stages++];
    st.n_kb = 0;
    st.n_main = n_main; st.chunk_bytes = n_main * 128;
    st.w_off = (int)woff; st.b_off = boff; st.epi = epi; st.relu = relu; st.n_act = n_act;
    st.alpha_dot = 0; st.write_encd = 0;
    st.rec_blk = plan.rec_blocks;
    plan.rec_blocks += n_act / 64;
    boff += (n_main + 3) & ~3;
    return st;
  };
  // adds one K-block to the stage
  auto add_kb = [&](TcStage& st, int a_src, PackSeg sg) {
    sg.dst_off = (long long)woff;
    sg.chunk_rows = st.n_main;
    sg.scale = (a_src == SRC_ENCX || a_src == SRC_ENCD || a_src == SRC_BIAS) ? kActScale : 1.0f;
    segs.push_back(sg);
    st.a_src[st.n_kb++] = a_src;
    woff += st.chunk_bytes;
  };
  // hidden layers get their bias through the MMA: as element 63 of the encoding weights where the layer reads an
  // encoding block anyway, else as one extra K=16 step (SRC_BIAS) at the end of the stage
  auto with_bias = [](PackSeg sg, const Lin& l) { sg.bias_off = l.b_off; return sg; };
  auto seg = [](const Lin& l, int row0, int col0, int ncols) {
    PackSeg s{};
    s.bias_off = -1;
    s.src_off = l.w_off; s.row0 = row0; s.n_rows = l.out; s.ld = l.in; s.col0 = col0; s.ncols = ncols;
    return s;
  };
  if (g.depth + 3 > kMaxStages) { *why = "too many layers"; return false; }
  int last_encx = 0;
  for (int i = 0; i < g.depth; ++i) {
    const Lin& l = g.lins[g.trunk[i]];
    TcStage& st = begin_stage(g.width, EPI_HIDDEN, !(i == 0 && !g.act0), g.width);
    bsegs.push_back(BiasSeg{st.b_off, l.out, l.b_off});
    if (i == 0) {
      add_kb(st, SRC_ENCX, with_bias(seg(l, 0, 0, g.Cx), l));
    } else {
      int off = 0;
      if (i - 1 == g.skip) {
        add_kb(st, SRC_ENCX, with_bias(seg(l, 0, 0, g.Cx), l));
        off = g.Cx;
        last_encx = i;
      }
      for (int j = 0; j < wb; ++j) add_kb(st, j, seg(l, 0, off + 64 * j, 64));
      if (off == 0) add_kb(st, SRC_BIAS, with_bias(seg(l, 0, 0, 0), l));
    }
  }
  if (g.use_viewdirs) {
    const Lin& la = g.lins[g.alpha];
    const Lin& lf = g.lins[g.feat];
    const Lin& ld = g.lins[g.dir];
    const Lin& lr = g.lins[g.rgb];
    if (g.dir_width % 64 != 0) { *why = "direction layer width must be a multiple of 64"; return false; }
    // alpha = <h_last, w_alpha> + b_alpha in the epilogue of the last trunk layer (fp32 weights next to the biases)
    plan.st[g.depth - 1].alpha_dot = 1;
    plan.alpha_w_off = boff;
    bsegs.push_back(BiasSeg{boff, g.width, la.w_off});
    boff += g.width;
    plan.alpha_b_off = boff;
    bsegs.push_back(BiasSeg{boff, 1, la.b_off});
    boff += 4;
    // the position-encoding block is dead after its last consumer: that stage's epilogue writes the direction encoding
    plan.st[last_encx].write_encd = g.Cd > 0 ? 1 : 0;
    TcStage& sh = begin_stage(g.feat_width, EPI_HIDDEN, g.feat_act ? 1 : 0, g.feat_width);
    bsegs.push_back(BiasSeg{sh.b_off, lf.out, lf.b_off});
    for (int j = 0; j < wb; ++j) add_kb(sh, j, seg(lf, 0, 64 * j, 64));
    add_kb(sh, SRC_BIAS, with_bias(seg(lf, 0, 0, 0), lf));
    TcStage& sd = begin_stage(g.dir_width, EPI_HIDDEN, 1, g.dir_width);
    bsegs.push_back(BiasSeg{sd.b_off, ld.out, ld.b_off});
    for (int j = 0; j < g.feat_width / 64; ++j) add_kb(sd, j, seg(ld, 0, 64 * j, 64));
    if (g.Cd > 0) add_kb(sd, SRC_ENCD, with_bias(seg(ld, 0, g.feat_width, g.Cd), ld));
    else add_kb(sd, SRC_BIAS, with_bias(seg(ld, 0, 0, 0), ld));
    TcStage& sr = begin_stage(16, EPI_RGB, 0, 0);
    bsegs.push_back(BiasSeg{sr.b_off, 3, lr.b_off});
    for (int j = 0; j < g.dir_width / 64; ++j) add_kb(sr, j, seg(lr, 0, 64 * j, 64));
  } else {
    const Lin& lo = g.lins[g.out4];
    TcStage& so = begin_stage(16, EPI_OUT4, 0, 0);
    bsegs.push_back(BiasSeg{so.b_off, 4, lo.b_off});
    for (int j = 0; j < wb; ++j) add_kb(so, j, seg(lo, 0, 64 * j, 64));
  }
  for (int s = 0; s < plan.n_stages; ++s)
    if (plan.st[s].n_kb > kMaxKB) { *why = "too many K blocks"; return false; }
  plan.n_bias = boff;
  // entry counts of the step program; each stage's first entry and its "wait for the direction encoding" flag, so that the
  // kernels can build the program one stage per lane
  plan.n_steps = 0;
  bool encd_pending = false;
  for (int s = 0; s < plan.n_stages; ++s) {
    TcStage& st = plan.st[s];
    st.step0 = plan.n_steps;
    st.encd_wait = 0;
    for (int kb = 0; kb < st.n_kb; ++kb) {
      const int sidx = st.a_src[kb];
      if ((sidx == SRC_ENCD || sidx == SRC_BIAS) && encd_pending) { st.encd_wait = 1; encd_pending = false; }
      plan.n_steps += sidx == SRC_BIAS ? 1 : 2;
    }
    if (st.write_encd) encd_pending = true;
  }
  w_bytes = woff;
  return true;
}

static size_t tc_smem_bytes(const TcPlan& plan) {
  size_t b = 1024;  // alignment slack
  b += (size_t)2 * (plan.n_act_blocks + 1) * kBlockBytes;
  b += (size_t)kFwdRing * kFwdSlotBytes;
  b += (size_t)((plan.n_bias + 3) & ~3) * sizeof(float);
  b += (size_t)2 * 2 * kTileM * sizeof(float);
  b += (2 * kFwdRing + 2 + 2 + 8 + 2 + 2) * sizeof(uint64_t) + 16;
  b += (size_t)plan.n_steps * sizeof(uint4);
  return b;
}

static size_t tc2_smem_bytes(const TcPlan& plan) {
  size_t b = 1024;  // alignment slack
  b += (size_t)2 * (plan.n_act_blocks + 1) * kBlockBytes;
  b += (size_t)kRing2 * kSlot2Bytes;
  b += (size_t)((plan.n_bias + 3) & ~3) * sizeof(float);
  b += (size_t)2 * 2 * kTileM * sizeof(float);
  b += (3 * kRing2 + 2 + 2 + 8 + 2 + 2) * sizeof(uint64_t) + 16;
  b += (size_t)plan.n_steps * sizeof(uint4);
  return b;
}

static size_t tc4_smem_bytes(const TcPlan& plan) {
  size_t b = 1024;
  b += (size_t)4 * (plan.n_act_blocks + 1) * kBlockBytes;
  b += (size_t)kRing4 * kSlot4Bytes;
  b += (size_t)((plan.n_bias + 3) & ~3) * sizeof(float);
  b += (2 * kRing4 + 4 + 4 + 8 + 4 + 4) * sizeof(uint64_t) + 16;
  b += (size_t)plan.n_steps * sizeof(uint4);
  return b;
}
// graphs the four-tile kernel takes: at most two activation blocks per tile and every stage N <= 128
static bool tc4_eligible(const TcPlan& plan) {
  if (plan.n_act_blocks > 2 || tc4_smem_bytes(plan) > 227 * 1024) return false;
  for (int s = 0; s < plan.n_stages; ++s)
    if (plan.st[s].n_main > 128) return false;
  return true;
}
static int g_fwd_4tile = 1;

// NB_FWD_2CTA: 1 = inference launches use the CTA-pair kernel (mlp_tc2_kernel), 0 = the single-CTA kernel
static int g_fwd_2cta = -1;
static int fwd_2cta_mode() {
  if (g_fwd_2cta < 0) g_fwd_2cta = getenv("NB_FWD_2CTA") ? atoi(getenv("NB_FWD_2CTA")) : 0;
  return g_fwd_2cta;
}
int set_fwd_2cta(int v) {
  const int old = fwd_2cta_mode();
  g_fwd_2cta = v != 0;
  return old;
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
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(mlp_tc2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(mlp_tc4_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
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
  // forward image + biases + (when the graph trains on tensor cores) the transposed image, in one launch
  const int n_t = p->bw_supported ? p->n_segs_t : 0;
  dim3 grid((unsigned)ceil_div(272 * 8, 256), (unsigned)(p->n_segs + p->n_bsegs + n_t));
  pack_all_kernel<<<grid, 256, 0, stream>>>(m->params, p->d_segs, p->n_segs, p->d_w, p->d_bsegs, p->n_bsegs, p->d_bias,
                                            p->d_segs_t, n_t, p->d_wt);
  NB_LAUNCHED();
  p->wt_fresh = p->bw_supported;
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
  unsigned char* mrec = nullptr;
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
    mrec = tc_train_mrec(p, ws, n_points);
  }
  TcSrc src{rays, ray_stride, z_vals, n_samples > 0 ? n_samples : 1, pts, viewdirs, points_per_dir > 0 ? points_per_dir : 1};
  const int64_t n_tiles = ceil_div(n_points, kTileM);
  if (rec == nullptr && g_fwd_4tile != 0 && tc4_eligible(p->plan)) {
    // narrow graph (every stage N <= 128): four tiles in flight per CTA
    const int grid4 = (int)std::min<int64_t>(ceil_div(n_tiles, 4), sm_count());
    mlp_tc4_kernel<<<grid4, kThreadsTc4, tc4_smem_bytes(p->plan), stream>>>(p->plan, p->d_w, p->d_bias, src, (long long)n_points, raw);
    NB_LAUNCHED();
    return NB_OK;
  }
  if (rec == nullptr && fwd_2cta_mode() != 0 && tc2_smem_bytes(p->plan) <= 227 * 1024) {
    // CTA-pair kernel: clusters of 2 CTAs, four tiles per cluster round
    const int clusters = (int)std::min<int64_t>(ceil_div(n_tiles, 4), sm_count() / 2);
    mlp_tc2_kernel<<<2 * clusters, kThreadsTc, tc2_smem_bytes(p->plan), stream>>>(p->plan, p->d_w, p->d_bias, src,
                                                                                 (long long)n_points, raw);
    NB_LAUNCHED();
    return NB_OK;
  }
  const int grid = (int)std::min<int64_t>(ceil_div(n_tiles, 2), sm_count());
  // Timing experiments behind the numbers in DESIGN.md (results are then garbage) exist only in builds made with
  // -DNB_TIMING_EXPERIMENTS: NB_FWD_DEBUG=1 producer skips the weight loads, 2 = epilogue skips the conversion
#ifdef NB_TIMING_EXPERIMENTS
  static const int dbg = getenv("NB_FWD_DEBUG") ? atoi(getenv("NB_FWD_DEBUG")) : 0;
#else
  const int dbg = 0;
#endif
  mlp_tc_kernel<<<grid, kThreadsTc, p->smem_bytes, stream>>>(p->plan, p->d_w, p->d_bias, src, (long long)n_points, raw, rec, mrec, dbg);
  NB_LAUNCHED();
  return NB_OK;
}

}  // namespace nb

namespace nb { extern int g_bwd_sm_reserve; }
extern "C" float nb_mlp_activation_scale(void) { return nb::kActScale; }

extern "C" int nb_set_option(const char* name, int value) {
  using namespace nb;
  NB_REQUIRE(name != nullptr, "nb_set_option: null name");
  if (strcmp(name, "fwd_2cta") == 0) return set_fwd_2cta(value);
  if (strcmp(name, "fwd_4tile") == 0) {
    const int old = g_fwd_4tile;
    g_fwd_4tile = value != 0;
    return old;
  }
  if (strcmp(name, "composite_lc") == 0) return set_composite_lc(value);
  if (strcmp(name, "sample_fine2") == 0) return set_sample_fine2(value);
  if (strcmp(name, "bwd_sm_reserve") == 0) {
    const int old = g_bwd_sm_reserve;
    g_bwd_sm_reserve = value < 0 ? 0 : (value > 64 ? 64 : value);
    return old;
  }
  set_error("nb_set_option: unknown option '%s'", name);
  return NB_ERR_INVALID;
}
