// Backward of the fused tensor-core MLP (train step, SURVEY.md 8a row a13: tf.GradientTape / loss.backward() through
// run_network + the network, nerf/run_nerf.py:807-826, nerf-pytorch/train_nerf.py:286) for sm_100a.
//
// The forward kernel (mlp_tc.cu) leaves, per 128-point tile, an "X record": every operand tile of the network
// (position / direction encodings and the 16-bit activations of every layer) as 16 KB blocks in the exact
// SWIZZLE_128B shared-memory image.  The same bytes serve three purposes without any re-layout:
//   * K-major A operand of the forward MMAs (what they were written for),
//   * ReLU masks of the dgrad chain (x > 0 element-wise, same swizzled position as the gradient element),
//   * MN-major operand of the wgrad MMAs (contraction over the 128 points = the ROWS of a block).
//
// Two kernels:
//   1. mlp_tc_dgrad_kernel -- the reversed layer chain for a tile, fused exactly like the forward: dZ tiles stay in
//      shared memory / TMEM between layers; A = dZ (K-major), B = transposed weight chunks streamed through a ring
//      by bulk async copies, D in TMEM; epilogue = tcgen05.ld -> ReLU mask -> fp16 -> next A operand.  Each dZ tile is
//      also bulk-stored to the per-tile "dZ record" for the wgrad pass.
//   2. mlp_tc_wgrad_kernel -- dW[out,in] = sum_points dZ[p,out] * X[p,in]: both operands MN-major straight from the two
//      records, accumulators resident in TMEM (up to 256x256 fp32 = all 512 columns) while a CTA sweeps its slice of
//      the points (split-K over CTAs), bias gradients as column sums by otherwise idle warps; partial results are
//      reduced deterministically by wgrad_reduce_kernel into the caller's flat gradient.
//      The N_out <= 4 heads (rgb / alpha / output_linear) are wgrad units too: their dZ is the scaled 16-bit d_raw tile.
//
// Gradients are computed in fp16 with a per-call power-of-two scale chosen from max|d_raw| (absmax_kernel), so the
// 5-bit exponent is centred on the data; the scale is removed in fp32 by the reduce kernel.
#include <stdlib.h>

#include <algorithm>
#include <vector>

#include "tc_common.cuh"

namespace nb {

constexpr int kThreadsBw = 224;   // W producer, MMA issuer, 4 epilogue warps, mask producer
constexpr int kMaskRing = 0;     // (the ReLU masks used to be 16 KB tiles of the X record streamed through a ring)
constexpr int kThreadsWg = 192;   // producer, MMA issuer, 4 column-sum / epilogue warps
constexpr int kWgMaxStages = 8;
constexpr int kHalfBytes = 64 * 128;  // 64 points of a block

// scale = 2^floor(log2(1024 / max|d_raw|)); 1 when the maximum is 0 or not finite
__device__ __forceinline__ float grad_scale_from(unsigned bits) {
  const float mx = __uint_as_float(bits);
  if (!(mx > 0.0f) || !(mx < 3.0e38f)) return 1.0f;
  int e;
  frexpf(1024.0f / mx, &e);  // 1024/mx = f * 2^e, f in [0.5, 1)
  e = max(-100, min(100, e - 1));
  return ldexpf(1.0f, e);
}

__global__ void absmax_kernel(const float* __restrict__ x, long long n, unsigned* __restrict__ out) {
  float m = 0.0f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float v = fabsf(x[i]);
    if (v < 3.0e38f) m = fmaxf(m, v);  // inf / nan do not steer the scale
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0 && m > 0.0f) atomicMax(out, __float_as_uint(m));
}

// ------------------------------------------------------------------------------------ transposed weight packing
__global__ void pack_weights_t_kernel(const float* __restrict__ params, const PackSegT* __restrict__ segs,
                                      unsigned char* __restrict__ out) {
  pack_weights_t_body(params, segs[blockIdx.y], out, blockIdx.x * blockDim.x + threadIdx.x);
}

// SMs the persistent backward kernels leave free (nb_set_option("bwd_sm_reserve")): in the data-parallel step the fine
// network's gradient all-reduce is enqueued on NCCL's stream while the coarse network's backward runs; both backward
// kernels hold one 215 KB CTA on every SM, so NCCL's CTAs could not become resident until a backward kernel retired.
int g_bwd_sm_reserve = 0;
static int bwd_sms() { return std::max(8, sm_count() - g_bwd_sm_reserve); }

// ------------------------------------------------------------------------------------ dgrad chain
// flags of the dgrad MMA issuer's step program (bits 0-1: gradient block to wait for)
enum : uint32_t { BF_WAIT_ACT = 4u, BF_K1 = 8u, BF_ACC = 16u, BF_LAST = 32u, BF_BUF1 = 64u };
// constant upper half of a SWIZZLE_128B K-major operand descriptor (see umma_desc)
constexpr uint64_t kDescHi128 = ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)2 << 61);

struct BwArgs {
  const unsigned char* wt;
  const float* d_raw;             // [n_points, 4]
  const unsigned* absmax_bits;
  const unsigned char* xrec;
  const unsigned char* mrec;      // 1-bit ReLU masks written by the forward epilogue: [tile][rec_blocks - 2][128 rows] x 8 B
  unsigned char* drec;
  long long n_points;
  int rec_blocks, n_g_blocks;
};

__global__ void __launch_bounds__(kThreadsBw, 1)
mlp_tc_dgrad_kernel(const __grid_constant__ BwPlan plan, const BwArgs a) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  unsigned char* gact = smem;                                   // dZ tiles (A operands)
  unsigned char* draw = gact + a.n_g_blocks * kBlockBytes;      // scaled d_raw tile (K = 16 used)
  unsigned char* mring = draw + kBlockBytes;                    // ReLU-mask tiles from the X record
  unsigned char* ring = mring + kMaskRing * kBlockBytes;        // transposed weight chunks
  uint64_t* bars = reinterpret_cast<uint64_t*>(ring + kRing * plan.ring_stage_bytes);
  uint64_t* w_full = bars;
  uint64_t* w_empty = w_full + kRing;
  uint64_t* m_full = w_empty + kRing;
  uint64_t* m_empty = m_full + kMaskRing;
  uint64_t* bar_draw = m_empty + kMaskRing;
  uint64_t* bar_act = bar_draw + 1;
  uint64_t* bar_acc = bar_act + 4;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_acc + 2);
  uint4* prog = reinterpret_cast<uint4*>(reinterpret_cast<unsigned char*>(tmem_slot) + 16 + ((16 - (smem_u32(tmem_slot) & 15u)) & 15u));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long n_tiles = (a.n_points + kTileM - 1) / kTileM;

  if (warp == 6 && lane < plan.n_stages) {
    // MMA issuer step program, one entry per K-block (lane s builds stage s): x = byte offset of the A operand block from
    // the gradient tiles' base (the d_raw tile follows them), y = instruction descriptor, z = flags, w = dZ-record block the
    // awaited gradient block is saved to
    const int s = lane;
    const BwStage& st = plan.st[s];
    int i = st.step0;
    for (int kb = 0; kb < st.n_kb; ++kb) {
      const int sidx = st.a_src[kb];
      uint4 e;
      e.x = (uint32_t)((sidx < 4 ? sidx : a.n_g_blocks) * kBlockBytes);
      e.y = umma_idesc(st.n_main);
      e.z = (s & 1) ? BF_BUF1 : 0u;
      e.w = 0u;
      if (sidx < 4) { e.z |= BF_WAIT_ACT | (uint32_t)sidx; e.w = (uint32_t)(plan.st[s - 1].drec_blk + sidx); }
      if (st.k_steps[kb] == 1) e.z |= BF_K1;
      if (kb != 0) e.z |= BF_ACC;
      if (kb == st.n_kb - 1) e.z |= BF_LAST;
      prog[i++] = e;
    }
  }
  if (threadIdx.x == 0) {
    for (int i = 0; i < kRing; ++i) {
      mbar_init(smem_u32(w_full + i), 1);
      mbar_init(smem_u32(w_empty + i), 1);
    }
    for (int i = 0; i < kMaskRing; ++i) {
      mbar_init(smem_u32(m_full + i), 1);
      mbar_init(smem_u32(m_empty + i), 128);
    }
    mbar_init(smem_u32(bar_draw), 128);
    for (int i = 0; i < 4; ++i) mbar_init(smem_u32(bar_act + i), 128);
    mbar_init(smem_u32(bar_acc), 1);
    mbar_init(smem_u32(bar_acc + 1), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  // columns 8..63 of the d_raw tile are never written again: zero them once (they are rows of an MN-major wgrad operand)
  for (int i = threadIdx.x; i < kBlockBytes / 16; i += blockDim.x) reinterpret_cast<uint4*>(draw)[i] = make_uint4(0u, 0u, 0u, 0u);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ============================ transposed-weight producer ============================
    if (lane == 0) {
      uint32_t slot = 0, phase = 0;
      for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        for (int s = 0; s < plan.n_stages; ++s) {
          const BwStage& st = plan.st[s];
          for (int kb = 0; kb < st.n_kb; ++kb) {
            mbar_wait(smem_u32(w_empty + slot), phase ^ 1);
            const uint32_t fb = smem_u32(w_full + slot);
            mbar_expect_tx(fb, (uint32_t)st.chunk_bytes);
            bulk_g2s(smem_u32(ring + slot * plan.ring_stage_bytes), a.wt + st.w_off + (size_t)kb * st.chunk_bytes,
                     (uint32_t)st.chunk_bytes, fb);
            if (++slot == kRing) { slot = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 6) {
    // (this warp built the step program above; the ReLU masks are now 8 bytes per row read straight from the 1-bit
    // record by the epilogue threads, so there is no mask tile to stream)
  } else if (warp == 1) {
    // ============================ MMA issuer ============================
    // Whole warp, warp-uniform control flow over a step program in shared memory (one entry per K-block); only the
    // tcgen05 / bulk-copy instructions are predicated on one elected lane (see mlp_tc.cu: the issue loop must stay
    // well below the tensor time of a step or the pipe starves).
    uint32_t slot = 0, phase = 0, draw_phase = 0, ph_act = 0;
    const uint32_t gact_s = smem_u32(gact), ring_s = smem_u32(ring), draw_s = smem_u32(draw);
    const uint32_t w_full_s = smem_u32(w_full), w_empty_s = smem_u32(w_empty), bar_act_s = smem_u32(bar_act);
    const uint32_t bar_acc_s = smem_u32(bar_acc);
    const int n_steps = plan.n_steps;
    const BwStage& last = plan.st[plan.n_stages - 1];
    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
      unsigned char* dt = a.drec + (size_t)tile * plan.drec_blocks * kBlockBytes;
      mbar_wait(smem_u32(bar_draw), draw_phase);
      draw_phase ^= 1;
      tc_fence_after();
      if (elect_one()) {
        bulk_s2g(dt + (size_t)plan.draw_blk * kBlockBytes, draw_s, kBlockBytes);
        bulk_commit();
      }
      __syncwarp();
      uint4 e = prog[0];
      for (int i = 0; i < n_steps; ++i) {
        const uint4 en = prog[i + 1 < n_steps ? i + 1 : i];  // prefetch
        const uint32_t fl = e.z;
        if (fl & BF_WAIT_ACT) {
          const uint32_t j = fl & 3u;
          mbar_wait(bar_act_s + 8 * j, (ph_act >> j) & 1u);
          ph_act ^= 1u << j;
          if (elect_one()) {
            bulk_s2g(dt + (size_t)e.w * kBlockBytes, gact_s + e.x, kBlockBytes);
            bulk_commit();
          }
          __syncwarp();
        }
        mbar_wait(w_full_s + 8 * slot, phase);
        tc_fence_after();
        const uint64_t a_desc = kDescHi128 | (uint64_t)(((gact_s + e.x) & 0x3FFFFu) >> 4);
        const uint64_t b_desc = kDescHi128 | (uint64_t)(((ring_s + slot * plan.ring_stage_bytes) & 0x3FFFFu) >> 4);
        const uint32_t d_main = tmem_base + ((fl & BF_BUF1) ? 256u : 0u);
        if (elect_one()) {
          tc_mma(d_main, a_desc, b_desc, e.y, (fl & BF_ACC) ? 1u : 0u);
          if (!(fl & BF_K1)) {
            tc_mma(d_main, a_desc + 2, b_desc + 2, e.y, 1u);
            tc_mma(d_main, a_desc + 4, b_desc + 4, e.y, 1u);
            tc_mma(d_main, a_desc + 6, b_desc + 6, e.y, 1u);
          }
          tc_commit(w_empty_s + 8 * slot);
          if (fl & BF_LAST) {
            bulk_wait_read0();  // record stores have read the tiles this stage's epilogue overwrites
            tc_commit(bar_acc_s + ((fl & BF_BUF1) ? 8u : 0u));
          }
        }
        __syncwarp();
        if (++slot == kRing) { slot = 0; phase ^= 1; }
        e = en;
      }
      // the last stage's gradient tiles feed no MMA; they only go to the record
      for (int j = 0; j < last.n_out; ++j) {
        mbar_wait(bar_act_s + 8 * j, (ph_act >> j) & 1u);
        ph_act ^= 1u << j;
        if (elect_one()) {
          bulk_s2g(dt + (size_t)(last.drec_blk + j) * kBlockBytes, gact_s + j * kBlockBytes, kBlockBytes);
          bulk_commit();
        }
        __syncwarp();
      }
    }
    if (elect_one()) bulk_wait0();
    __syncwarp();
  } else {
    // ============================ epilogue warps (thread == row) ============================
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const uint32_t lane_addr = ((uint32_t)(q * 32)) << 16;
    const float scale = grad_scale_from(*a.absmax_bits);
    uint32_t acc_phase[2] = {0, 0};
    const int sw = row & 7;
    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
      {
        const long long p = tile * kTileM + row;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (p < a.n_points) v = reinterpret_cast<const float4*>(a.d_raw)[p];
        uint4 pk;
        pk.x = pack_h2(v.x * scale, v.y * scale, false);
        pk.y = pack_h2(v.z * scale, v.w * scale, false);
        pk.z = 0u; pk.w = 0u;
        unsigned char* rp = draw + row * 128;
        *reinterpret_cast<uint4*>(rp + ((0 ^ sw) << 4)) = pk;
        *reinterpret_cast<uint4*>(rp + ((1 ^ sw) << 4)) = make_uint4(0u, 0u, 0u, 0u);
      }
      tc_fence_before();
      fence_async_smem();
      mbar_arrive(smem_u32(bar_draw));

      for (int s = 0; s < plan.n_stages; ++s) {
        const BwStage& st = plan.st[s];
        const int buf = s & 1;
        // this stage's ReLU masks (8 B per row and 64-column block) are requested before the wait for the stage's
        // accumulator: the MMAs in flight hide the HBM latency
        const bool masked = st.mask_blk >= 0;
        const uint2* mrow = reinterpret_cast<const uint2*>(a.mrec) +
                            ((size_t)tile * (a.rec_blocks - 2) + (masked ? st.mask_blk - 2 : 0)) * kTileM + row;
        uint2 mk0, mk1, mk2, mk3;
        mk0 = mk1 = mk2 = mk3 = make_uint2(0xffffffffu, 0xffffffffu);
        if (masked) {
          mk0 = __ldg(mrow);
          if (st.n_out > 1) mk1 = __ldg(mrow + kTileM);
          if (st.n_out > 2) mk2 = __ldg(mrow + 2 * kTileM);
          if (st.n_out > 3) mk3 = __ldg(mrow + 3 * kTileM);
        }
        mbar_wait(smem_u32(bar_acc + buf), acc_phase[buf]);
        acc_phase[buf] ^= 1;
        tc_fence_after();
        const uint32_t t_main = tmem_base + lane_addr + buf * 256;
        for (int j = 0; j < st.n_out; ++j) {
          uint32_t v[64];
          const uint2 mk = j == 0 ? mk0 : j == 1 ? mk1 : j == 2 ? mk2 : mk3;
          NB_TMEM_LD32(t_main + j * 64, v);
          {
            uint32_t* v2 = v + 32;
            NB_TMEM_LD32(t_main + j * 64 + 32, v2);
          }
          NB_TMEM_WAIT64(v);
          unsigned char* rp = gact + j * kBlockBytes + row * 128;
          if (masked) {
            // element i of the block passes where bit 31 - (i & 31) of mask word i >> 5 is set (forward epilogue)
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              if (!(mk.x & (0x80000000u >> i))) v[i] = 0u;
              if (!(mk.y & (0x80000000u >> i))) v[32 + i] = 0u;
            }
          }
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            uint4 pk;
            pk.x = pack_h2(__uint_as_float(v[8 * c + 0]), __uint_as_float(v[8 * c + 1]), false);
            pk.y = pack_h2(__uint_as_float(v[8 * c + 2]), __uint_as_float(v[8 * c + 3]), false);
            pk.z = pack_h2(__uint_as_float(v[8 * c + 4]), __uint_as_float(v[8 * c + 5]), false);
            pk.w = pack_h2(__uint_as_float(v[8 * c + 6]), __uint_as_float(v[8 * c + 7]), false);
            *reinterpret_cast<uint4*>(rp + ((c ^ sw) << 4)) = pk;
          }
          tc_fence_before();
          fence_async_smem();
          mbar_arrive(smem_u32(bar_act + j));
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

// ------------------------------------------------------------------------------------ wgrad (split-K over CTAs)
// pipeline depth of a unit whose stage holds n_slots half-blocks: as many as fit next to the barriers
__host__ __device__ inline int wg_stages(int n_slots) {
  const int s = (216 * 1024) / (n_slots * kHalfBytes);
  return s > kWgMaxStages ? kWgMaxStages : s;
}

struct WgArgs {
  const unsigned char* xrec;
  const unsigned char* drec;
  float* part_w;   // [grid][65536]
  float* part_b;   // [grid][4 row groups][256]
  long long n_tiles;
  int rec_blocks, drec_blocks;
  int part0[kMaxUnits + 1];  // CTA b works on unit u with part0[u] <= b < part0[u+1]
  int dbg;                   // NB_WG_DEBUG bit 0: no loads, bit 1: no MMAs, bit 2: no column sums (timing experiments)
};
constexpr int kPartFloats = 65536;

__global__ void __launch_bounds__(kThreadsWg, 1)
mlp_tc_wgrad_kernel(const __grid_constant__ WgPlan plan, const __grid_constant__ WgArgs a) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  int ui = 0;
  while (ui + 1 < plan.n_units && (int)blockIdx.x >= a.part0[ui + 1]) ++ui;
  const WgUnit& u = plan.u[ui];
  const int part = blockIdx.x - a.part0[ui], nparts = a.part0[ui + 1] - a.part0[ui];
  const long long tile0 = a.n_tiles * part / nparts, tile1 = a.n_tiles * (part + 1) / nparts;
  const int n_slots = u.m_blocks + u.n_x;
  const int stage_bytes = n_slots * kHalfBytes;
  const int kWgStages = wg_stages(n_slots);
  const int Ntot = 64 * u.n_x;
  unsigned char* stages = smem;
  uint64_t* bars = reinterpret_cast<uint64_t*>(stages + kWgStages * stage_bytes);
  uint64_t* bar_full = bars;
  uint64_t* bar_empty = bars + kWgMaxStages;
  uint64_t* bar_done = bars + 2 * kWgMaxStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_done + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int i = 0; i < kWgStages; ++i) {
      mbar_init(smem_u32(bar_full + i), 1);
      mbar_init(smem_u32(bar_empty + i), 1 + 128);
    }
    mbar_init(smem_u32(bar_done), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  // dZ blocks the record does not hold (layer narrower than the 128-row MMA) are zeros
  if (u.m_load < u.m_blocks) {
    for (int s = 0; s < kWgStages; ++s) {
      uint4* z = reinterpret_cast<uint4*>(stages + s * stage_bytes + u.m_load * kHalfBytes);
      const int n16 = (u.m_blocks - u.m_load) * kHalfBytes / 16;
      for (int i = threadIdx.x; i < n16; i += blockDim.x) z[i] = make_uint4(0u, 0u, 0u, 0u);
    }
    fence_async_smem();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const long long n_steps = (tile1 - tile0) * 2;  // half tiles

  if (warp == 0) {
    if (lane == 0) {
      uint32_t slot = 0, phase = 0;
      for (long long it = 0; it < n_steps; ++it) {
        const long long tile = tile0 + (it >> 1);
        const size_t hoff = (size_t)(it & 1) * kHalfBytes;
        const unsigned char* dzt = a.drec + ((size_t)tile * a.drec_blocks + u.dz_blk) * kBlockBytes + hoff;
        const unsigned char* xt = a.xrec + (size_t)tile * a.rec_blocks * kBlockBytes + hoff;
        mbar_wait(smem_u32(bar_empty + slot), phase ^ 1);
        const uint32_t fb = smem_u32(bar_full + slot);
        if (NB_DBG(a.dbg & 1)) { mbar_arrive(fb); if (++slot == kWgStages) { slot = 0; phase ^= 1; } continue; }
        mbar_expect_tx(fb, (uint32_t)((u.m_load + u.n_x) * kHalfBytes));
        const uint32_t sbase = smem_u32(stages + slot * stage_bytes);
        for (int b = 0; b < u.m_load; ++b) bulk_g2s(sbase + b * kHalfBytes, dzt + (size_t)b * kBlockBytes, kHalfBytes, fb);
        for (int x = 0; x < u.n_x; ++x)
          bulk_g2s(sbase + (u.m_blocks + x) * kHalfBytes, xt + (size_t)u.x_blk[x] * kBlockBytes, kHalfBytes, fb);
        if (++slot == kWgStages) { slot = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      uint32_t slot = 0, phase = 0;
      const int n0 = Ntot > 256 ? 256 : Ntot, n1 = Ntot - n0;
      const uint32_t id0 = umma_idesc_mn(n0), id1 = umma_idesc_mn(n1 > 0 ? n1 : 16);
      for (long long it = 0; it < n_steps; ++it) {
        mbar_wait(smem_u32(bar_full + slot), phase);
        tc_fence_after();
        const uint32_t sbase = smem_u32(stages + slot * stage_bytes);
        const uint32_t xb = sbase + u.m_blocks * kHalfBytes;
        for (int mh = 0; mh < (NB_DBG(a.dbg & 2) ? 0 : u.m_blocks / 2); ++mh) {
          const uint32_t ab = sbase + 2 * mh * kHalfBytes;
          const uint32_t d = tmem_base + mh * Ntot;
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const uint32_t acc = (it | k) != 0 ? 1u : 0u;
            const uint64_t ad = umma_desc_mn(ab + k * 2048, kHalfBytes, 1024);
            tc_mma(d, ad, umma_desc_mn(xb + k * 2048, kHalfBytes, 1024), id0, acc);
            if (n1 > 0) tc_mma(d + 256, ad, umma_desc_mn(xb + 4 * kHalfBytes + k * 2048, kHalfBytes, 1024), id1, acc);
          }
        }
        tc_commit(smem_u32(bar_empty + slot));
        if (++slot == kWgStages) { slot = 0; phase ^= 1; }
      }
      tc_commit(smem_u32(bar_done));
    }
  } else {
    // ---- bias gradient: column sums of the dZ tiles while they sit in shared memory.  Warp w sums rows
    // [16w, 16w+16) of every stage; lane l owns the 16-byte chunk l of the 256-column row (block l>>3, chunk l&7).
    const int t = threadIdx.x - 64;
    const int rg = t >> 5, blk = lane >> 3, chunk = lane & 7;
    const bool active = u.b_off >= 0 && blk < u.m_load && !NB_DBG(a.dbg & 4);
    float cs[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) cs[i] = 0.0f;
    uint32_t slot = 0, phase = 0;
    for (long long it = 0; it < n_steps; ++it) {
      mbar_wait(smem_u32(bar_full + slot), phase);
      if (active) {
        const uint32_t base = smem_u32(stages + slot * stage_bytes + blk * kHalfBytes + rg * 16 * 128);
#pragma unroll
        for (int r = 0; r < 16; ++r) {
          uint4 v;
          asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
                       : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                       : "r"(base + r * 128 + ((chunk ^ (r & 7)) << 4)));
          const float2 f0 = __half22float2(*reinterpret_cast<const __half2*>(&v.x));
          const float2 f1 = __half22float2(*reinterpret_cast<const __half2*>(&v.y));
          const float2 f2 = __half22float2(*reinterpret_cast<const __half2*>(&v.z));
          const float2 f3 = __half22float2(*reinterpret_cast<const __half2*>(&v.w));
          cs[0] += f0.x; cs[1] += f0.y; cs[2] += f1.x; cs[3] += f1.y;
          cs[4] += f2.x; cs[5] += f2.y; cs[6] += f3.x; cs[7] += f3.y;
        }
      }
      mbar_arrive(smem_u32(bar_empty + slot));
      if (++slot == kWgStages) { slot = 0; phase ^= 1; }
    }
    {
      float4* pb = reinterpret_cast<float4*>(a.part_b + ((size_t)blockIdx.x * 4 + rg) * 256 + lane * 8);
      pb[0] = make_float4(cs[0], cs[1], cs[2], cs[3]);
      pb[1] = make_float4(cs[4], cs[5], cs[6], cs[7]);
    }
    // ---- accumulators -> partial buffer
    mbar_wait(smem_u32(bar_done), 0);
    tc_fence_after();
    const int q = warp & 3;
    const uint32_t lane_addr = ((uint32_t)(q * 32)) << 16;
    float* pw = a.part_w + (size_t)blockIdx.x * kPartFloats;
    for (int mh = 0; mh < u.m_blocks / 2; ++mh) {
      const int row = mh * 128 + q * 32 + lane;
      float4* dst = reinterpret_cast<float4*>(pw + (size_t)row * Ntot);
      for (int c0 = 0; c0 < Ntot; c0 += 32) {
        uint32_t v[32];
        NB_TMEM_LD32(tmem_base + lane_addr + mh * Ntot + c0, v);
        tc_wait_ld();
        if (n_steps > 0) {
#pragma unroll
          for (int i = 0; i < 8; ++i)
            dst[(c0 >> 2) + i] = make_float4(__uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1]),
                                             __uint_as_float(v[4 * i + 2]), __uint_as_float(v[4 * i + 3]));
        } else {
#pragma unroll
          for (int i = 0; i < 8; ++i) dst[(c0 >> 2) + i] = make_float4(0.f, 0.f, 0.f, 0.f);
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

// grad[w] += (sum over the unit's CTAs of the partial) / scale ; same for the bias.  Deterministic order (partial p0 first).
// Four consecutive accumulator elements per thread (16-byte loads of the partials: same row, same X block).
__global__ void wgrad_reduce_kernel(const __grid_constant__ WgPlan plan, const __grid_constant__ WgArgs a,
                                    const unsigned* __restrict__ absmax_bits, float* __restrict__ grad) {
  const WgUnit& u = plan.u[blockIdx.y];
  const int Ntot = 64 * u.n_x, M = 64 * u.m_blocks;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int n4 = M * Ntot / 4;
  const int p0 = a.part0[blockIdx.y], p1 = a.part0[blockIdx.y + 1];
  const float inv = 1.0f / grad_scale_from(*absmax_bits);
  if (t < n4) {
    const int e = 4 * t;
    const int row = e / Ntot, c = e - row * Ntot;
    const int xs = c >> 6, cc = c & 63;
    const int wr = row - u.row0;
    if (wr < 0 || wr >= u.out_rows || cc >= u.x_valid[xs]) return;
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    const float4* src = reinterpret_cast<const float4*>(a.part_w) + t;
#pragma unroll 4
    for (int p = p0; p < p1; ++p) {
      const float4 v = src[(size_t)p * (kPartFloats / 4)];
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
    // X-record blocks 0 / 1 are the encodings (true values); every other block holds kActScale * activation
    const float xinv = u.x_blk[xs] >= 2 ? kInvActScale : 1.0f;
    float* g = grad + u.w_off + (long long)wr * u.ld + u.x_col0[xs] + cc;
    const int nv = u.x_valid[xs] - cc;   // >= 1
    g[0] += s.x * inv * xinv;
    if (nv > 1) g[1] += s.y * inv * xinv;
    if (nv > 2) g[2] += s.z * inv * xinv;
    if (nv > 3) g[3] += s.w * inv * xinv;
  } else if (t < n4 + 256) {
    const int j = t - n4, wr = j - u.row0;
    if (u.b_off < 0 || wr < 0 || wr >= u.out_rows) return;
    float s = 0.0f;
    for (int p = 4 * p0; p < 4 * p1; ++p) s += a.part_b[(size_t)p * 256 + j];
    grad[u.b_off + wr] += s * inv;
  }
}

// ------------------------------------------------------------------------------------ host: plans
static bool relu_after_trunk(const Graph& g, int i) { return !(i == 0 && !g.act0); }

// Builds the dgrad-chain plan, the wgrad units and the head segments from the graph and the forward plan
// (p->plan must already hold the X-record block of every forward stage).
bool bw_build_plan(const Graph& g, PackedBf16* p, std::vector<PackSegT>& segs, const char** why) {
  const TcPlan& fp = p->plan;
  BwPlan& bw = p->bw;
  WgPlan& wg = p->wg;
  memset(&bw, 0, sizeof(bw));
  memset(&wg, 0, sizeof(wg));
  *why = nullptr;
  const int D = g.depth, wb = g.width / 64;
  size_t woff = 0;
  int ring_max = 0;
  auto h_blk = [&](int i) { return fp.st[i].rec_blk; };  // activations of trunk layer i
  auto begin = [&](int n_main, int mask_blk) -> BwStage& {
    BwStage& st = bw.st[bw.n_stages++];
    st.n_kb = 0;
    st.n_main = n_main;
    st.chunk_bytes = n_main * 128;
    st.w_off = (int)woff;
    st.mask_blk = mask_blk;
    st.n_out = n_main / 64;
    st.drec_blk = bw.drec_blocks;
    bw.drec_blocks += st.n_out;
    ring_max = std::max(ring_max, st.chunk_bytes);
    return st;
  };
  auto add_kb = [&](BwStage& st, int a_src, int k_steps, const Lin& l, int col0, int krow0, int kcount, int kdst0) {
    PackSegT sg{};
    sg.dst_off = (long long)woff;
    sg.src_off = l.w_off;
    sg.n_rows = st.n_main;
    sg.ld = l.in;
    sg.col0 = col0;
    sg.krow0 = krow0;
    sg.kcount = kcount;
    sg.kdst0 = kdst0;
    segs.push_back(sg);
    st.a_src[st.n_kb] = a_src;
    st.k_steps[st.n_kb] = k_steps;
    st.n_kb++;
    woff += st.chunk_bytes;
  };
  auto add_unit = [&](int dz_blk, int dz_cols, std::initializer_list<int> xblks, std::initializer_list<int> xcol0,
                      std::initializer_list<int> xvalid, const Lin& l, bool bias) -> bool {
    if (wg.n_units >= kMaxUnits) return false;
    WgUnit& u = wg.u[wg.n_units++];
    u.dz_blk = dz_blk;
    u.m_load = dz_cols / 64;
    u.m_blocks = u.m_load <= 2 ? 2 : 4;
    u.n_x = (int)xblks.size();
    int i = 0;
    for (int v : xblks) u.x_blk[i++] = v;
    i = 0;
    for (int v : xcol0) u.x_col0[i++] = v;
    i = 0;
    for (int v : xvalid) u.x_valid[i++] = v;
    u.row0 = 0;
    u.out_rows = l.out;
    u.w_off = l.w_off;
    u.ld = l.in;
    u.b_off = bias ? l.b_off : -1;
    u.cost = u.m_load + u.n_x;  // the kernel is HBM-bound: cost = half-blocks loaded per stage
    return (u.m_blocks / 2) * 64 * u.n_x <= 512;
  };
  // head layer (N_out <= 4): dZ = channels [ch0, ch0 + l.out) of the d_raw tile, X = n_xb blocks from x_blk0
  auto add_head = [&](const Lin& l, int ch0, int x_blk0, int n_xb) -> bool {
    if (wg.n_units >= kMaxUnits || n_xb > kMaxXSrc) return false;
    WgUnit& u = wg.u[wg.n_units++];
    u.dz_blk = bw.draw_blk;
    u.m_load = 1;
    u.m_blocks = 2;
    u.n_x = n_xb;
    for (int j = 0; j < n_xb; ++j) { u.x_blk[j] = x_blk0 + j; u.x_col0[j] = 64 * j; u.x_valid[j] = 64; }
    u.row0 = ch0;
    u.out_rows = l.out;
    u.w_off = l.w_off;
    u.ld = l.in;
    u.b_off = l.b_off;
    u.cost = u.m_load + u.n_x;
    return true;
  };
  if (D + 3 > kMaxStages) { *why = "too many layers"; return false; }
  if (wb != 2 && wb != 4) { *why = "width must be 128 or 256"; return false; }
  std::vector<int> dz_trunk(D, -1);
  int dz_dir = -1, dz_feat = -1;
  const int maskL = relu_after_trunk(g, D - 1) ? h_blk(D - 1) : -1;
  if (g.use_viewdirs) {
    const Lin& la = g.lins[g.alpha];
    const Lin& lf = g.lins[g.feat];
    const Lin& ld = g.lins[g.dir];
    const Lin& lr = g.lins[g.rgb];
    const int feat_blk = fp.st[D].rec_blk, dirh_blk = fp.st[D + 1].rec_blk;
    const int fb = g.feat_width / 64, db = g.dir_width / 64;
    // d(dir layer pre-activation) = d_rgb . W_rgb, masked by relu(dir layer)
    BwStage& s0 = begin(g.dir_width, dirh_blk);
    add_kb(s0, SRC_DRAW, 1, lr, 0, 0, 3, 0);
    dz_dir = s0.drec_blk;
    // d(feat) = d_dir . W_dir[:, :feat]
    BwStage& s1 = begin(g.feat_width, g.feat_act ? feat_blk : -1);
    for (int j = 0; j < db; ++j) add_kb(s1, j, 4, ld, 0, 64 * j, 64, 0);
    dz_feat = s1.drec_blk;
    // d(h_last) = d_feat . W_feat + d_alpha . W_alpha, masked by relu(h_last)
    BwStage& s2 = begin(g.width, maskL);
    for (int j = 0; j < fb; ++j) add_kb(s2, j, 4, lf, 0, 64 * j, 64, 0);
    add_kb(s2, SRC_DRAW, 1, la, 0, 0, 1, 3);
    dz_trunk[D - 1] = s2.drec_blk;
  } else {
    const Lin& lo = g.lins[g.out4];
    BwStage& s0 = begin(g.width, maskL);
    add_kb(s0, SRC_DRAW, 1, lo, 0, 0, 4, 0);
    dz_trunk[D - 1] = s0.drec_blk;
  }
  for (int i = D - 1; i >= 1; --i) {
    const Lin& l = g.lins[g.trunk[i]];
    const int off = (i - 1 == g.skip) ? g.Cx : 0;
    BwStage& st = begin(g.width, relu_after_trunk(g, i - 1) ? h_blk(i - 1) : -1);
    for (int j = 0; j < wb; ++j) add_kb(st, j, 4, l, off, 64 * j, 64, 0);
    dz_trunk[i - 1] = st.drec_blk;
  }
  for (int s = 0; s < bw.n_stages; ++s)
    if (bw.st[s].n_kb > kMaxKB) { *why = "too many K blocks"; return false; }
  bw.ring_stage_bytes = (int)round_up(ring_max, 1024);
  bw.n_steps = 0;
  for (int s = 0; s < bw.n_stages; ++s) {
    bw.st[s].step0 = bw.n_steps;
    bw.n_steps += bw.st[s].n_kb;
  }
  bw.draw_blk = bw.drec_blocks++;
  p->wt_bytes = woff;

  // ---- wgrad units
  bool ok = true;
  if (!g.use_viewdirs) ok = ok && add_head(g.lins[g.out4], 0, h_blk(D - 1), wb);
  if (g.use_viewdirs) {
    const Lin& lf = g.lins[g.feat];
    const Lin& ld = g.lins[g.dir];
    const int feat_blk = fp.st[D].rec_blk;
    ok = ok && add_head(g.lins[g.rgb], 0, fp.st[D + 1].rec_blk, g.dir_width / 64);
    ok = ok && add_head(g.lins[g.alpha], 3, h_blk(D - 1), wb);
    // dir layer: X = [feat | dir encoding]
    {
      if (wg.n_units >= kMaxUnits) ok = false;
      WgUnit& u = wg.u[wg.n_units++];
      u.dz_blk = dz_dir;
      u.m_load = g.dir_width / 64;
      u.m_blocks = 2;
      u.n_x = 0;
      for (int j = 0; j < g.feat_width / 64; ++j) {
        u.x_blk[u.n_x] = feat_blk + j; u.x_col0[u.n_x] = 64 * j; u.x_valid[u.n_x] = 64; u.n_x++;
      }
      if (g.Cd > 0) { u.x_blk[u.n_x] = 1; u.x_col0[u.n_x] = g.feat_width; u.x_valid[u.n_x] = g.Cd; u.n_x++; }
      u.row0 = 0; u.out_rows = ld.out; u.w_off = ld.w_off; u.ld = ld.in; u.b_off = ld.b_off; u.cost = u.m_load + u.n_x;
      if (u.m_load > 2 || u.n_x > kMaxXSrc || 64 * u.n_x > 512) ok = false;
    }
    // feat / bottleneck: X = h_last
    if (wb == 4) ok = ok && add_unit(dz_feat, g.feat_width, {h_blk(D - 1), h_blk(D - 1) + 1, h_blk(D - 1) + 2, h_blk(D - 1) + 3},
                                     {0, 64, 128, 192}, {64, 64, 64, 64}, lf, true);
    else ok = ok && add_unit(dz_feat, g.feat_width, {h_blk(D - 1), h_blk(D - 1) + 1}, {0, 64}, {64, 64}, lf, true);
  }
  for (int i = D - 1; i >= 0; --i) {
    const Lin& l = g.lins[g.trunk[i]];
    if (i == 0) {
      ok = ok && add_unit(dz_trunk[0], g.width, {0}, {0}, {g.Cx}, l, true);
      break;
    }
    const bool cat = (i - 1 == g.skip);
    const int off = cat ? g.Cx : 0, hb = h_blk(i - 1);
    if (wb == 4) ok = ok && add_unit(dz_trunk[i], g.width, {hb, hb + 1, hb + 2, hb + 3}, {off, off + 64, off + 128, off + 192},
                                     {64, 64, 64, 64}, l, true);
    else ok = ok && add_unit(dz_trunk[i], g.width, {hb, hb + 1}, {off, off + 64}, {64, 64}, l, true);
    if (cat) ok = ok && add_unit(dz_trunk[i], g.width, {0}, {0}, {g.Cx}, l, false);
  }
  if (!ok) { *why = "wgrad units do not fit"; return false; }
  return true;
}

static size_t bw_smem_bytes(const BwPlan& bw, int n_g_blocks) {
  size_t b = 1024;
  b += (size_t)(n_g_blocks + 1 + kMaskRing) * kBlockBytes;
  b += (size_t)kRing * bw.ring_stage_bytes;
  b += (2 * kRing + 2 * kMaskRing + 1 + 4 + 2) * sizeof(uint64_t) + 16;
  b += 32 + (size_t)bw.n_steps * sizeof(uint4);
  return b;
}
static size_t wg_smem_bytes() { return 1024 + (size_t)216 * 1024 + (2 * kWgMaxStages + 1) * 8 + 16; }

// ------------------------------------------------------------------------------------ host: setup / workspace / launch
void bw_free(PackedBf16* p) {
  if (p == nullptr) return;
  cudaFree(p->d_wt);
  cudaFree(p->d_segs_t);
  p->d_wt = nullptr;
  p->d_segs_t = nullptr;
}

// Called once per model from packed_build (after the forward plan exists).
int bw_setup(nb_model* m, cudaStream_t stream) {
  PackedBf16* p = m->packed;
  std::vector<PackSegT> segs;
  const char* why = nullptr;
  p->bw_supported = bw_build_plan(m->g, p, segs, &why);
  if (!p->bw_supported) return NB_OK;  // inference still works; tc_backward reports the reason
  p->n_g_blocks = m->g.width / 64;
  if (m->g.use_viewdirs) p->n_g_blocks = std::max(p->n_g_blocks, m->g.feat_width / 64);
  p->bw_smem_bytes = bw_smem_bytes(p->bw, p->n_g_blocks);
  if (p->bw_smem_bytes > 227 * 1024 || p->n_g_blocks > 4) {
    p->bw_supported = false;
    return NB_OK;
  }
  p->n_segs_t = (int)segs.size();
  cudaError_t e = cudaMalloc(&p->d_wt, std::max<size_t>(p->wt_bytes, 16));
  if (e == cudaSuccess) e = cudaMalloc(&p->d_segs_t, sizeof(PackSegT) * segs.size());
  if (e == cudaSuccess) e = cudaMemcpyAsync(p->d_segs_t, segs.data(), sizeof(PackSegT) * segs.size(), cudaMemcpyHostToDevice, stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(mlp_tc_dgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(mlp_tc_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  if (e != cudaSuccess) {
    set_error("tensor-core backward setup failed: %s", cudaGetErrorString(e));
    return NB_ERR_CUDA;
  }
  p->wt_fresh = false;
  return NB_OK;
}

struct BwLayout {
  size_t off_scale = 0, off_xrec = 0, off_mrec = 0, off_drec = 0, off_pw = 0, off_pb = 0, total = 0;
  int64_t n_tiles = 0;
  int max_ctas = 0;
};
static BwLayout bw_layout(const PackedBf16* p, int64_t n_points) {
  BwLayout L;
  L.n_tiles = ceil_div(std::max<int64_t>(n_points, 1), kTileM);
  L.max_ctas = sm_count() + kMaxUnits;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off += (size_t)round_up((int64_t)bytes, 1024); return o; };
  L.off_scale = take(256);
  L.off_xrec = take((size_t)L.n_tiles * p->plan.rec_blocks * kBlockBytes);
  L.off_mrec = take((size_t)L.n_tiles * (p->plan.rec_blocks - 2) * kTileM * 8);
  L.off_drec = take((size_t)L.n_tiles * p->bw.drec_blocks * kBlockBytes);
  L.off_pw = take((size_t)L.max_ctas * kPartFloats * sizeof(float));
  L.off_pb = take((size_t)L.max_ctas * 4 * 256 * sizeof(float));
  L.total = off;
  return L;
}

unsigned char* tc_train_mrec(const PackedBf16* p, void* ws, int64_t n_points) {
  unsigned char* base = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(ws) + 1023) & ~(uintptr_t)1023);
  return base + bw_layout(p, n_points).off_mrec;
}
size_t tc_train_workspace_bytes(const PackedBf16* p, int64_t n_points) { return bw_layout(p, n_points).total + 1024; }
unsigned char* tc_train_xrec(const PackedBf16* p, void* ws, int64_t n_points) {
  unsigned char* base = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(ws) + 1023) & ~(uintptr_t)1023);
  return base + bw_layout(p, n_points).off_xrec;
}

// CTAs per unit proportional to MMA cost, at least 1, at most one per tile
static void wg_assign(const WgPlan& wg, int64_t n_tiles, int n_sm, int* part0) {
  int total = 0;
  for (int i = 0; i < wg.n_units; ++i) total += wg.u[i].cost;
  int used = 0;
  int parts[kMaxUnits];
  for (int i = 0; i < wg.n_units; ++i) {
    int64_t pz = std::max<int64_t>(1, (int64_t)n_sm * wg.u[i].cost / std::max(total, 1));
    pz = std::min<int64_t>(pz, n_tiles);
    parts[i] = (int)pz;
    used += parts[i];
  }
  // hand the remainder to the most expensive-per-CTA units
  while (used < n_sm) {
    int best = -1;
    double bestv = 0.0;
    for (int i = 0; i < wg.n_units; ++i) {
      if (parts[i] >= n_tiles) continue;
      const double v = (double)wg.u[i].cost / parts[i];
      if (v > bestv) { bestv = v; best = i; }
    }
    if (best < 0) break;
    parts[best]++;
    used++;
  }
  part0[0] = 0;
  for (int i = 0; i < wg.n_units; ++i) part0[i + 1] = part0[i] + parts[i];
}

int tc_backward(const nb_model* m, int64_t n_points, float* d_raw, void* ws, size_t ws_bytes, float* grad,
                cudaStream_t stream) {
  PackedBf16* p = m->packed;
  if (p == nullptr) {
    set_error("nb_mlp_backward: nb_model_pack(NB_PREC_BF16) has not been called");
    return NB_ERR_INVALID;
  }
  if (!p->bw_supported) {
    set_error("tensor-core backward does not support this graph (train it with NB_PREC_FP32)");
    return NB_ERR_UNSUPPORTED;
  }
  if (n_points <= 0) return NB_OK;
  const BwLayout L = bw_layout(p, n_points);
  if (ws == nullptr || ws_bytes < L.total + 1024) {
    set_error("nb_mlp_backward: workspace of %zu bytes given, %zu needed", ws_bytes, L.total + 1024);
    return NB_ERR_WORKSPACE;
  }
  unsigned char* base = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(ws) + 1023) & ~(uintptr_t)1023);
  unsigned* absmax = reinterpret_cast<unsigned*>(base + L.off_scale);
  if (!p->wt_fresh) {
    dim3 grid((unsigned)ceil_div(256 * 8, 256), (unsigned)p->n_segs_t);
    pack_weights_t_kernel<<<grid, 256, 0, stream>>>(m->params, p->d_segs_t, p->d_wt);
    NB_LAUNCHED();
    p->wt_fresh = true;
  }
  NB_CUDA(cudaMemsetAsync(absmax, 0, 256, stream));
  absmax_kernel<<<std::min<int64_t>(ceil_div(n_points * 4, 1024), 4 * sm_count()), 256, 0, stream>>>(d_raw, n_points * 4, absmax);
  NB_LAUNCHED();
  BwArgs ba{};
  ba.wt = p->d_wt;
  ba.d_raw = d_raw;
  ba.absmax_bits = absmax;
  ba.xrec = base + L.off_xrec;
  ba.mrec = base + L.off_mrec;
  ba.drec = base + L.off_drec;
  ba.n_points = n_points;
  ba.rec_blocks = p->plan.rec_blocks;
  ba.n_g_blocks = p->n_g_blocks;
  const int grid = (int)std::min<int64_t>(L.n_tiles, bwd_sms());
  mlp_tc_dgrad_kernel<<<grid, kThreadsBw, p->bw_smem_bytes, stream>>>(p->bw, ba);
  NB_LAUNCHED();
  WgArgs wa{};
  wa.xrec = ba.xrec;
  wa.drec = ba.drec;
  wa.part_w = reinterpret_cast<float*>(base + L.off_pw);
  wa.part_b = reinterpret_cast<float*>(base + L.off_pb);
  wa.n_tiles = L.n_tiles;
  wa.rec_blocks = p->plan.rec_blocks;
  wa.drec_blocks = p->bw.drec_blocks;
  wg_assign(p->wg, L.n_tiles, bwd_sms(), wa.part0);
#ifdef NB_TIMING_EXPERIMENTS
  {
    static const int dbg = getenv("NB_WG_DEBUG") ? atoi(getenv("NB_WG_DEBUG")) : 0;
    wa.dbg = dbg;
  }
#endif
  const int wgrid = wa.part0[p->wg.n_units];
  mlp_tc_wgrad_kernel<<<wgrid, kThreadsWg, wg_smem_bytes(), stream>>>(p->wg, wa);
  NB_LAUNCHED();
  dim3 rgrid((unsigned)ceil_div(kPartFloats / 4 + 256, 256), (unsigned)p->wg.n_units);
  wgrad_reduce_kernel<<<rgrid, 256, 0, stream>>>(p->wg, wa, absmax, grad);
  NB_LAUNCHED();
  return NB_OK;
}

}  // namespace nb

extern "C" int nb_mlp_record_layout(const nb_model_t* m, int64_t n_points, const void* workspace, int64_t* xrec_offset,
                                    int* xrec_blocks, int64_t* drec_offset, int* drec_blocks) {
  using namespace nb;
  NB_REQUIRE(m != nullptr && m->packed != nullptr && m->packed->bw_supported,
             "nb_mlp_record_layout: model is not packed for the tensor-core train path");
  const BwLayout L = bw_layout(m->packed, n_points);
  const uintptr_t w = reinterpret_cast<uintptr_t>(workspace);
  const uintptr_t base = (w + 1023) & ~(uintptr_t)1023;
  if (xrec_offset) *xrec_offset = (int64_t)(base - w + L.off_xrec);
  if (drec_offset) *drec_offset = (int64_t)(base - w + L.off_drec);
  if (xrec_blocks) *xrec_blocks = m->packed->plan.rec_blocks;
  if (drec_blocks) *drec_blocks = m->packed->bw.drec_blocks;
  return NB_OK;
}

extern "C" int nb_mlp_mask_record_layout(const nb_model_t* m, int64_t n_points, const void* workspace, int64_t* mrec_offset,
                                         int* mrec_blocks) {
  using namespace nb;
  NB_REQUIRE(m != nullptr && m->packed != nullptr && m->packed->bw_supported,
             "nb_mlp_mask_record_layout: model is not packed for the tensor-core train path");
  const BwLayout L = bw_layout(m->packed, n_points);
  const uintptr_t w = reinterpret_cast<uintptr_t>(workspace);
  const uintptr_t base = (w + 1023) & ~(uintptr_t)1023;
  if (mrec_offset) *mrec_offset = (int64_t)(base - w + L.off_mrec);
  if (mrec_blocks) *mrec_blocks = m->packed->plan.rec_blocks - 2;
  return NB_OK;
}
