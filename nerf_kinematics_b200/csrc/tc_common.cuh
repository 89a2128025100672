// Shared pieces of the tcgen05 (5th-gen tensor core) MLP kernels: plans, packed-weight images and the PTX wrappers.
//   mlp_tc.cu      forward  (inference and, with a saved-activation record, the forward half of a train step)
//   mlp_tc_bwd.cu  backward (fused dgrad chain, split-K wgrad, head gradients)
#pragma once
#include <cuda_fp16.h>
#include <math.h>

#include "model.h"

// Kernel-side timing hooks (skip loads / MMAs / conversions) exist only in -DNB_TIMING_EXPERIMENTS builds; a release
// build cannot be switched into returning garbage by an environment variable.
#ifdef NB_TIMING_EXPERIMENTS
#define NB_DBG(x) (x)
#else
#define NB_DBG(x) false
#endif

namespace nb {

// Activation scale of the 16-bit path.  Hidden activations are carried as fp16 (11 significand bits, range 65 504);
// shipped checkpoints reach ~1.1e5 in their trunk (wheel_robot6 at 1280x720) and would clamp.  ReLU networks are
// positively homogeneous, so the packed image multiplies every weight that reads an ENCODING block, and every bias, by
// kActScale = 2^-5: all hidden activations become kActScale * h (exact: a power of two; range 2.1e6, values below
// 2e-3 lose relative -- not absolute, <= 2e-6 -- precision), hidden-to-hidden weights are untouched, and the heads'
// fp32 epilogue multiplies by kInvActScale.  The backward sees scaled X records: the wgrad reduce divides the
// sub-matrices whose X operand is an activation block by kActScale (dz and the bias gradients are unaffected).
constexpr float kActScale = 0.03125f;
constexpr float kInvActScale = 32.0f;

constexpr int kMaxStages = 14;
constexpr int kMaxKB = 6;
constexpr int kRing = 4;   // dgrad weight ring (4 x 32 KB since the mask ring is gone)
constexpr int kTileM = 128;
constexpr int kBlockBytes = kTileM * 128;  // one 128-row x 64-element 16-bit K-block, SWIZZLE_128B
enum { SRC_ENCX = 4, SRC_ENCD = 5, SRC_DRAW = 6, SRC_BIAS = 7 };
enum { EPI_HIDDEN = 0, EPI_RGB = 2, EPI_OUT4 = 3 };
// forward kernel: weight producer warp, 2 MMA issuer warps (one per tile), a spare warp, 2 tiles x 8 epilogue warps
constexpr int kThreadsTc = 640;
constexpr int kFwdRing = 3;                 // ring slots of the forward kernel
constexpr int kFwdSlotBytes = 256 * 64;     // one slot = <= 256 weight rows (N) x 32 K-elements, SWIZZLE_64B

// ---- forward plan
struct TcStage {
  int n_kb;
  int a_src[kMaxKB];  // 0..3 activation block, SRC_ENCX, SRC_ENCD
  int n_main;         // N of the stage's MMAs (multiple of 16, <= 256)
  int chunk_bytes;    // n_main * 128
  int w_off;          // byte offset of K-block 0 in the packed weights; K-block k at w_off + k*chunk_bytes, stored as
                      // two half-blocks of 32 K-elements (n_main rows x 64 B, SWIZZLE_64B), one ring slot each
  int b_off;          // float offset of this stage's biases
  int epi, relu;
  int n_act;          // columns stored back as activations (multiple of 64)
  int alpha_dot;      // 1: the epilogue also forms alpha = <activation row, alpha weights> (+ bias) in fp32
  int write_encd;     // 1: the position-encoding block is dead after this stage's MMAs; its epilogue re-fills it
                      //    with the direction encoding
  int rec_blk;        // train: first block of this stage's activations in the per-tile record (X record)
  int step0;          // index of the stage's first entry in the MMA issuers' step program
  int encd_wait;      // 1: the stage's first reader of the direction-encoding block waits for its re-fill
};

struct TcPlan {
  int n_stages, n_act_blocks, Cx, Cd, Lx, Ld, n_bias;
  int n_steps;        // ring slots consumed per tile pair = 2 * sum of the stages' K-blocks
  int alpha_w_off, alpha_b_off;  // float offsets (bias buffer) of the fp32 alpha weights [width] and bias; -1 = none
  int rec_blocks;     // X record: [encx, encd, stage activations...] blocks of kBlockBytes per 128-point tile
  TcStage st[kMaxStages];
};

// ---- backward (dgrad chain) plan: stage s computes dX = (dZ . W) * relu'(X) for one layer, reversed order
struct BwStage {
  int n_kb;
  int a_src[kMaxKB];    // 0..3 gradient block, SRC_DRAW
  int k_steps[kMaxKB];  // 16-element K steps of this block (4, or 1 for the d_raw block)
  int n_main;           // N = width of the layer input whose gradient is produced
  int chunk_bytes;      // n_main * 128
  int w_off;            // byte offset into the transposed packed weights
  int mask_blk;         // first X-record block of the ReLU mask, -1 = none (linear layer input)
  int n_out;            // n_main / 64 gradient blocks written
  int drec_blk;         // first block of this stage's output in the per-tile dZ record
  int step0;            // index of the stage's first entry in the MMA issuer's step program
};
struct BwPlan {
  int n_stages, drec_blocks, ring_stage_bytes;
  int n_steps;          // K-blocks per tile over all stages (entries of the MMA issuer's step program)
  int draw_blk;         // dZ-record block of the scaled 16-bit d_raw tile (operand of the head wgrad units)
  BwStage st[kMaxStages];
};

// ---- wgrad work: unit = one (dZ, X) pair of record block ranges -> one weight sub-matrix
constexpr int kMaxUnits = 20;
constexpr int kMaxXSrc = 6;
struct WgUnit {
  int dz_blk;        // first dZ-record block
  int m_load;        // dZ blocks present in the record (1, 2 or 4)
  int m_blocks;      // 2 or 4 (M = 64 * m_blocks; blocks past m_load are zeros)
  int n_x;           // X blocks (N = 64 * n_x; M*N accumulators must fit 512 TMEM columns)
  int x_blk[kMaxXSrc];     // X-record block of each
  int x_col0[kMaxXSrc];    // destination column of each block's element 0 in the weight matrix
  int x_valid[kMaxXSrc];   // valid columns of each block (<= 64)
  int row0;          // accumulator rows [row0, row0 + out_rows) are weight rows [0, out_rows)
  int out_rows;      // valid rows (layer out)
  long long w_off;   // float offset of weight[0,0]
  int ld;            // weight leading dimension (layer in)
  long long b_off;   // float offset of the bias, -1 = do not produce a bias gradient from this unit
  int cost;          // relative MMA cost per tile
};
struct WgItem {
  int unit, tile0, tile1, part;  // tiles [tile0, tile1) of unit; part = index of this item within its unit
};
struct WgPlan {
  int n_units;
  WgUnit u[kMaxUnits];
};

struct PackSeg {
  long long dst_off;  // bytes, start of the chunk in the packed image
  long long src_off;  // floats, start of the source matrix in the flat parameters
  long long bias_off; // floats, the layer's bias vector (written to chunk element 63), -1 = none
  int row0, n_rows;   // chunk rows [row0, row0+n_rows) <- source rows [0, n_rows)
  int chunk_rows;     // rows of the whole chunk (the stage's N)
  int ld, col0, ncols;  // source leading dim; chunk element k <- source column col0+k for k < ncols, else 0
  float scale;          // kActScale for chunks that multiply an encoding block (and the bias riding in them), else 1
};
// transposed chunk for the dgrad chain: chunk row n (layer input), element kdst0+k <- W[krow0+k, col0+n], k < kcount
struct PackSegT {
  long long dst_off, src_off;
  int n_rows, ld, col0, krow0, kcount, kdst0;
};
struct BiasSeg {
  int dst, count;
  long long src_off;
};
struct PackedBf16 {
  TcPlan plan;
  unsigned char* d_w = nullptr;
  float* d_bias = nullptr;
  PackSeg* d_segs = nullptr;
  BiasSeg* d_bsegs = nullptr;
  int n_segs = 0, n_bsegs = 0;
  size_t w_bytes = 0;
  size_t smem_bytes = 0;
  bool supported = false;
  // backward
  BwPlan bw;
  WgPlan wg;
  unsigned char* d_wt = nullptr;  // transposed images for the dgrad chain
  PackSegT* d_segs_t = nullptr;
  int n_segs_t = 0;
  size_t wt_bytes = 0;
  size_t bw_smem_bytes = 0;
  bool wt_fresh = false;          // d_wt matches the current parameters
  bool bw_supported = false;
  int n_g_blocks = 0;
};

// Operands are IEEE fp16 (tcgen05 kind::f16 takes f16 or bf16 at the same rate): 11 significand bits keep
// the rendered image ~5x closer to the fp32 reference than bf16 does (measured: 2.2e-3 vs 1.07e-2 max-abs on the
// lego checkpoint); conversions saturate at +-65504 instead of overflowing to inf.  Element 2i sits in the low half.
__device__ __forceinline__ uint32_t pack_h2(float lo, float hi, bool relu) {
  uint32_t d;
  if (relu) asm("cvt.rn.relu.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  else asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  return d;
}

// one 16-byte chunk (8 elements) of a transposed weight chunk for the dgrad chain: chunk row n (layer input), element
// kdst0+k <- W[krow0+k, col0+n]; SWIZZLE_128B K-major image (used by pack_weights_t_kernel and the combined pack launch)
__device__ __forceinline__ void pack_weights_t_body(const float* __restrict__ params, const PackSegT& sg,
                                                    unsigned char* __restrict__ out, int idx) {
  if (idx >= sg.n_rows * 8) return;
  const int n = idx >> 3, c = idx & 7;
  float v[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const int kk = c * 8 + e - sg.kdst0;
    v[e] = (kk >= 0 && kk < sg.kcount) ? params[sg.src_off + (long long)(sg.krow0 + kk) * sg.ld + sg.col0 + n] : 0.0f;
  }
  uint4 pk;
  pk.x = pack_h2(v[0], v[1], false);
  pk.y = pack_h2(v[2], v[3], false);
  pk.z = pack_h2(v[4], v[5], false);
  pk.w = pack_h2(v[6], v[7], false);
  *reinterpret_cast<uint4*>(out + sg.dst_off + (long long)n * 128 + ((c ^ (n & 7)) << 4)) = pk;
}

// ------------------------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra DONE;\n"
      "bra WAIT_LOOP;\n"
      "DONE:\n"
      "}\n" ::"r"(bar),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
// L2 eviction-priority hints for the bulk copies: the packed weights are re-read by every CTA for every tile
// (evict_last), the per-tile records stream through once per kernel (evict_first)
__device__ __forceinline__ uint64_t l2_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t l2_evict_last() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ void bulk_g2s_hint(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar, uint64_t pol) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar), "l"(pol)
               : "memory");
}
__device__ __forceinline__ void bulk_s2g_hint(void* dst, uint32_t src, uint32_t bytes, uint64_t pol) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;" ::"l"(dst), "r"(src), "r"(bytes),
               "l"(pol)
               : "memory");
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_mma(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(acc)
      : "memory");
}
// ---- CTA-pair (cta_group::2) variants: one MMA covers M = 256 rows, 128 in each CTA of a 2-CTA cluster; every CTA
// supplies its own A rows and HALF of the B rows (N/2) from its own shared memory, so the B-operand reads and the weight
// ring fill per SM are half of the single-CTA kernel's.  Issued by the leader CTA (cluster rank 0) only.
__device__ __forceinline__ void tc_mma2(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(acc)
      : "memory");
}
// arrives (once the MMAs issued so far have completed) on the barrier at this shared-memory offset in BOTH CTAs
__device__ __forceinline__ void tc_commit2(uint32_t bar) {
  const unsigned short mask = 3;
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
               "h"(mask)
               : "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of the same shared-memory offset in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa_u32(uint32_t saddr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
  return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  // default (CTA-scope release) semantics on purpose: the .release.cluster form costs a MEMBAR.ALL.GPU per arrival.  What
  // the arrival publishes -- operand rows in the ARRIVING CTA's own shared memory, made visible to the async proxy by
  // the fence.proxy.async before it -- is read back by that same SM's tensor core, after the leader has seen the
  // arrival and issued the MMA; nothing written here is read through another SM's generic loads.
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// wait on a barrier the peer CTA arrives on (CTA-scope acquire: the .acquire.cluster form invalidates L1 on every poll)
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "WAIT_LOOP_C:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra DONE_C;\n"
      "bra WAIT_LOOP_C;\n"
      "DONE_C:\n"
      "}\n" ::"r"(bar),
      "r"(parity)
      : "memory");
}
// kind::f16 instruction descriptor of a CTA-pair MMA: M = 256 (128 rows per CTA), N = full width
__device__ __forceinline__ uint32_t umma_idesc2(int n) {
  return (1u << 4) | (0u << 7) | (0u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);
}

// One lane of a converged warp (warp-uniform issue loops predicate their tcgen05 / bulk-copy instructions on it, so
// descriptors and loop state stay in uniform registers).
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n.reg .pred p;\nelect.sync _|p, 0xffffffff;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void tc_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

#define NB_TMEM_LD32(taddr, v)                                                                                      \
  asm volatile(                                                                                                     \
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "                                                                     \
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28," \
      "%29,%30,%31}, [%32];"                                                                                        \
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), \
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),      \
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),     \
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])                   \
      : "r"(taddr))

// tcgen05.wait::ld that the compiler sees as (re)defining the 64 registers of two x32 loads: plain C++ arithmetic on those
// registers (mask bit packing, conditional zeroing) would otherwise be free to move above the wait -- only asm volatile
// statements are ordered with respect to each other -- and read them before the asynchronous load has landed.
#define NB_R8(v, o) "+r"(v[o]), "+r"(v[o + 1]), "+r"(v[o + 2]), "+r"(v[o + 3]), "+r"(v[o + 4]), "+r"(v[o + 5]), "+r"(v[o + 6]), "+r"(v[o + 7])
#define NB_TMEM_WAIT64(v)                                                                                           \
  do {                                                                                                              \
    asm volatile("tcgen05.wait::ld.sync.aligned;" : NB_R8(v, 0), NB_R8(v, 8), NB_R8(v, 16), NB_R8(v, 24)::"memory"); \
    asm volatile("" : NB_R8(v, 32), NB_R8(v, 40), NB_R8(v, 48), NB_R8(v, 56)::"memory");                          \
  } while (0)

__device__ __forceinline__ void tmem_ld4(uint32_t taddr, uint32_t (&v)[4]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3])
               : "r"(taddr));
}
__device__ __forceinline__ uint32_t tmem_ld1(uint32_t taddr) {
  uint32_t v;
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(v) : "r"(taddr));
  return v;
}

// UMMA shared-memory descriptor: K-major operand, SWIZZLE_128B, 8-row groups 1024 B apart (SBO), version 1.
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr) {
  uint64_t d = (uint64_t)((saddr & 0x3FFFFu) >> 4);
  d |= (uint64_t)1 << 16;            // leading byte offset (unused for swizzled K-major), 16 B units
  d |= (uint64_t)(1024 >> 4) << 32;  // stride byte offset
  d |= (uint64_t)1 << 46;            // descriptor version (sm_100)
  d |= (uint64_t)2 << 61;            // SWIZZLE_128B
  return d;
}
// Same for a SWIZZLE_64B K-major operand: rows of 64 B (32 elements), 8-row groups 512 B apart.
__device__ __forceinline__ uint64_t umma_desc_sw64(uint32_t saddr) {
  uint64_t d = (uint64_t)((saddr & 0x3FFFFu) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(512 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)4 << 61;            // SWIZZLE_64B
  return d;
}
// Instruction descriptor kind::f16: D=f32 (bit 4), A/B format 0 = f16 (1 would be bf16), both K-major, M=128.
__device__ __forceinline__ uint32_t umma_idesc(int n) {
  return (1u << 4) | (0u << 7) | (0u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(kTileM >> 4) << 24);
}

// bulk async store shared -> global (completes in a bulk group of the issuing thread)
__device__ __forceinline__ void bulk_s2g(void* dst, uint32_t src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// MN-major operand (the contraction index runs over ROWS of a 128-byte-row SWIZZLE_128B block): 64 MN elements are
// contiguous in a row, 8 rows form a swizzle atom; LBO = byte stride between 64-element MN blocks, SBO = byte
// stride between 8-row K groups (cute::UMMA make_umma_desc<Major::MN>, canonical ((8,n),(8,k)):((1,LBO),(8,SBO))).
__device__ __forceinline__ uint64_t umma_desc_mn(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = (uint64_t)((saddr & 0x3FFFFu) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// kind::f16 instruction descriptor with both operands MN-major (bits 15, 16), f16 inputs, f32 accumulate, M = 128.
__device__ __forceinline__ uint32_t umma_idesc_mn(int n) {
  return (1u << 4) | (1u << 15) | (1u << 16) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(kTileM >> 4) << 24);
}

// per-model host entry points (mlp_tc.cu / mlp_tc_bwd.cu)
int bw_setup(nb_model* m, cudaStream_t stream);  // plans + buffers of the backward; needs p->plan
void bw_free(PackedBf16* p);
size_t tc_train_workspace_bytes(const PackedBf16* p, int64_t n_points);
unsigned char* tc_train_xrec(const PackedBf16* p, void* ws, int64_t n_points);  // X record inside the workspace
unsigned char* tc_train_mrec(const PackedBf16* p, void* ws, int64_t n_points);  // 1-bit ReLU mask record (8 B per row and block)

}  // namespace nb
