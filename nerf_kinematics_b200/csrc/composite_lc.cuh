// Alpha compositing, lane-contiguous form (included by volume.cu; S == 32 * K samples per ray, K in {2, 4, 6}: the 64 / 128 /
// 192 samples of every BASELINE config).
//
// The strided kernels in volume.cu (lane owns samples lane, lane + 32, ...) are issue-bound, not HBM-bound: K transmittance
// scans, K suffix scans and two shuffles per sample for the neighbour depth, 370-1 070 warp instructions per ray.  Here a lane
// owns K CONSECUTIVE samples: the products / suffix sums run sequentially inside the lane and ONE warp scan joins the 32 lane
// totals; the neighbour depth is just the next word of shared memory.  The data gets there the Blackwell way: a ray's raw
// [S, 4] and z [S] rows are contiguous in HBM, so one elected lane asks the bulk-copy engine for them
// (cp.async.bulk global -> shared, completion on a per-warp mbarrier) -- the NEXT ray's request is in flight while the current
// ray is composited (double buffer per warp, persistent grid), no load instructions, no registers tied up by prefetched
// samples -- and the per-sample outputs (weights, d_raw) leave through shared memory and one bulk store per ray.
// Per-sample formulas and their association are those of composite_fwd_kernel / composite_bwd_ray (raw2outputs,
// nerf/run_nerf.py:93-163); only the order in which the products / sums of a ray are associated differs.

__device__ __forceinline__ void bulk_wait_read1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }

constexpr int kLcBarBytes = 256;  // up to 4 mbarriers per warp at the start of the block's shared memory

template <int K>
struct LcLayout {
  static constexpr int S = 32 * K;
  static constexpr int in_bytes = S * 20;                    // raw [S] float4 | z [S] float  (one buffer)
  // input buffers per warp = rays requested ahead + 1: a short ray is composited in less time than its copy takes to land,
  // so 64-sample rays keep three requests in flight (measured with two buffers: 0.48 ms per 1 M rays, slower than the
  // strided kernel's 0.43), 192-sample rays one
  static constexpr int n_in = K <= 2 ? 4 : (K <= 4 ? 3 : 2);
  static constexpr int fwd_warp_bytes = n_in * in_bytes + 2 * S * 4;    // + weights staging x 2
  static constexpr int bwd_warp_bytes = n_in * in_bytes + 2 * S * 16;   // + d_raw staging x 2
};

// one elected lane: request ray `ray` into input buffer `dst` (shared-space address), completion on `bar`
template <int K>
__device__ __forceinline__ void lc_request(const float4* __restrict__ raw, const float* __restrict__ z, int64_t ray,
                                           uint32_t dst, uint32_t bar) {
  constexpr int S = 32 * K;
  mbar_expect_tx(bar, S * 20);
  bulk_g2s(dst, raw + ray * S, S * 16, bar);
  bulk_g2s(dst + S * 16, z + ray * S, S * 4, bar);
}

// depths of this lane's samples and the one after them (the next lane's first; unused for the ray's last sample)
template <int K>
__device__ __forceinline__ void lc_load_z(const float* sz, int lane, float (&zs)[K + 1]) {
#pragma unroll
  for (int j = 0; j < K; j += 2) {
    const float2 t = *reinterpret_cast<const float2*>(sz + K * lane + j);
    zs[j] = t.x;
    zs[j + 1] = t.y;
  }
  zs[K] = lane < 31 ? sz[K * lane + K] : 0.0f;
}

// a ray's direction, fetched one iteration ahead of its use (a dependent global load per ray would otherwise sit on every
// iteration's critical path: the samples themselves are already in shared memory by then)
__device__ __forceinline__ float3 lc_load_dir(const float* __restrict__ rays, int stride, int64_t ray) {
  const float* p = rays + ray * stride;
  return make_float3(p[3], p[4], p[5]);
}
__device__ __forceinline__ float lc_norm(const float3 d) { return sqrtf((d.x * d.x + d.y * d.y) + d.z * d.z); }

// exclusive product scan of the lane totals
__device__ __forceinline__ float lc_excl_prod(float P, int lane) {
  float incl = P;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const float t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl = incl * t;
  }
  float excl = __shfl_up_sync(0xffffffffu, incl, 1);
  if (lane == 0) excl = 1.0f;
  return excl;
}

// ------------------------------------------------------------------------------------ forward
template <int K, bool kNoise>
__global__ void __launch_bounds__(kThreads)
composite_fwd_lc_kernel(const float4* __restrict__ raw, const float* __restrict__ z, const float* __restrict__ rays,
                        int stride, int64_t n, const float* __restrict__ noise, Rng rng, int white,
                        float* __restrict__ rgb, float* __restrict__ disp, float* __restrict__ acc,
                        float* __restrict__ depth, float* __restrict__ weights) {
  constexpr int S = 32 * K;
  using L = LcLayout<K>;
  extern __shared__ __align__(128) unsigned char lc_smem[];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int64_t warps = (int64_t)gridDim.x * kWarpsPerBlock;
  int64_t ray = (int64_t)blockIdx.x * kWarpsPerBlock + wib;
  if (ray >= n) return;
  unsigned char* base = lc_smem + kLcBarBytes + wib * L::fwd_warp_bytes;
  constexpr int NB = L::n_in;
  const uint32_t bar0 = smem_u32(lc_smem) + wib * 32;
  const uint32_t in0 = smem_u32(base);
  if (lane == 0) {
#pragma unroll
    for (int i = 0; i < NB; ++i) mbar_init(bar0 + 8 * i, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
#pragma unroll
    for (int i = 0; i < NB - 1; ++i)
      if (ray + i * warps < n) lc_request<K>(raw, z, ray + i * warps, in0 + i * L::in_bytes, bar0 + 8 * i);
  }
  __syncwarp();
  uint32_t phase = 0;  // bit b = parity to wait for on buffer b
  int b = 0, bn = NB - 1;  // buffer of this ray / of the ray requested now (NB - 1 rays ahead)
  float3 dir_next = lc_load_dir(rays, stride, ray);
  for (int it = 0; ray < n; ray += warps, ++it) {
    const int64_t nxt = ray + (NB - 1) * warps;
    // buffer bn was read by the previous iteration, which ended with __syncwarp()
    if (nxt < n && lane == 0) lc_request<K>(raw, z, nxt, in0 + bn * L::in_bytes, bar0 + 8 * bn);
    const float3 dir = dir_next;
    if (ray + warps < n) dir_next = lc_load_dir(rays, stride, ray + warps);
    mbar_wait(bar0 + 8 * b, (phase >> b) & 1u);
    phase ^= 1u << b;
    const float4* sraw = reinterpret_cast<const float4*>(base + b * L::in_bytes);
    const float* sz = reinterpret_cast<const float*>(base + b * L::in_bytes + S * 16);
    float zs[K + 1];
    lc_load_z<K>(sz, lane, zs);
    const float dnorm = lc_norm(dir);
    float4 rw[K];
    float al[K], pj[K];
    float P = 1.0f;
#pragma unroll
    for (int j = 0; j < K; ++j) {
      const int s = K * lane + j;
      rw[j] = sraw[s];
      if (kNoise) {
        if (noise != nullptr) rw[j].w = rw[j].w + ld_stream(noise + ray * S + s);
        else rw[j].w = rw[j].w + rng_normal(rng, ray, s);
      }
      const float dist = ((j == K - 1 && lane == 31) ? 1e10f : (zs[j + 1] - zs[j])) * dnorm;
      const float sigma = fmaxf(rw[j].w, 0.0f);
      const float alpha = 1.0f - exp_neg(-sigma * dist);
      al[j] = alpha;
      pj[j] = P;
      P = P * ((1.0f - alpha) + 1e-10f);
    }
    const float E = lc_excl_prod(P, lane);
    float cr = 0.f, cg = 0.f, cb = 0.f, cd = 0.f, ca = 0.f;
    float wv[K];
#pragma unroll
    for (int j = 0; j < K; ++j) {
      const float w = al[j] * (E * pj[j]);
      cr += w * sigmoidf_(rw[j].x);
      cg += w * sigmoidf_(rw[j].y);
      cb += w * sigmoidf_(rw[j].z);
      cd += w * zs[j];
      ca += w;
      wv[j] = w;
    }
    if (weights != nullptr) {
      float* so = reinterpret_cast<float*>(base + NB * L::in_bytes + (it & 1) * S * 4);
      if (lane == 0) bulk_wait_read1();  // the store that last read this staging buffer (two rays ago) is done with it
      __syncwarp();
#pragma unroll
      for (int j = 0; j < K; j += 2) *reinterpret_cast<float2*>(so + K * lane + j) = make_float2(wv[j], wv[j + 1]);
      fence_async_smem();
      __syncwarp();
      if (lane == 0) {
        bulk_s2g(weights + ray * S, smem_u32(so), S * 4);
        bulk_commit();
      }
    }
    cr = warp_sum(cr); cg = warp_sum(cg); cb = warp_sum(cb); cd = warp_sum(cd); ca = warp_sum(ca);
    if (lane == 0) {
      if (white) {
        const float bg = 1.0f - ca;
        cr += bg; cg += bg; cb += bg;
      }
      rgb[ray * 3 + 0] = cr;
      rgb[ray * 3 + 1] = cg;
      rgb[ray * 3 + 2] = cb;
      if (acc != nullptr) acc[ray] = ca;
      if (depth != nullptr) depth[ray] = cd;
      if (disp != nullptr) {
        const float q = cd / ca;  // NaN (acc == 0) propagates like torch.max / tf.maximum
        disp[ray] = (q != q) ? q : 1.0f / fmaxf(1e-10f, q);
      }
    }
    __syncwarp();  // every lane is done with input buffer b before it is requested again
    bn = b;
    b = b + 1 == NB ? 0 : b + 1;
  }
  if (lane == 0) bulk_wait_read0();  // shared memory must outlive the stores reading it
}

// ------------------------------------------------------------------------------------ forward, R rays per warp
// Short rays (64 samples: every coarse pass) leave the one-ray-per-warp kernel above issue-bound (~370 warp instructions per
// 1.5 KB: the scans, reductions, barrier handling and the per-ray epilogue cost the same for 64 samples as for 192).  Here a
// warp composites R CONSECUTIVE rays at once: LPR = 32 / R lanes per ray, K consecutive samples per lane (S = LPR * K), the
// scans and reductions run segmented (shuffle width LPR) and serve all R rays with the same instructions.  R consecutive
// rays' raw / z rows are contiguous in HBM, so a group is still ONE pair of bulk copies and one bulk store (the last group
// of an n not divisible by R is shorter; its missing rays compute on stale shared memory and store nothing).  A ray's
// arithmetic does not depend on its place in the group: results are bit-identical for every chunking of a frame.
// Measured per 1 M 64-sample rays: 0.427 ms (one ray per warp) -> 0.321 ms = 0.785 of the HBM peak with R = 2, K = 4 and two
// input buffers (three: 0.344, four: 0.382 -- fewer resident warps; R = 4, K = 8: 0.43-0.55, 80-112 registers).  The same
// change does nothing for the backward (0.70-0.72 against 0.724), which is not issue-bound at 64 samples.
template <int K, int R, int NB, bool kNoise>
__global__ void __launch_bounds__(kThreads)
composite_fwd_lcm_kernel(const float4* __restrict__ raw, const float* __restrict__ z, const float* __restrict__ rays,
                         int stride, int64_t n, const float* __restrict__ noise, Rng rng, int white,
                         float* __restrict__ rgb, float* __restrict__ disp, float* __restrict__ acc,
                         float* __restrict__ depth, float* __restrict__ weights) {
  constexpr int SW = 32 * K, S = SW / R, LPR = 32 / R;
  constexpr int in_bytes = SW * 20;
  constexpr int warp_bytes = NB * in_bytes + 2 * SW * 4;
  extern __shared__ __align__(128) unsigned char lc_smem[];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int sub = lane / LPR, sl = lane % LPR;
  const int64_t ngrp = (n + R - 1) / R;
  const int64_t warps = (int64_t)gridDim.x * kWarpsPerBlock;
  int64_t grp = (int64_t)blockIdx.x * kWarpsPerBlock + wib;
  if (grp >= ngrp) return;
  unsigned char* base = lc_smem + kLcBarBytes + wib * warp_bytes;
  const uint32_t bar0 = smem_u32(lc_smem) + wib * 32;
  const uint32_t in0 = smem_u32(base);
  auto request = [&](int64_t g, int buf) {   // one elected lane
    const int64_t left = n - g * R;
    const uint32_t nv = left < R ? (uint32_t)left : (uint32_t)R;
    const uint32_t bar = bar0 + 8 * buf, dst = in0 + buf * in_bytes;
    mbar_expect_tx(bar, nv * (S * 20));
    bulk_g2s(dst, raw + g * SW, nv * (S * 16), bar);
    bulk_g2s(dst + SW * 16, z + g * SW, nv * (S * 4), bar);
  };
  if (lane == 0) {
#pragma unroll
    for (int i = 0; i < NB; ++i) mbar_init(bar0 + 8 * i, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
#pragma unroll
    for (int i = 0; i < NB - 1; ++i)
      if (grp + i * warps < ngrp) request(grp + i * warps, i);
  }
  __syncwarp();
  uint32_t phase = 0;
  int b = 0, bn = NB - 1;
  auto ray_of = [&](int64_t g) {   // this lane's ray of group g, clamped into the array for the loads of a short group
    const int64_t r = g * R + sub;
    return r < n ? r : n - 1;
  };
  float3 dir_next = lc_load_dir(rays, stride, ray_of(grp));
  for (int it = 0; grp < ngrp; grp += warps, ++it) {
    const int64_t nxt = grp + (NB - 1) * warps;
    if (nxt < ngrp && lane == 0) request(nxt, bn);
    const float3 dir = dir_next;
    if (grp + warps < ngrp) dir_next = lc_load_dir(rays, stride, ray_of(grp + warps));
    const int64_t ray = grp * R + sub;
    const bool valid = ray < n;
    mbar_wait(bar0 + 8 * b, (phase >> b) & 1u);
    phase ^= 1u << b;
    const float4* sraw = reinterpret_cast<const float4*>(base + b * in_bytes);
    const float* sz = reinterpret_cast<const float*>(base + b * in_bytes + SW * 16);
    float zs[K + 1];
#pragma unroll
    for (int j = 0; j < K; j += 2) {
      const float2 t = *reinterpret_cast<const float2*>(sz + K * lane + j);
      zs[j] = t.x;
      zs[j + 1] = t.y;
    }
    zs[K] = sl < LPR - 1 ? sz[K * lane + K] : 0.0f;
    const float dnorm = lc_norm(dir);
    float4 rw[K];
    float al[K], pj[K];
    float P = 1.0f;
#pragma unroll
    for (int j = 0; j < K; ++j) {
      const int s = K * sl + j;   // sample index within the ray
      rw[j] = sraw[K * lane + j];
      if (kNoise) {
        if (noise != nullptr) rw[j].w = rw[j].w + (valid ? ld_stream(noise + ray * S + s) : 0.0f);
        else rw[j].w = rw[j].w + rng_normal(rng, ray, s);
      }
      const float dist = ((j == K - 1 && sl == LPR - 1) ? 1e10f : (zs[j + 1] - zs[j])) * dnorm;
      const float sigma = fmaxf(rw[j].w, 0.0f);
      const float alpha = 1.0f - exp_neg(-sigma * dist);
      al[j] = alpha;
      pj[j] = P;
      P = P * ((1.0f - alpha) + 1e-10f);
    }
    // exclusive product scan of the lane totals within the ray's LPR lanes
    float incl = P;
#pragma unroll
    for (int o = 1; o < LPR; o <<= 1) {
      const float t = __shfl_up_sync(0xffffffffu, incl, o, LPR);
      if (sl >= o) incl = incl * t;
    }
    float E = __shfl_up_sync(0xffffffffu, incl, 1, LPR);
    if (sl == 0) E = 1.0f;
    float cr = 0.f, cg = 0.f, cb = 0.f, cd = 0.f, ca = 0.f;
    float wv[K];
#pragma unroll
    for (int j = 0; j < K; ++j) {
      const float w = al[j] * (E * pj[j]);
      cr += w * sigmoidf_(rw[j].x);
      cg += w * sigmoidf_(rw[j].y);
      cb += w * sigmoidf_(rw[j].z);
      cd += w * zs[j];
      ca += w;
      wv[j] = w;
    }
    if (weights != nullptr) {
      float* so = reinterpret_cast<float*>(base + NB * in_bytes + (it & 1) * SW * 4);
      if (lane == 0) bulk_wait_read1();
      __syncwarp();
#pragma unroll
      for (int j = 0; j < K; j += 2) *reinterpret_cast<float2*>(so + K * lane + j) = make_float2(wv[j], wv[j + 1]);
      fence_async_smem();
      __syncwarp();
      if (lane == 0) {
        const int64_t left = n - grp * R;
        const uint32_t nv = left < R ? (uint32_t)left : (uint32_t)R;
        bulk_s2g(weights + grp * SW, smem_u32(so), nv * (S * 4));
        bulk_commit();
      }
    }
#pragma unroll
    for (int o = LPR / 2; o > 0; o >>= 1) {
      cr += __shfl_xor_sync(0xffffffffu, cr, o);
      cg += __shfl_xor_sync(0xffffffffu, cg, o);
      cb += __shfl_xor_sync(0xffffffffu, cb, o);
      cd += __shfl_xor_sync(0xffffffffu, cd, o);
      ca += __shfl_xor_sync(0xffffffffu, ca, o);
    }
    if (sl == 0 && valid) {
      if (white) {
        const float bg = 1.0f - ca;
        cr += bg; cg += bg; cb += bg;
      }
      rgb[ray * 3 + 0] = cr;
      rgb[ray * 3 + 1] = cg;
      rgb[ray * 3 + 2] = cb;
      if (acc != nullptr) acc[ray] = ca;
      if (depth != nullptr) depth[ray] = cd;
      if (disp != nullptr) {
        const float q = cd / ca;  // NaN (acc == 0) propagates like torch.max / tf.maximum
        disp[ray] = (q != q) ? q : 1.0f / fmaxf(1e-10f, q);
      }
    }
    __syncwarp();  // every lane is done with input buffer b before it is requested again
    bn = b;
    b = b + 1 == NB ? 0 : b + 1;
  }
  if (lane == 0) bulk_wait_read0();
}

// ------------------------------------------------------------------------------------ backward
// Three sweeps over the lane's K samples, the samples re-read from shared memory instead of kept in registers:
//   1. exp(-sigma dist) and the local transmittance products; one product scan (and, only when d_disp is given, the two sums
//      acc and depth that its gradient needs);
//   2. (reverse) colours, g = <d_rgb, c> + gA + gD z, the local suffix sums of g w; one suffix scan;
//   3. the gradients (composite_bwd_ray's expressions), staged in shared memory and bulk-stored as one [S, 4] row.
template <int K, bool kNoise>
__global__ void __launch_bounds__(kThreads)
composite_bwd_lc_kernel(const float4* __restrict__ raw, const float* __restrict__ z, const float* __restrict__ rays,
                        int stride, int64_t n, const float* __restrict__ noise, Rng rng, int white,
                        const float* __restrict__ d_rgb, const float* __restrict__ d_disp,
                        const float* __restrict__ d_acc, const float* __restrict__ d_depth, float4* __restrict__ d_raw) {
  constexpr int S = 32 * K;
  using L = LcLayout<K>;
  extern __shared__ __align__(128) unsigned char lc_smem[];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int64_t warps = (int64_t)gridDim.x * kWarpsPerBlock;
  int64_t ray = (int64_t)blockIdx.x * kWarpsPerBlock + wib;
  if (ray >= n) return;
  unsigned char* base = lc_smem + kLcBarBytes + wib * L::bwd_warp_bytes;
  constexpr int NB = L::n_in;
  const uint32_t bar0 = smem_u32(lc_smem) + wib * 32;
  const uint32_t in0 = smem_u32(base);
  if (lane == 0) {
#pragma unroll
    for (int i = 0; i < NB; ++i) mbar_init(bar0 + 8 * i, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
#pragma unroll
    for (int i = 0; i < NB - 1; ++i)
      if (ray + i * warps < n) lc_request<K>(raw, z, ray + i * warps, in0 + i * L::in_bytes, bar0 + 8 * i);
  }
  __syncwarp();
  uint32_t phase = 0;  // bit b = parity to wait for on buffer b
  int b = 0, bn = NB - 1;  // buffer of this ray / of the ray requested now (NB - 1 rays ahead)
  // per-ray scalars (direction, upstream gradients) are fetched one iteration ahead, like the samples
  float3 dir_next = lc_load_dir(rays, stride, ray);
  float3 g_next = make_float3(0.f, 0.f, 0.f);
  float gA_next = 0.f, gD_next = 0.f;
  auto fetch_grads = [&](int64_t r) {
    if (d_rgb != nullptr) g_next = make_float3(d_rgb[r * 3], d_rgb[r * 3 + 1], d_rgb[r * 3 + 2]);
    if (d_acc != nullptr) gA_next = d_acc[r];
    if (d_depth != nullptr) gD_next = d_depth[r];
  };
  fetch_grads(ray);
  for (int it = 0; ray < n; ray += warps, ++it) {
    const int64_t nxt = ray + (NB - 1) * warps;
    // buffer bn was read by the previous iteration, which ended with __syncwarp()
    if (nxt < n && lane == 0) lc_request<K>(raw, z, nxt, in0 + bn * L::in_bytes, bar0 + 8 * bn);
    const float gr = g_next.x, gg = g_next.y, gb = g_next.z;
    float gA = gA_next, gD = gD_next;
    const float dnorm = lc_norm(dir_next);
    if (ray + warps < n) {
      dir_next = lc_load_dir(rays, stride, ray + warps);
      fetch_grads(ray + warps);
    }
    if (white) gA -= (gr + gg + gb);
    mbar_wait(bar0 + 8 * b, (phase >> b) & 1u);
    phase ^= 1u << b;
    const float4* sraw = reinterpret_cast<const float4*>(base + b * L::in_bytes);
    const float* sz = reinterpret_cast<const float*>(base + b * L::in_bytes + S * 16);
    float zs[K + 1];
    lc_load_z<K>(sz, lane, zs);
    // ---- sweep 1
    float ev[K], pj[K];
    uint32_t pos = 0;  // bit j: the (noisy) sigma input of sample j is positive
    float P = 1.0f;
#pragma unroll
    for (int j = 0; j < K; ++j) {
      const int s = K * lane + j;
      float sg = sraw[s].w;
      if (kNoise) {
        if (noise != nullptr) sg = sg + ld_stream(noise + ray * S + s);
        else sg = sg + rng_normal(rng, ray, s);
      }
      if (sg > 0.0f) pos |= 1u << j;
      const float dist = ((j == K - 1 && lane == 31) ? 1e10f : (zs[j + 1] - zs[j])) * dnorm;
      const float e = exp_neg(-fmaxf(sg, 0.0f) * dist);
      ev[j] = e;
      pj[j] = P;
      P = P * ((1.0f - (1.0f - e)) + 1e-10f);
    }
    const float E = lc_excl_prod(P, lane);
    if (d_disp != nullptr) {  // warp-uniform
      float ca = 0.f, cd = 0.f;
#pragma unroll
      for (int j = 0; j < K; ++j) {
        const float w = (1.0f - ev[j]) * (E * pj[j]);
        ca += w;
        cd += w * zs[j];
      }
      ca = warp_sum(ca);
      cd = warp_sum(cd);
      const float q = cd / ca;
      if (q > 1e-10f) {
        const float gq = -d_disp[ray] / (q * q);
        gD += gq / ca;
        gA -= gq * cd / (ca * ca);
      }
    }
    // ---- sweep 2 (reverse): g and the local suffix sums of g * w
    float gv[K], ls[K], cR[K], cG[K], cB[K];
    float run = 0.0f;
#pragma unroll
    for (int j = K - 1; j >= 0; --j) {
      const float4 r = sraw[K * lane + j];
      cR[j] = sigmoidf_(r.x);
      cG[j] = sigmoidf_(r.y);
      cB[j] = sigmoidf_(r.z);
      const float g = (gr * cR[j] + gg * cG[j] + gb * cB[j]) + gA + gD * zs[j];
      gv[j] = g;
      ls[j] = run;
      run += g * ((1.0f - ev[j]) * (E * pj[j]));
    }
    float incl = run;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const float t = __shfl_down_sync(0xffffffffu, incl, o);
      if (lane + o < 32) incl += t;
    }
    float X = __shfl_down_sync(0xffffffffu, incl, 1);  // sum of g * w over the lanes after this one
    if (lane == 31) X = 0.0f;
    // ---- sweep 3: gradients -> staging buffer -> one bulk store
    float4* so = reinterpret_cast<float4*>(base + NB * L::in_bytes + (it & 1) * S * 16);
    if (lane == 0) bulk_wait_read1();
    __syncwarp();
#pragma unroll
    for (int j = 0; j < K; ++j) {
      const float e = ev[j];
      const float T = E * pj[j];
      const float w = (1.0f - e) * T;
      const float om = (1.0f - (1.0f - e)) + 1e-10f;
      const float dist = ((j == K - 1 && lane == 31) ? 1e10f : (zs[j + 1] - zs[j])) * dnorm;
      const float d_alpha = gv[j] * T - (X + ls[j]) / om;
      float4 o4;
      o4.x = w * gr * cR[j] * (1.0f - cR[j]);
      o4.y = w * gg * cG[j] * (1.0f - cG[j]);
      o4.z = w * gb * cB[j] * (1.0f - cB[j]);
      o4.w = ((pos >> j) & 1u) ? d_alpha * (dist * e) : 0.0f;
      so[K * lane + j] = o4;
    }
    fence_async_smem();
    __syncwarp();
    if (lane == 0) {
      bulk_s2g(d_raw + ray * S, smem_u32(so), S * 16);
      bulk_commit();
    }
    __syncwarp();
    bn = b;
    b = b + 1 == NB ? 0 : b + 1;
  }
  if (lane == 0) bulk_wait_read0();
}

