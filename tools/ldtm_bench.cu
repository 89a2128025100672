// Micro-benchmark: what does reading an accumulator out of TMEM cost?  Every CTA of a full grid (148) runs W warps, each
// issuing `tcgen05.ld.32x32b.x32` (32 lanes x 32 columns x 4 B = 4 KB per instruction) on its own lane quarter in a loop,
// with `tcgen05.wait::ld` after every `per_wait` loads; optionally one more warp keeps the tensor pipe busy with
// N = 256 MMAs into the other half of TMEM (does the epilogue's read compete with the MMA's accumulate traffic?).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I nerf_kinematics_b200/csrc -I include \
//        tools/ldtm_bench.cu -o tools/bin/ldtm_bench && tools/bin/ldtm_bench
#include <cstdio>
#include <cuda_runtime.h>

#include "tc_common.cuh"
using namespace nb;
constexpr uint64_t kDescHiSw128 = ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)2 << 61);

__global__ void __launch_bounds__(32 * 17, 1) bench(int n_ld_warps, int reps, int per_wait, int with_mma, long long* out,
                                                     uint32_t* sink) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  __shared__ int stop;
  for (int i = threadIdx.x; i < 96 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) {
    stop = 0;
    mbar_init(smem_u32(&bar), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  fence_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  const int warp = threadIdx.x >> 5;
  long long t0 = 0, t1 = 0;
  if (warp < n_ld_warps) {
    const int q = warp & 3;
    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + ((warp >> 2) & 3) * 32;
    uint32_t acc = 0;
    uint32_t v[32];
    t0 = clock64();
    for (int r = 0; r < reps; r += per_wait) {
      for (int k = 0; k < per_wait; ++k) {
        NB_TMEM_LD32(taddr + ((r + k) & 3) * 32, v);
      }
      tc_wait_ld();
      acc += v[0] ^ v[31];
    }
    t1 = clock64();
    if (acc == 0x12345u) sink[0] = acc;
    if ((threadIdx.x & 31) == 0) out[blockIdx.x * 16 + warp] = t1 - t0;
    if (warp == 0 && (threadIdx.x & 31) == 0) *((volatile int*)&stop) = 1;
  } else if (warp == 16 && with_mma) {
    const uint32_t a0 = smem_u32(smem), b0 = smem_u32(smem + 32 * 1024);
    const uint64_t a_desc = kDescHiSw128 | (uint64_t)((a0 & 0x3FFFFu) >> 4);
    const uint64_t b_desc = kDescHiSw128 | (uint64_t)((b0 & 0x3FFFFu) >> 4);
    const uint32_t idesc = umma_idesc(256);
    uint32_t phase = 0;
    long long n = 0;
    while (*((volatile int*)&stop) == 0) {
      if (elect_one()) {
        for (int k = 0; k < 16; ++k) tc_mma(tmem_base + 256, a_desc + 2 * (k & 3), b_desc + 2 * (k & 3), idesc, k ? 1u : 0u);
        tc_commit(smem_u32(&bar));
      }
      __syncwarp();
      mbar_wait(smem_u32(&bar), phase);
      phase ^= 1;
      n += 16;
    }
    if ((threadIdx.x & 31) == 0) out[148 * 16 + blockIdx.x] = n;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
}

int main() {
  long long* out;
  uint32_t* sink;
  cudaMalloc(&out, (148 * 17) * sizeof(long long));
  cudaMalloc(&sink, 64);
  cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  const int reps = 4096;
  printf("tcgen05.ld.32x32b.x32 (4 KB per warp instruction), 148 CTAs, %d loads per warp\n", reps);
  printf("%8s %9s %5s | %14s %12s %14s | %s\n", "ld warps", "per wait", "mma", "cyc / ld /warp", "B/clk per SM", "B/clk per SMSP",
         "MMAs (N=256) per CTA meanwhile, cycles per MMA");
  for (int with_mma = 0; with_mma < 2; ++with_mma)
    for (int w : {1, 4, 8, 16})
      for (int per_wait : {1, 2, 4}) {
        cudaMemset(out, 0, (148 * 17) * sizeof(long long));
        bench<<<148, 32 * 17, 100 * 1024>>>(w, reps, per_wait, with_mma, out, sink);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); return 1; }
        static long long h[148 * 17];
        cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
        double cyc = 0;
        for (int b = 0; b < 148; ++b) {
          long long m = 0;
          for (int k = 0; k < w; ++k) m = h[b * 16 + k] > m ? h[b * 16 + k] : m;
          cyc += (double)m;
        }
        cyc /= 148;
        double mm = 0;
        for (int b = 0; b < 148; ++b) mm += (double)h[148 * 16 + b];
        mm /= 148;
        const double bytes = (double)w * reps * 4096.0;
        printf("%8d %9d %5d | %14.1f %12.1f %14.1f | %.0f  %.1f\n", w, per_wait, with_mma, cyc / reps, bytes / cyc,
               bytes / cyc / (w < 4 ? w : 4), mm, mm > 0 ? cyc / mm : 0.0);
      }
  return 0;
}
