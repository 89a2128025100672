// Micro-benchmark: cycles per tcgen05.mma (kind::f16, M=128, K=16, SS mode, SWIZZLE_128B K-major operands in shared
// memory) issued back to back by one thread, for N in {64,128,256}; every CTA of a full grid runs the same loop.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I nerf_kinematics_b200/csrc tools/umma_bench.cu -o /tmp/umma_bench
#include <cstdio>
#include <cuda_runtime.h>
#include "tc_common.cuh"
using namespace nb;

__global__ void __launch_bounds__(640, 1) bench(int n, int reps, int per_commit, int ablocks, int sw64, int alt, int mode, long long* out) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar;
  __shared__ uint64_t bar_never;
  __shared__ uint64_t bar_dummy;
  __shared__ volatile int stop;
  __shared__ uint32_t tmem_slot;
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bar), 1);
    mbar_init(smem_u32(&bar_never), 1);
    mbar_init(smem_u32(&bar_dummy), 1);
    stop = 0;
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
  if (threadIdx.x == 0) {
    const uint32_t a0 = smem_u32(smem), b0 = smem_u32(smem + 64 * 1024);
    const uint32_t idesc = umma_idesc(n);
    uint32_t phase = 0;
    long long t0 = clock64();
    for (int r = 0; r < reps; ++r) {
      for (int i = 0; i < per_commit; ++i) {
        const int blk = i >> 2, k = i & 3;
        const uint64_t bd = sw64 ? umma_desc_sw64(b0 + (blk & 1) * 32768 + (k & 1) * 32) : umma_desc(b0 + (blk & 1) * 32768 + k * 32);
        tc_mma(tmem_base + (alt ? (i & 1) * 256 : 0), umma_desc(a0 + (blk % ablocks) * kBlockBytes + k * 32), bd, idesc,
               i > 1 ? 1u : 0u);
        if ((mode & 1) && (i & 3) == 3) tc_commit(smem_u32(&bar_dummy));   // commit every 4 MMAs, never waited on
        if ((mode & 2) && (i & 3) == 3) tc_fence_after();
      }
      tc_commit(smem_u32(&bar));
      mbar_wait(smem_u32(&bar), phase);
      phase ^= 1;
    }
    long long t1 = clock64();
    if (blockIdx.x == 0) out[0] = t1 - t0;
    stop = 1;
    mbar_arrive(smem_u32(&bar_never));
  } else if ((mode & 4) && threadIdx.x >= 128 && threadIdx.x < 256) {
    // TMEM-reading warps: back-to-back tcgen05.ld while the MMAs run
    const uint32_t lane_addr = ((uint32_t)((threadIdx.x >> 5) & 3) * 32) << 16;
    uint32_t acc = 0;
    while (!stop) {
      uint32_t v[32];
      NB_TMEM_LD32(tmem_base + lane_addr, v);
      tc_wait_ld();
      acc += v[0];
    }
    if (acc == 0x12345) out[1] = acc;
  } else if (threadIdx.x >= 128) {
    // polling warps: spin on a barrier that completes only when the issuer is done (like waiting epilogue warps)
    mbar_wait(smem_u32(&bar_never), 0);
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
}

__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n.reg .pred p;\nelect.sync _|p, 0xffffffff;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(pred));
  return pred != 0;
}

// variant: the whole warp runs the issue loop (warp-uniform control flow, descriptors in uniform registers);
// only the tcgen05 instructions are predicated on one elected lane.
__global__ void __launch_bounds__(128, 1) bench_uniform(int n, int reps, int per_commit, long long* out) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) {
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
  if (threadIdx.x < 32) {
    const uint32_t a0 = smem_u32(smem), b0 = smem_u32(smem + 64 * 1024);
    const uint32_t idesc = umma_idesc(n);
    const uint64_t ad0 = umma_desc(a0), bd0 = umma_desc(b0);
    uint32_t phase = 0;
    long long t0 = clock64();
    for (int r = 0; r < reps; ++r) {
      for (int i = 0; i < per_commit; i += 4) {
        const uint64_t ad = ad0 + (uint64_t)(((i >> 2) & 3) * (kBlockBytes >> 4));
        const uint64_t bd = bd0 + (uint64_t)(((i >> 2) & 1) * (32768 >> 4));
        if (elect_one()) {
          tc_mma(tmem_base, ad, bd, idesc, i != 0 ? 1u : 0u);
          tc_mma(tmem_base, ad + 2, bd + 2, idesc, 1u);
          tc_mma(tmem_base, ad + 4, bd + 4, idesc, 1u);
          tc_mma(tmem_base, ad + 6, bd + 6, idesc, 1u);
        }
        __syncwarp();
      }
      if (elect_one()) tc_commit(smem_u32(&bar));
      __syncwarp();
      mbar_wait(smem_u32(&bar), phase);
      phase ^= 1;
    }
    long long t1 = clock64();
    if (blockIdx.x == 0 && threadIdx.x == 0) out[0] = t1 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
}

int main() {
  long long* d;
  cudaMalloc(&d, 16);
  cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaFuncSetAttribute(bench_uniform, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  for (int n : {64, 128, 256})
    for (int pc : {4, 16, 64}) {
      bench_uniform<<<148, 128, 200 * 1024>>>(n, 256, pc, d);
      cudaEvent_t e0, e1;
      cudaEventCreate(&e0); cudaEventCreate(&e1);
      cudaEventRecord(e0);
      bench_uniform<<<148, 128, 200 * 1024>>>(n, 256, pc, d);
      cudaEventRecord(e1);
      cudaError_t err = cudaDeviceSynchronize();
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      long long cyc; cudaMemcpy(&cyc, d, 8, cudaMemcpyDeviceToHost);
      printf("uniform issuer N=%3d mma/commit=%2d: %.1f cycles/MMA (floor %d) %.3f ms %.0f TFLOP/s %s\n", n, pc,
             (double)cyc / (256.0 * pc), n / 2, ms, 2.0 * 128 * n * 16 * 256.0 * pc * 148 / (ms * 1e-3) / 1e12, cudaGetErrorString(err));
    }
  const int ns[] = {64, 128, 256};
  const int pcs[] = {4, 16, 64};
  (void)ns; (void)pcs;
  for (int threads : {640})
   for (int sw64 : {1})
    for (int alt : {1})
     for (int mode : {0, 1, 3, 4, 7})
      for (int pc : {64}) {
        const int grid = 148, n = 256;
        const int reps = 256;
        bench<<<grid, threads, 200 * 1024>>>(n, reps, pc, 4, sw64, alt, mode, d);
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0); cudaEventCreate(&e1);
        cudaEventRecord(e0);
        bench<<<grid, threads, 200 * 1024>>>(n, reps, pc, 4, sw64, alt, mode, d);
        cudaEventRecord(e1);
        cudaError_t err = cudaDeviceSynchronize();
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        long long cyc; cudaMemcpy(&cyc, d, 8, cudaMemcpyDeviceToHost);
        const double per = (double)cyc / ((double)reps * pc);
        const double tf = 2.0 * 128 * n * 16 * (double)reps * pc * grid / (ms * 1e-3) / 1e12;
        printf("mode %d poll warps %2d sw64=%d alt=%d | grid %3d N=%3d mma/commit=%2d: %.1f cycles/MMA (floor %d)  %.3f ms  %.0f TFLOP/s  %s\n", mode, (threads - 128) / 32, sw64, alt, grid, n, pc, per, n / 2, ms, tf,
               cudaGetErrorString(err));
      }
  return 0;
}
