// Micro-benchmark behind the MMA-issuer design of csrc/mlp_tc.cu: cycles per tcgen05.mma (kind::f16, M=128, K=16, SS mode,
// K-major operands in shared memory) for two ways of issuing them, every CTA of a full grid (148) running the same loop.
//
//   issuer "thread"  : `if (threadIdx.x == 0)` -- one divergent thread rebuilds both 64-bit descriptors per MMA (what the
//                      first forward kernel did): the instruction stream of that thread, not the tensor pipe, sets the pace
//   issuer "uniform" : the whole warp runs the loop in warp-uniform control flow, descriptors advance in uniform
//                      registers and only the tcgen05 instructions are predicated on an elected lane
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I nerf_kinematics_b200/csrc -I include \
//        tools/umma_bench.cu -o tools/bin/umma_bench && tools/bin/umma_bench
//
// Measured on B200 (profiles/r1b_kernel_timings_cuda_events.txt): uniform issuer 52 / 68 / 132 cycles per MMA at
// N = 64 / 128 / 256 (floors 32 / 64 / 128); single thread 91 / 91 / 136-237; commit -> mbarrier -> wait round trip ~300.
#include <cstdio>
#include <cuda_runtime.h>

#include "tc_common.cuh"
using namespace nb;

template <bool kUniform>
__global__ void __launch_bounds__(128, 1) bench(int n, int reps, int per_commit, int sw64, long long* out) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;  // fp16 ones
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
  const uint32_t a0 = smem_u32(smem), b0 = smem_u32(smem + 64 * 1024);
  const uint32_t idesc = umma_idesc(n);
  if (kUniform) {
    if (threadIdx.x < 32) {
      const uint64_t ad0 = umma_desc(a0), bd0 = sw64 ? umma_desc_sw64(b0) : umma_desc(b0);
      uint32_t phase = 0;
      const long long t0 = clock64();
      for (int r = 0; r < reps; ++r) {
        for (int i = 0; i < per_commit; i += 4) {
          const uint64_t ad = ad0 + (uint64_t)(((i >> 2) & 3) * (kBlockBytes >> 4));
          const uint64_t bd = bd0 + (uint64_t)(((i >> 2) & 1) * (32768 >> 4));
          if (elect_one()) {
            tc_mma(tmem_base, ad, bd, idesc, i != 0 ? 1u : 0u);
            tc_mma(tmem_base, ad + 2, bd + 2, idesc, 1u);
            tc_mma(tmem_base, ad + 4, bd + (sw64 ? 0 : 4), idesc, 1u);
            tc_mma(tmem_base, ad + 6, bd + (sw64 ? 2 : 6), idesc, 1u);
          }
          __syncwarp();
        }
        if (elect_one()) tc_commit(smem_u32(&bar));
        __syncwarp();
        mbar_wait(smem_u32(&bar), phase);
        phase ^= 1;
      }
      if (blockIdx.x == 0 && threadIdx.x == 0) out[0] = clock64() - t0;
    }
  } else if (threadIdx.x == 0) {
    uint32_t phase = 0;
    const long long t0 = clock64();
    for (int r = 0; r < reps; ++r) {
      for (int i = 0; i < per_commit; ++i) {
        const int blk = i >> 2, k = i & 3;
        const uint64_t bd = sw64 ? umma_desc_sw64(b0 + (blk & 1) * 32768 + (k & 1) * 32) : umma_desc(b0 + (blk & 1) * 32768 + k * 32);
        tc_mma(tmem_base, umma_desc(a0 + (blk & 3) * kBlockBytes + k * 32), bd, idesc, i != 0 ? 1u : 0u);
      }
      tc_commit(smem_u32(&bar));
      mbar_wait(smem_u32(&bar), phase);
      phase ^= 1;
    }
    if (blockIdx.x == 0) out[0] = clock64() - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
}

template <bool kUniform>
static void run(const char* name, int n, int pc, int sw64, long long* d) {
  const int grid = 148, reps = 256, smem = 200 * 1024;
  cudaFuncSetAttribute(bench<kUniform>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  bench<kUniform><<<grid, 128, smem>>>(n, reps, pc, sw64, d);  // warm-up
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  cudaEventRecord(e0);
  bench<kUniform><<<grid, 128, smem>>>(n, reps, pc, sw64, d);
  cudaEventRecord(e1);
  const cudaError_t err = cudaDeviceSynchronize();
  float ms = 0.f;
  cudaEventElapsedTime(&ms, e0, e1);
  long long cyc = 0;
  cudaMemcpy(&cyc, d, 8, cudaMemcpyDeviceToHost);
  printf("%-7s issuer  B %-6s N=%3d  %2d MMAs/commit: %6.1f cycles/MMA (floor %3d)  %7.0f TFLOP/s  %s\n", name,
         sw64 ? "SW64" : "SW128", n, pc, (double)cyc / ((double)reps * pc), n / 2,
         2.0 * 128 * n * 16 * (double)reps * pc * grid / (ms * 1e-3) / 1e12, cudaGetErrorString(err));
}

int main() {
  long long* d;
  cudaMalloc(&d, 16);
  for (int n : {64, 128, 256})
    for (int pc : {4, 16, 64}) {
      run<true>("uniform", n, pc, 0, d);
      run<false>("thread", n, pc, 0, d);
    }
  run<true>("uniform", 256, 64, 1, d);
  run<false>("thread", 256, 64, 1, d);
  return 0;
}
