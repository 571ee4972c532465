// Probe: what does ONE tcgen05.mma.cta_group::1.kind::f16 cost on sm_100a as a function of its N, with the A operand in
// tensor memory (the form qmm_tc.cu / qmm_sk.cu issue: M = 128 weight rows, N = tokens, K = 16) or in shared memory?
// Design input for the quantised GEMM: at decode the token count (N of the MMA) is 16..64, and the measured per-k-block
// time of the kernel grows with it.
//
// One CTA per SM issues ITERS groups of 8 MMAs (4 consecutive K slices into accumulator D0, 4 into D1 -- the order the GEMM
// uses) from one thread, commits, waits for completion; cycles per MMA = elapsed / (8 * ITERS).  Operands are whatever
// bits the memories hold (values do not matter for timing; NaNs do not slow the tensor pipe).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I mlx_parallm_b200/csrc -o scripts/probe_mma scripts/probe_mma.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "tc_ptx.cuh"

using namespace b200p;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

__device__ __forceinline__ void umma_f16_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accum) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accum)
      : "memory");
}

// mode 0: A in TMEM, two accumulators alternating every 4 MMAs (the GEMM's order); 1: A in TMEM, ONE accumulator;
// mode 2: A in shared memory (SS form), two accumulators
__global__ void __launch_bounds__(128, 1) probe(int N, int iters, int mode, unsigned long long* out) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar = base, slot = base + 8;
  const uint32_t b_tile = base + 1024;                 // [256 x 64] 16-bit, SWIZZLE_128B: 32 KB
  const uint32_t a_tile = b_tile + 32768;              // [128 x 64] 16-bit: 16 KB
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) { mbar_init(bar, 1); fence_barrier_init(); }
  if (warp == 0) tmem_alloc(slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(slot));
  if (threadIdx.x == 0) {
    const uint32_t idesc = (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((128u >> 4) << 24);
    const uint32_t d0 = tmem_base, d1 = tmem_base + (uint32_t)(mode == 1 ? 0 : N);
    const uint32_t a0 = tmem_base + 448;               // 64 columns of A at the top of TMEM
    const uint64_t bd = make_desc_sw128(b_tile), ad = make_desc_sw128(a_tile);
    uint32_t ph = 0;
    // warm-up group
    umma_kblock_f16_ts(d0, d1, a0, bd, idesc, 0u);
    umma_commit(bar);
    mbar_wait(bar, ph); ph ^= 1;
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      if (mode == 2) {
#pragma unroll
        for (int k4 = 0; k4 < 4; ++k4) umma_f16_ss(d0, ad + 2 * k4, bd + 2 * k4, idesc, 1u);
#pragma unroll
        for (int k4 = 0; k4 < 4; ++k4) umma_f16_ss(d1, ad + 2 * k4, bd + 2 * k4, idesc, 1u);
      } else {
        umma_kblock_f16_ts(d0, d1, a0, bd, idesc, 1u);
      }
    }
    umma_commit(bar);
    mbar_wait(bar, ph);
    const long long t1 = clock64();
    out[blockIdx.x] = (unsigned long long)(t1 - t0);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem_base, 512); }
}

int main() {
  int sms = 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  unsigned long long* out;
  CK(cudaMallocManaged(&out, sizeof(unsigned long long) * sms));
  const int smem = 1024 + 1024 + 32768 + 16384;
  CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  const int iters = 2000;
  const char* names[3] = {"A in TMEM, D0/D1 alternating", "A in TMEM, one accumulator", "A in SMEM, D0/D1 alternating"};
  for (int mode = 0; mode < 3; ++mode) {
    printf("%s\n", names[mode]);
    for (int N : {16, 32, 64, 128, 256}) {
      if (mode != 1 && N > 128) continue;              // two accumulators of N columns + the A slot must fit 512 columns
      for (int grid : {1, sms}) {
        probe<<<grid, 128, smem>>>(N, iters, mode, out);
        CK(cudaDeviceSynchronize());
        double mx = 0;
        for (int i = 0; i < grid; ++i) mx = out[i] > mx ? (double)out[i] : mx;
        const double per = mx / (8.0 * iters);
        printf("  N=%3d grid=%3d: %.1f cycles per MMA (M=128,K=16)  -> %.0f MAC/cycle/SM, %.1f cycles per 64-k block of 2x128 rows\n", N, grid,
               per, 128.0 * N * 16 / per, 8 * per);
      }
    }
  }
  return 0;
}
