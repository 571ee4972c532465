// Probe 2 (follow-up of probe_mma.cu): is the 51-cycle floor of a tcgen05.mma (M=128, K=16, N <= 64) a property of the
// tensor pipe or of ONE issuing thread?  (a) two warps of a CTA issue at the same time, each into its own accumulator from its
// own A columns; (b) M = 64 instructions (half the lanes) at N = 64..256.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I mlx_parallm_b200/csrc -o scripts/probe_mma2 scripts/probe_mma2.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "tc_ptx.cuh"

using namespace b200p;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

__device__ __forceinline__ void umma_f16_ts1(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accum) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n"
      "}\n" ::"r"(d_tmem),
      "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accum)
      : "memory");
}

// issuers = 1 or 2 warps; each issues iters * 4 MMAs (K slices 0..3 of its A slot) into its own accumulator
__global__ void __launch_bounds__(128, 1) probe(int N, int Mrows, int iters, int issuers, int commits, unsigned long long* out) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar0 = base, bar1 = base + 8, slot = base + 16, scratch = base + 64;   // scratch: 4 barriers per issuer, never waited on
  const uint32_t b_tile = base + 1024;                 // [256 x 64] 16-bit, SWIZZLE_128B: 32 KB
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) { mbar_init(bar0, 1); mbar_init(bar1, 1); for (int i = 0; i < 8; ++i) mbar_init(scratch + 8 * i, 1); fence_barrier_init(); }
  if (warp == 0) tmem_alloc(slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(slot));
  __shared__ long long tt[4];
  if ((threadIdx.x & 31) == 0 && warp < issuers) {
    const uint32_t idesc = (1u << 4) | ((uint32_t)(N >> 3) << 17) | (((uint32_t)Mrows >> 4) << 24);
    const uint32_t d = tmem_base + (uint32_t)(warp * 256);          // accumulators at columns 0 / 256 (N <= 192 with 2 issuers)
    const uint32_t a = tmem_base + 448 + (uint32_t)(warp * 32);     // 32 columns of A each
    const uint64_t bd = make_desc_sw128(b_tile);
    const uint32_t bar = warp == 0 ? bar0 : bar1;
    uint32_t ph = 0;
    for (int k4 = 0; k4 < 4; ++k4) umma_f16_ts1(d, a + 8 * k4, bd + 2 * k4, idesc, k4 ? 1u : 0u);
    umma_commit(bar);
    mbar_wait(bar, ph); ph ^= 1;
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
      for (int k4 = 0; k4 < 4; ++k4) umma_f16_ts1(d, a + 8 * k4, bd + 2 * k4, idesc, 1u);
      for (int c = 0; c < commits; ++c) umma_commit(scratch + 8 * (4 * warp + c));
    }
    umma_commit(bar);
    mbar_wait(bar, ph);
    const long long t1 = clock64();
    tt[2 * warp] = t0; tt[2 * warp + 1] = t1;
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 0) {
    long long s = tt[0], e = tt[1];
    if (issuers == 2) { s = min(s, tt[2]); e = max(e, tt[3]); }
    out[blockIdx.x] = (unsigned long long)(e - s);
  }
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem_base, 512); }
}

int main() {
  int sms = 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  unsigned long long* out;
  CK(cudaMallocManaged(&out, sizeof(unsigned long long) * sms));
  const int smem = 1024 + 1024 + 32768;
  CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  const int iters = 2000;
  for (int Mrows : {128}) {
    for (int issuers : {1, 2})
    for (int commits : {0, 1, 2, 3}) {
      printf("M=%d, %d issuing warp(s), A in TMEM, one accumulator per issuer, %d tcgen05.commit after every 4 MMAs\n", Mrows, issuers, commits);
      for (int N : {16, 64}) {
        probe<<<sms, 128, smem>>>(N, Mrows, iters, issuers, commits, out);
        CK(cudaDeviceSynchronize());
        double mx = 0;
        for (int i = 0; i < sms; ++i) mx = out[i] > mx ? (double)out[i] : mx;
        const double per = mx / (4.0 * iters * issuers);
        printf("  N=%3d: %.1f cycles per MMA (all issuers together: one MMA retired every %.1f cycles) -> %.0f MAC/cycle/SM\n", N,
               mx / (4.0 * iters), per, (double)Mrows * N * 16 / per);
      }
    }
  }
  return 0;
}
