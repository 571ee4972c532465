// Probe: can a kernel's weight stream be pre-staged in L2 by the PREVIOUS kernel (cp.async.bulk.prefetch.L2), so HBM
// stays busy across kernel boundaries?  Streams matrix i with 1D bulk copies while prefetching matrix i+1.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/probe_l2pf scripts/probe_l2pf.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count)); }
__device__ __forceinline__ void mbar_arrive(uint32_t bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory"); }
__device__ __forceinline__ void mbar_expect(uint32_t bar, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile("{\n.reg .pred p;\nW:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}\n" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk1d(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void prefetch_l2(const void* src, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(bytes) : "memory");
}
constexpr int CH = 32768, ST = 6;
// stream `cur` (units of CH) through smem; optionally prefetch the same offsets of `nxt` into L2 as we go
__global__ void __launch_bounds__(64, 1) stream(const uint8_t* cur, const uint8_t* nxt, long long units) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t ring = (smem_u32(smem) + 1023u) & ~1023u;
  const uint32_t bars = ring + ST * CH;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < ST; ++s) { mbar_init(bars + 8 * s, 1); mbar_init(bars + 8 * (ST + s), 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const long long u0 = units * blockIdx.x / gridDim.x, u1 = units * (blockIdx.x + 1) / gridDim.x;
  if (warp == 0) {
    if (lane == 0) {
      int si = 0, ph = 0;
      for (long long u = u0; u < u1; ++u) {
        mbar_wait(bars + 8 * (ST + si), ph ^ 1);
        mbar_expect(bars + 8 * si, CH);
        bulk1d(ring + si * CH, cur + u * CH, CH, bars + 8 * si);
        if (nxt) prefetch_l2(nxt + u * CH, CH);
        if (++si == ST) { si = 0; ph ^= 1; }
      }
    }
  } else {
    int si = 0, ph = 0;
    for (long long u = u0; u < u1; ++u) {
      mbar_wait(bars + 8 * si, ph);
      __syncwarp();
      if (lane == 0) mbar_arrive(bars + 8 * (ST + si));
      if (++si == ST) { si = 0; ph ^= 1; }
    }
  }
}
// prefetch-only kernel: a few CTAs walk the matrix
__global__ void prefetch_only(const uint8_t* nxt, long long units) {
  const long long u0 = units * blockIdx.x / gridDim.x, u1 = units * (blockIdx.x + 1) / gridDim.x;
  if (threadIdx.x == 0) for (long long u = u0; u < u1; ++u) prefetch_l2(nxt + u * CH, CH);
}
__global__ void spin(long long cycles) { long long t0 = clock64(); while (clock64() - t0 < cycles) {} }

int main() {
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  CK(cudaFuncSetAttribute(stream, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  const int smem = 1024 + ST * CH + 256;
  for (long long mb : {14, 29, 59, 88}) {
    const long long bytes = mb * 1024 * 1024 / CH * CH, units = bytes / CH;
    const int NM = (int)(640 / mb) + 2;
    uint8_t* w; CK(cudaMalloc(&w, bytes * NM)); CK(cudaMemset(w, 1, bytes * NM));
    float ms;
    // a. HBM-fed, back to back
    for (int rep = 0; rep < 2; ++rep) {
      CK(cudaEventRecord(e0));
      for (int m = 0; m < NM; ++m) stream<<<148, 64, smem>>>(w + m * bytes, nullptr, units);
      CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaEventElapsedTime(&ms, e0, e1));
    }
    printf("%3lld MB  a. HBM-fed back-to-back            : %6.1f us/matrix = %5.0f GB/s\n", mb, ms * 1e3 / NM, bytes * NM / ms / 1e6);
    // b. prefetch kernel, 40 us pause, then stream (timed alone)
    float tb = 0;
    for (int m = 0; m < NM; ++m) {
      prefetch_only<<<32, 32>>>(w + m * bytes, units);
      spin<<<1, 1>>>(80000);
      CK(cudaEventRecord(e0));
      stream<<<148, 64, smem>>>(w + m * bytes, nullptr, units);
      CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaEventElapsedTime(&ms, e0, e1));
      tb += ms;
    }
    printf("%3lld MB  b. L2-fed (prefetched, then timed) : %6.1f us/matrix = %5.0f GB/s\n", mb, tb * 1e3 / NM, bytes * NM / tb / 1e6);
    // c. pipeline: stream m while prefetching m+1 from inside the same kernel
    for (int rep = 0; rep < 2; ++rep) {
      prefetch_only<<<32, 32>>>(w, units);
      spin<<<1, 1>>>(80000);
      CK(cudaEventRecord(e0));
      for (int m = 0; m < NM; ++m) stream<<<148, 64, smem>>>(w + m * bytes, m + 1 < NM ? w + (m + 1) * bytes : nullptr, units);
      CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaEventElapsedTime(&ms, e0, e1));
    }
    printf("%3lld MB  c. stream m + in-kernel prefetch m+1: %6.1f us/matrix = %5.0f GB/s\n", mb, ms * 1e3 / NM, bytes * NM / ms / 1e6);
    // d. pipeline with a separate concurrent prefetch kernel on a second stream
    cudaStream_t s2; CK(cudaStreamCreateWithFlags(&s2, cudaStreamNonBlocking));
    cudaEvent_t ev[64]; for (int i = 0; i < 64; ++i) CK(cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming));
    for (int rep = 0; rep < 2; ++rep) {
      prefetch_only<<<32, 32>>>(w, units);
      spin<<<1, 1>>>(80000);
      CK(cudaEventRecord(e0));
      for (int m = 0; m < NM; ++m) {
        CK(cudaEventRecord(ev[m % 64], 0));
        CK(cudaStreamWaitEvent(s2, ev[m % 64], 0));
        if (m + 1 < NM) prefetch_only<<<16, 32, 0, s2>>>(w + (m + 1) * bytes, units);
        stream<<<148, 64, smem>>>(w + m * bytes, nullptr, units);
      }
      CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaEventElapsedTime(&ms, e0, e1));
      CK(cudaDeviceSynchronize());
    }
    printf("%3lld MB  d. stream m || side-stream prefetch : %6.1f us/matrix = %5.0f GB/s\n", mb, ms * 1e3 / NM, bytes * NM / ms / 1e6);
    fflush(stdout);
    CK(cudaFree(w));
  }
  return 0;
}
