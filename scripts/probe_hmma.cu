// Probe: throughput of legacy mma.sync.m16n8k16 (f16 in, f32 accumulate) on sm_100a -- design input for a pipeline-free GEMM
// for small weight shards at 64 rows.  W warps per SM each run ITERS x 8 independent MMAs (8 accumulators).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/probe_hmma scripts/probe_hmma.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

__global__ void probe(int iters, float* sink, unsigned long long* cyc) {
  float c[8][4];
  for (int n = 0; n < 8; ++n) for (int j = 0; j < 4; ++j) c[n][j] = 0.f;
  uint32_t a[4] = {threadIdx.x, threadIdx.x * 3u, 0x3c003c00u, 0x3c003c00u}, b0 = 0x3c003c00u, b1 = threadIdx.x;
  __syncthreads();
  const long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int n = 0; n < 8; ++n)
      asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                   : "+f"(c[n][0]), "+f"(c[n][1]), "+f"(c[n][2]), "+f"(c[n][3])
                   : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
  __syncthreads();
  const long long t1 = clock64();
  float s = 0.f;
  for (int n = 0; n < 8; ++n) for (int j = 0; j < 4; ++j) s += c[n][j];
  sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = (unsigned long long)(t1 - t0);
}

int main() {
  int sms = 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  float* sink; unsigned long long* cyc;
  CK(cudaMalloc(&sink, sizeof(float) * sms * 1024));
  CK(cudaMallocManaged(&cyc, sizeof(unsigned long long) * sms));
  const int iters = 4000;
  for (int warps : {4, 8, 16, 32}) {
    probe<<<sms, warps * 32>>>(iters, sink, cyc);
    CK(cudaDeviceSynchronize());
    double mx = 0;
    for (int i = 0; i < sms; ++i) mx = cyc[i] > mx ? (double)cyc[i] : mx;
    const double mmas = (double)iters * 8 * warps;
    printf("%2d warps per SM: %.2f cycles per mma.sync.m16n8k16 per SM -> %.0f MAC/cycle/SM\n", warps, mx / mmas, mmas * 2048.0 / mx);
  }
  return 0;
}
