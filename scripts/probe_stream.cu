// Probe: how fast can 148 persistent CTAs stream a 4-bit weight matrix from HBM into shared memory, as a function
// of the access pattern?  (Design input for the qmm tensor path: MLX row-major layout vs a tile-contiguous repack.)
//   mode 0  TMA 2D boxes [128 rows x 128 B] (SWIZZLE_128B) out of a row-major [N x K/2] byte matrix (pitch K/2)
//   mode 1  1D bulk copies of contiguous CHUNK-byte pieces
//   mode 2  mode 1 + `xbytes` of L2-resident extra reads per chunk (the activation tile re-read)
//   mode 3  mode 0 + the same extra reads
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/probe_stream scripts/probe_stream.cu -lcuda
#include <cuda.h>
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
__device__ __forceinline__ void tma2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(bar) : "memory");
}
__device__ __forceinline__ void bulk1d(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

struct P {
  const uint8_t* w;      // matrix base
  const uint8_t* x;      // small L2-resident buffer (>= 1 MB)
  long long units;       // number of stage-sized units
  int stage_bytes;       // weight bytes per stage (32 KB: two 128x128 boxes, or one contiguous chunk)
  int xbytes;            // extra L2 bytes per stage
  int stages;
  int mode;
  int num_st;            // stages along K per tile (mode 0/3)
};

__global__ void __launch_bounds__(64, 1) probe(const __grid_constant__ CUtensorMap tm, const P p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t ring = (smem_u32(smem) + 1023u) & ~1023u;
  const int per = p.stage_bytes + p.xbytes;
  const uint32_t bars = ring + p.stages * per;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) { mbar_init(bars + 8 * s, 1); mbar_init(bars + 8 * (p.stages + s), 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const long long u0 = p.units * blockIdx.x / gridDim.x, u1 = p.units * (blockIdx.x + 1) / gridDim.x;
  if (warp == 0) {
    if (lane == 0) {
      int si = 0, ph = 0;
      for (long long u = u0; u < u1; ++u) {
        mbar_wait(bars + 8 * (p.stages + si), ph ^ 1);
        const uint32_t dst = ring + si * per;
        mbar_expect(bars + 8 * si, (uint32_t)per);
        if (p.mode == 0 || p.mode == 3) {
          const int tile = (int)(u / p.num_st), st = (int)(u % p.num_st);
          tma2d(dst, &tm, st * 128, tile * 256, bars + 8 * si);
          tma2d(dst + 16384, &tm, st * 128, tile * 256 + 128, bars + 8 * si);
        } else {
          bulk1d(dst, p.w + u * (long long)p.stage_bytes, (uint32_t)p.stage_bytes, bars + 8 * si);
        }
        if (p.xbytes) bulk1d(dst + p.stage_bytes, p.x + ((u * 7) % 16) * p.xbytes, (uint32_t)p.xbytes, bars + 8 * si);
        if (++si == p.stages) { si = 0; ph ^= 1; }
      }
    }
  } else {
    int si = 0, ph = 0;
    for (long long u = u0; u < u1; ++u) {
      mbar_wait(bars + 8 * si, ph);
      __syncwarp();
      if (lane == 0) mbar_arrive(bars + 8 * (p.stages + si));
      if (++si == p.stages) { si = 0; ph ^= 1; }
    }
  }
}

int main(int argc, char** argv) {
  const long long N = 28672, K = 4096, pitch = K / 2, bytes = N * pitch;
  const int NM = 4;
  uint8_t* w;
  CK(cudaMalloc(&w, bytes * NM));
  CK(cudaMemset(w, 1, bytes * NM));
  uint8_t* x;
  CK(cudaMalloc(&x, 4 << 20));
  CK(cudaMemset(x, 2, 4 << 20));
  CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  struct Case { int mode, stage_bytes, xbytes, stages; const char* name; };
  Case cases[] = {
      {0, 32768, 0, 3, "tma2d 2x[128x128B] pitch 2048, 3 stages"},
      {0, 32768, 0, 6, "tma2d 2x[128x128B] pitch 2048, 6 stages"},
      {1, 32768, 0, 3, "bulk1d 32KB contiguous, 3 stages"},
      {1, 32768, 0, 6, "bulk1d 32KB contiguous, 6 stages"},
      {1, 16384, 0, 12, "bulk1d 16KB contiguous, 12 stages"},
      {1, 65536, 0, 3, "bulk1d 64KB contiguous, 3 stages"},
      {2, 32768, 8192, 5, "bulk1d 32KB + 8KB L2 reads, 5 stages"},
      {2, 32768, 16384, 4, "bulk1d 32KB + 16KB L2 reads, 4 stages"},
      {2, 32768, 32768, 3, "bulk1d 32KB + 32KB L2 reads, 3 stages"},
      {3, 32768, 32768, 3, "tma2d + 32KB L2 reads, 3 stages"},
      {3, 32768, 8192, 5, "tma2d + 8KB L2 reads, 5 stages"},
  };
  for (auto& c : cases) {
    for (int grid : {148, 296}) {
      if (grid == 296 && (c.stages * (c.stage_bytes + c.xbytes) + 2048) * 2 > 227 * 1024) continue;
      float best = 1e9f;
      for (int rep = 0; rep < 3; ++rep) {
        CK(cudaEventRecord(e0));
        for (int m = 0; m < NM; ++m) {
          CUtensorMap tm;
          cuuint64_t dims[2] = {(cuuint64_t)pitch, (cuuint64_t)N};
          cuuint64_t strides[1] = {(cuuint64_t)pitch};
          cuuint32_t box[2] = {128u, 128u}, estr[2] = {1u, 1u};
          CUresult r = cuTensorMapEncodeTiled(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, w + m * bytes, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                              CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
          if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 1; }
          P p;
          p.w = w + m * bytes; p.x = x; p.stage_bytes = c.stage_bytes; p.xbytes = c.xbytes; p.stages = c.stages; p.mode = c.mode;
          p.units = bytes / c.stage_bytes; p.num_st = (int)(pitch / 128);
          const int smem = 1024 + c.stages * (c.stage_bytes + c.xbytes) + 256;
          probe<<<grid, 64, smem>>>(tm, p);
        }
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
      }
      CK(cudaGetLastError());
      const double us = best * 1e3 / NM;
      printf("%-48s grid %3d: %7.1f us per 58.7 MB matrix  = %6.0f GB/s HBM (+%5.0f GB/s L2 extra)\n", c.name, grid, us, bytes / us / 1e3,
             (double)bytes / c.stage_bytes * c.xbytes / us / 1e3);
      fflush(stdout);
    }
  }
  return 0;
}
