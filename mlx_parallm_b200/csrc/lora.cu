// LoRA at inference (row f4): the low-rank update of mlx_lm's LoRALinear,
//
//     y[m, col0 + n] += scale * sum_r ( sum_k x[m,k] * A[k,r] ) * B[r,n]          (mlx_lm tuner LoRALinear.__call__;
//                                                                                 adapters on q_proj / v_proj of the last
//                                                                                 layers: lora_init.py:63-73, A.9)
//
// applied on top of the quantised projection's output (the fp32 accumulation buffer of the fused q|k|v projection in the
// decode step, or its 16-bit output in prefill).  Rank r <= 16 (padded with zeros), both small GEMMs on mma.sync m16n8k16 with
// fp32 accumulation: a CTA takes 16 token rows and a slice of 512 output columns; its 8 warps split K for t = x A (partials
// summed through shared memory), then each warp applies t B to its n-tiles -- one MMA per 8 columns, because the rank IS the
// MMA's k.  Operands are stored transposed by the host (At [16, K], Bt [N, 16]) so that every fragment is a contiguous
// 32-bit load; with PERM8 activations the host permutes At's k axis the same way.
// Bytes per launch: M*K*2 (x, L2) + 16*K*2 + N*16*2 (adapters) + M*N*4 -- latency-bound at decode sizes.
#include "common.cuh"
#include "mma_util.cuh"

namespace b200p {

constexpr int LR_WARPS = 8;
constexpr int LR_THREADS = LR_WARPS * 32;
constexpr int LR_R = 16;                        // padded rank
constexpr int LR_NSLICE = 512;                  // output columns per CTA

template <typename T, bool OUT_F32>
__global__ void __launch_bounds__(LR_THREADS)
lora_add_kernel(const T* __restrict__ x, const T* __restrict__ At, const T* __restrict__ Bt, void* __restrict__ yv, int M,
                int64_t N, int64_t K, int64_t ldx, int64_t ldy, int64_t col0, float scale) {
  __shared__ float part[LR_WARPS][16][LR_R + 1];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int m0 = blockIdx.y * 16;
  const int64_t n_begin = (int64_t)blockIdx.x * LR_NSLICE;
  pdl_wait();
  // ---- t = x[m0:m0+16, :] @ A  : this warp's K slice ------------------------------------------------
  const int r0 = m0 + g, r1 = m0 + g + 8;
  const uint32_t* x0 = reinterpret_cast<const uint32_t*>(x + (int64_t)(r0 < M ? r0 : 0) * ldx);
  const uint32_t* x1 = reinterpret_cast<const uint32_t*>(x + (int64_t)(r1 < M ? r1 : 0) * ldx);
  const uint32_t* a0 = reinterpret_cast<const uint32_t*>(At + (int64_t)g * K);          // rank g     (n-tile 0)
  const uint32_t* a1 = reinterpret_cast<const uint32_t*>(At + (int64_t)(g + 8) * K);    // rank g + 8 (n-tile 1)
  float c0[4] = {0.f, 0.f, 0.f, 0.f}, c1[4] = {0.f, 0.f, 0.f, 0.f};
  const int64_t ksteps = K / 16;
  for (int64_t ks = warp; ks < ksteps; ks += LR_WARPS) {
    const int64_t o = ks * 8 + t;                 // uint32 index of (k = 16 ks + 2t, +1)
    uint32_t af[4];
    af[0] = r0 < M ? x0[o] : 0u;
    af[1] = r1 < M ? x1[o] : 0u;
    af[2] = r0 < M ? x0[o + 4] : 0u;
    af[3] = r1 < M ? x1[o + 4] : 0u;
    mma16816<T>(c0, af, a0[o], a0[o + 4]);
    mma16816<T>(c1, af, a1[o], a1[o + 4]);
  }
  // c0: ranks 2t, 2t+1 (rows g, g+8); c1: ranks 8 + 2t, 8 + 2t + 1
  part[warp][g][2 * t] = c0[0];      part[warp][g][2 * t + 1] = c0[1];
  part[warp][g + 8][2 * t] = c0[2];  part[warp][g + 8][2 * t + 1] = c0[3];
  part[warp][g][8 + 2 * t] = c1[0];      part[warp][g][8 + 2 * t + 1] = c1[1];
  part[warp][g + 8][8 + 2 * t] = c1[2];  part[warp][g + 8][8 + 2 * t + 1] = c1[3];
  __syncthreads();
  // every lane builds its A fragment of t (16 rows x 16 ranks), scaled, in the storage type
  auto tsum = [&](int row, int r) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < LR_WARPS; ++w) s += part[w][row][r];
    return s * scale;
  };
  uint32_t tf[4];
  tf[0] = pack2<T>(tsum(g, 2 * t), tsum(g, 2 * t + 1));
  tf[1] = pack2<T>(tsum(g + 8, 2 * t), tsum(g + 8, 2 * t + 1));
  tf[2] = pack2<T>(tsum(g, 2 * t + 8), tsum(g, 2 * t + 9));
  tf[3] = pack2<T>(tsum(g + 8, 2 * t + 8), tsum(g + 8, 2 * t + 9));
  pdl_launch_dependents();
  // ---- y[:, n-tiles of this warp] += t @ B ------------------------------------------------------------
  const int64_t n_end = min(N, n_begin + LR_NSLICE);
  for (int64_t n0 = n_begin + warp * 8; n0 < n_end; n0 += LR_WARPS * 8) {
    const int64_t n = n0 + g;                                      // B-fragment column of this lane
    const uint32_t* bp = reinterpret_cast<const uint32_t*>(Bt + (n < N ? n : 0) * LR_R);
    const uint32_t b0 = n < N ? bp[t] : 0u, b1 = n < N ? bp[t + 4] : 0u;
    float d[4] = {0.f, 0.f, 0.f, 0.f};
    mma16816<T>(d, tf, b0, b1);
    // d[0], d[1] = (row g, cols n0 + 2t, +1); d[2], d[3] = (row g + 8, same cols)
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int64_t col = n0 + 2 * t + j;
      if (col >= N) continue;
      if (OUT_F32) {
        float* y = reinterpret_cast<float*>(yv);
        if (r0 < M) y[(int64_t)r0 * ldy + col0 + col] += d[j];
        if (r1 < M) y[(int64_t)r1 * ldy + col0 + col] += d[2 + j];
      } else {
        T* y = reinterpret_cast<T*>(yv);
        // (scale * z).astype(x.dtype) added to the projection's output, rounded to the storage type (A.9)
        if (r0 < M) { T* p = y + (int64_t)r0 * ldy + col0 + col; *p = Tr<T>::from_f(Tr<T>::to_f(*p) + Tr<T>::to_f(Tr<T>::from_f(d[j]))); }
        if (r1 < M) { T* p = y + (int64_t)r1 * ldy + col0 + col; *p = Tr<T>::from_f(Tr<T>::to_f(*p) + Tr<T>::to_f(Tr<T>::from_f(d[2 + j]))); }
      }
    }
  }
}

}  // namespace b200p

using namespace b200p;

extern "C" int b200p_lora_add(const void* x, const void* At, const void* Bt, void* y, int64_t M, int64_t N, int64_t K,
                              int64_t ldx, int64_t ldy, int64_t col0, float scale, int dtype, int out_f32, void* stream) {
  B200P_REQUIRE(x && At && Bt && y, B200P_ERR_ARG, "null pointer");
  B200P_REQUIRE(dtype == B200P_F16 || dtype == B200P_BF16, B200P_ERR_ARG, "LoRA update needs 16-bit activations / adapters");
  B200P_REQUIRE(M >= 0 && N > 0 && K > 0 && K % 16 == 0 && col0 >= 0 && ldy >= col0 + N && ldx >= K, B200P_ERR_ARG, "bad shape");
  B200P_REQUIRE(aligned(x, 4) && aligned(At, 4) && aligned(Bt, 4) && ldx % 2 == 0, B200P_ERR_ALIGN, "operands must be 4-byte aligned");
  if (M == 0) return B200P_OK;
  PdlClass pdl_cls(64);
  dim3 grid((unsigned)((N + LR_NSLICE - 1) / LR_NSLICE), (unsigned)((M + 15) / 16));
  cudaStream_t st = (cudaStream_t)stream;
#define LGO(T, F)                                                                                                   \
  launch_pdl(lora_add_kernel<T, F>, grid, dim3(LR_THREADS), 0, st, (const T*)x, (const T*)At, (const T*)Bt, y, (int)M, N, K, \
             ldx, ldy, col0, scale)
  if (dtype == B200P_F16) { if (out_f32) LGO(__half, true); else LGO(__half, false); }
  else { if (out_f32) LGO(__nv_bfloat16, true); else LGO(__nv_bfloat16, false); }
#undef LGO
  B200P_LAUNCH_CHECK("lora_add_kernel");
  return B200P_OK;
}
