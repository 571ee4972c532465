// Generic fused-dequant GEMV on CUDA cores: y[m, n] = sum_k x[m,k] * (q[n,k]*s[n,g] + b[n,g]).
//
// This is the any-dtype / any-shape path of b200p_qmm (float32 models, odd shapes, and M below the
// tensor-path threshold).  One warp owns two output rows and streams their packed weights with
// 128-bit no-allocate loads; the 32 lanes split K; fp32 accumulation; warp-shuffle reduction.
// Algorithmic bytes per launch: N*K*bits/8 + 2*N*(K/gs)*sizeof(T)  (+ M*K*sizeof(T) of x, L1/L2-resident).
#include "common.cuh"

namespace b200p {

template <typename T> struct XVec;  // 8 consecutive activations as floats
template <> struct XVec<float> {
  static __device__ __forceinline__ void load(const float* p, float (&v)[8]) {
    float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
};
template <> struct XVec<__half> {
  static __device__ __forceinline__ void load(const __half* p, float (&v)[8]) {
    uint4 u = *reinterpret_cast<const uint4*>(p);
    const __half2* h = reinterpret_cast<const __half2*>(&u);
#pragma unroll
    for (int i = 0; i < 4; ++i) { float2 f = __half22float2(h[i]); v[2 * i] = f.x; v[2 * i + 1] = f.y; }
  }
};
template <> struct XVec<__nv_bfloat16> {
  static __device__ __forceinline__ void load(const __nv_bfloat16* p, float (&v)[8]) {
    uint4 u = *reinterpret_cast<const uint4*>(p);
    const uint32_t* w = reinterpret_cast<const uint32_t*>(&u);
#pragma unroll
    for (int i = 0; i < 4; ++i) { v[2 * i] = __uint_as_float(w[i] << 16); v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u); }
  }
};

__device__ __forceinline__ float silu_f(float g) { return g / (1.0f + __expf(-g)); }

// grid.x = ceil(pairs / 4) with pairs = N/2 ; block = 128 (4 warps, one row pair each)
template <typename T, int BITS, int MB, int EPI>
__global__ void __launch_bounds__(128)
qmv_kernel(const T* __restrict__ x, const uint32_t* __restrict__ w, const T* __restrict__ scales,
           const T* __restrict__ biases, void* __restrict__ yv, int m_valid, int64_t N, int64_t K,
           int64_t ldx, int64_t ldy, int gs, int flags) {
  constexpr int WPL = (32 / BITS) * 4;  // weights per lane per 128-bit load
  const int lane = threadIdx.x & 31;
  const int64_t pair = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
  const int64_t half_n = N >> 1;
  if (pair >= half_n) return;
  // SILU_MUL pairs gate row r with up row r + N/2 so the epilogue is warp-local
  const int64_t r0 = (EPI == B200P_EPI_SILU_MUL) ? pair : 2 * pair;
  const int64_t r1 = (EPI == B200P_EPI_SILU_MUL) ? pair + half_n : 2 * pair + 1;
  const int64_t wrow = K * BITS / 32;  // uint32 words per row
  const int64_t ng = K / gs;
  const uint32_t* w0 = w + r0 * wrow;
  const uint32_t* w1 = w + r1 * wrow;

  float acc0[MB], acc1[MB];
#pragma unroll
  for (int m = 0; m < MB; ++m) acc0[m] = acc1[m] = 0.f;

  pdl_wait();
  for (int64_t k = (int64_t)lane * WPL; k < K; k += 32 * WPL) {
    uint4 q0 = ldg_stream(w0 + k * BITS / 32);
    uint4 q1 = ldg_stream(w1 + k * BITS / 32);
    const int64_t g = k / gs;
    const float s0 = Tr<T>::to_f(scales[r0 * ng + g]), b0 = Tr<T>::to_f(biases[r0 * ng + g]);
    const float s1 = Tr<T>::to_f(scales[r1 * ng + g]), b1 = Tr<T>::to_f(biases[r1 * ng + g]);
    const uint32_t* p0 = reinterpret_cast<const uint32_t*>(&q0);
    const uint32_t* p1 = reinterpret_cast<const uint32_t*>(&q1);
#pragma unroll
    for (int c = 0; c < WPL / 8; ++c) {  // chunks of 8 weights
      float wa[8], wb[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        uint32_t qa, qb;
        if (BITS == 4) {
          qa = (p0[c] >> (4 * j)) & 0xfu;
          qb = (p1[c] >> (4 * j)) & 0xfu;
        } else {
          qa = (p0[2 * c + (j >> 2)] >> (8 * (j & 3))) & 0xffu;
          qb = (p1[2 * c + (j >> 2)] >> (8 * (j & 3))) & 0xffu;
        }
        wa[j] = fmaf(u8_to_float(qa), s0, b0);
        wb[j] = fmaf(u8_to_float(qb), s1, b1);
      }
#pragma unroll
      for (int m = 0; m < MB; ++m) {
        if (m < m_valid) {
          float xv[8];
          XVec<T>::load(x + (int64_t)m * ldx + k + 8 * c, xv);
          if (flags & B200P_X_PERM8) {   // stored order [0,4,1,5,2,6,3,7] -> natural
            const float t1 = xv[1], t2 = xv[2], t3 = xv[3], t4 = xv[4], t5 = xv[5], t6 = xv[6];
            xv[4] = t1; xv[1] = t2; xv[5] = t3; xv[2] = t4; xv[6] = t5; xv[3] = t6;
          }
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            acc0[m] = fmaf(xv[j], wa[j], acc0[m]);
            acc1[m] = fmaf(xv[j], wb[j], acc1[m]);
          }
        }
      }
    }
  }
  pdl_launch_dependents();
#pragma unroll
  for (int m = 0; m < MB; ++m) {
    acc0[m] = warp_sum(acc0[m]);
    acc1[m] = warp_sum(acc1[m]);
  }
  if (lane == 0) {
    const bool yp = (flags & B200P_Y_PERM8) != 0;   // 16-bit outputs in the PERM8 column order
    const int64_t c0 = yp ? ((r0 & ~7ll) | (((r0 & 3) << 1) | ((r0 >> 2) & 1))) : r0;
    const int64_t c1 = yp ? ((r1 & ~7ll) | (((r1 & 3) << 1) | ((r1 >> 2) & 1))) : r1;
#pragma unroll
    for (int m = 0; m < MB; ++m) {
      if (m >= m_valid) break;
      if (EPI == B200P_EPI_STORE) {
        T* y = reinterpret_cast<T*>(yv) + (int64_t)m * ldy;
        y[c0] = Tr<T>::from_f(acc0[m]);
        y[c1] = Tr<T>::from_f(acc1[m]);
      } else if (EPI == B200P_EPI_STORE_F32) {
        float* y = reinterpret_cast<float*>(yv) + (int64_t)m * ldy;
        y[r0] = acc0[m];
        y[r1] = acc1[m];
      } else if (EPI == B200P_EPI_ADD_RESIDUAL) {
        float* y = reinterpret_cast<float*>(yv) + (int64_t)m * ldy;
        y[r0] += acc0[m];
        y[r1] += acc1[m];
      } else {
        T* y = reinterpret_cast<T*>(yv) + (int64_t)m * ldy;
        y[c0] = Tr<T>::from_f(silu_f(acc0[m]) * acc1[m]);
      }
    }
  }
}

template <typename T, int BITS, int EPI>
static int launch_qmv_mb(const T* x, const uint32_t* w, const T* s, const T* b, void* y, int64_t M, int64_t N,
                         int64_t K, int64_t ldx, int64_t ldy, int gs, int flags, cudaStream_t st) {
  const size_t ysz = (EPI == B200P_EPI_STORE_F32 || EPI == B200P_EPI_ADD_RESIDUAL) ? sizeof(float) : sizeof(T);
  dim3 grid((unsigned)((N / 2 + 3) / 4)), block(128);
  for (int64_t m0 = 0; m0 < M; m0 += 4) {
    int mv = (int)((M - m0) < 4 ? (M - m0) : 4);
    const T* xp = x + m0 * ldx;
    void* yp = reinterpret_cast<char*>(y) + m0 * ldy * ysz;
    if (mv == 1) launch_pdl(qmv_kernel<T, BITS, 1, EPI>, grid, block, 0, st, xp, w, s, b, yp, mv, N, K, ldx, ldy, gs, flags);
    else if (mv == 2) launch_pdl(qmv_kernel<T, BITS, 2, EPI>, grid, block, 0, st, xp, w, s, b, yp, mv, N, K, ldx, ldy, gs, flags);
    else launch_pdl(qmv_kernel<T, BITS, 4, EPI>, grid, block, 0, st, xp, w, s, b, yp, mv, N, K, ldx, ldy, gs, flags);
  }
  B200P_LAUNCH_CHECK("qmv_kernel");
  return B200P_OK;
}

template <typename T>
static int launch_qmv_t(const void* x, const uint32_t* w, const void* s, const void* b, void* y, int64_t M,
                        int64_t N, int64_t K, int64_t ldx, int64_t ldy, int gs, int bits, int epi, int flags, cudaStream_t st) {
#define GO(B_, E_) return launch_qmv_mb<T, B_, E_>((const T*)x, w, (const T*)s, (const T*)b, y, M, N, K, ldx, ldy, gs, flags, st)
  if (bits == 4) {
    switch (epi) { case 0: GO(4, 0); case 1: GO(4, 1); case 2: GO(4, 2); case 3: GO(4, 3); }
  } else {
    switch (epi) { case 0: GO(8, 0); case 1: GO(8, 1); case 2: GO(8, 2); case 3: GO(8, 3); }
  }
#undef GO
  set_error("bad epilogue %d", epi);
  return B200P_ERR_ARG;
}

int qmv_launch(const void* x, const uint32_t* w, const void* s, const void* b, void* y, int64_t M, int64_t N,
               int64_t K, int64_t ldx, int64_t ldy, int gs, int bits, int dtype, int epi, int flags, cudaStream_t st) {
  B200P_DISPATCH_DTYPE(dtype, T, return (launch_qmv_t<T>(x, w, s, b, y, M, N, K, ldx, ldy, gs, bits, epi, flags, st)));
  return B200P_OK;
}

}  // namespace b200p
