// mx.quantize(w, group_size, bits): dense 16/32-bit weights -> MLX affine codes + per-group scales / biases
// (nn.quantize at utils.py:886-905, the converter at utils.py:934-981).  Row f3 of the scope table.
//
// HBM-bound byte work: N*K*sizeof(T) in, N*K*bits/8 + 2*N*(K/gs)*sizeof(T) out, one pass.  One thread owns the codes of
// one output word (8 at 4 bits, 4 at 8 bits: one 16- or 8-byte load); the gs/8 (gs/4) threads of a group reduce min / max
// with xor-shuffles.  Every arithmetic step is rounded to the storage type exactly where the oracle rounds it
// (oracle/mlx_ops.py: quantize), so codes, scales and biases are BIT-EXACT against it.
#include "common.cuh"

namespace b200p {

template <typename T> __device__ __forceinline__ float rt(float v) { return Tr<T>::to_f(Tr<T>::from_f(v)); }

template <typename T, int BITS>
__global__ void __launch_bounds__(256)
quantize_kernel(const T* __restrict__ w, uint32_t* __restrict__ wq, T* __restrict__ scales, T* __restrict__ biases, int64_t N,
                int64_t K, int gs) {
  constexpr int PER = 32 / BITS;                       // codes per word = elements per thread
  const int tpg = gs / PER;                            // threads per group: 4..32, a power of two
  const int64_t words_per_row = K / PER;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = idx < N * words_per_row;           // dead threads still take part in the shuffles
  pdl_wait();
  const int64_t row = live ? idx / words_per_row : 0, wi = live ? idx - row * words_per_row : 0;
  float v[PER];
  if (live) {
    const T* src = w + row * K + wi * PER;
    if (sizeof(T) * PER == 16) {
      const uint4 u = *reinterpret_cast<const uint4*>(src);
      const T* p = reinterpret_cast<const T*>(&u);
#pragma unroll
      for (int j = 0; j < PER; ++j) v[j] = Tr<T>::to_f(p[j]);
    } else {
#pragma unroll
      for (int j = 0; j < PER; ++j) v[j] = Tr<T>::to_f(src[j]);
    }
  } else {
#pragma unroll
    for (int j = 0; j < PER; ++j) v[j] = 0.f;
  }
  float mx = v[0], mn = v[0];
#pragma unroll
  for (int j = 1; j < PER; ++j) { mx = fmaxf(mx, v[j]); mn = fminf(mn, v[j]); }
  for (int o = 1; o < tpg; o <<= 1) {
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
  }
  const float n_bins = (float)((1 << BITS) - 1);
  const bool mask = fabsf(mn) > fabsf(mx);
  float s = fmaxf(rt<T>(__fdiv_rn(mx - mn, n_bins)), 1e-7f);
  s = mask ? s : -s;
  const float edge = mask ? mn : mx;
  const float q0 = rintf(rt<T>(__fdiv_rn(edge, s)));
  if (q0 != 0.f) s = rt<T>(__fdiv_rn(edge, q0));
  const float bias = (q0 == 0.f) ? 0.f : edge;
  s = rt<T>(s);
  const float b = rt<T>(bias);
  uint32_t word = 0;
#pragma unroll
  for (int j = 0; j < PER; ++j) {
    float q = rintf(rt<T>(__fdiv_rn(v[j] - b, s)));
    q = fminf(fmaxf(q, 0.f), n_bins);
    word |= ((uint32_t)q) << (j * BITS);
  }
  if (live) {
    wq[idx] = word;
    if ((wi % tpg) == 0) {
      const int64_t g = row * (K / gs) + (wi * PER) / gs;
      scales[g] = Tr<T>::from_f(s);
      biases[g] = Tr<T>::from_f(b);
    }
  }
}

}  // namespace b200p

using namespace b200p;

extern "C" int b200p_quantize(const void* w, uint32_t* wq, void* scales, void* biases, int64_t N, int64_t K,
                              int group_size, int bits, int dtype, void* stream) {
  B200P_REQUIRE(N >= 0 && K > 0, B200P_ERR_ARG, "bad shape");
  if (N == 0) return B200P_OK;
  B200P_REQUIRE(w && wq && scales && biases, B200P_ERR_ARG, "null pointer");
  B200P_REQUIRE(bits == 4 || bits == 8, B200P_ERR_ARG, "bits must be 4 or 8, got %d", bits);
  B200P_REQUIRE((group_size == 32 || group_size == 64 || group_size == 128) && K % group_size == 0, B200P_ERR_ARG,
                "group_size must be 32, 64 or 128 and divide K=%lld", (long long)K);
  B200P_REQUIRE(aligned(w, 16) && aligned(wq, 4), B200P_ERR_ALIGN, "w must be 16-byte aligned");
  const int per = 32 / bits;
  const int64_t total = N * (K / per);
  dim3 grid((unsigned)((total + 255) / 256)), block(256);
  cudaStream_t st = (cudaStream_t)stream;
#define QGO(T)                                                                                                        \
  do {                                                                                                                \
    if (bits == 4) launch_pdl(quantize_kernel<T, 4>, grid, block, 0, st, (const T*)w, wq, (T*)scales, (T*)biases, N, K, group_size); \
    else launch_pdl(quantize_kernel<T, 8>, grid, block, 0, st, (const T*)w, wq, (T*)scales, (T*)biases, N, K, group_size);          \
  } while (0)
  B200P_DISPATCH_DTYPE(dtype, T, QGO(T));
#undef QGO
  B200P_LAUNCH_CHECK("quantize_kernel");
  return B200P_OK;
}
