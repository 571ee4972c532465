// Generic cached attention on CUDA cores (any storage dtype, head_dim <= 256, any GQA ratio,
// L >= 1 query tokens per row with the causal bound keys < lengths[b] + l + 1).
//
// Used for float32 models, unusual head dims and as the chunked-prefill attention; the split-K
// tensor kernel in attn_decode.cu is the hot decode path for 16-bit D=128 models.
// One CTA (4 warps) per (b, l, q-head); warps interleave keys; lanes split head_dim; online softmax
// per warp, merged through shared memory.
#include "common.cuh"

namespace b200p {

template <typename T, int EPL>
__device__ __forceinline__ void load_vec(const T* p, float (&v)[EPL], bool act) {
  if (!act) {
#pragma unroll
    for (int j = 0; j < EPL; ++j) v[j] = 0.f;
    return;
  }
#pragma unroll
  for (int j = 0; j < EPL; ++j) v[j] = Tr<T>::to_f(p[j]);
}

template <typename T, int EPL>
__global__ void __launch_bounds__(128)
attn_generic_kernel(const T* __restrict__ q, const T* __restrict__ kc, const T* __restrict__ vc,
                    const int32_t* __restrict__ lengths, T* __restrict__ out, int L, int Hq, int Hkv, int D,
                    int64_t Tcap, float scale, int out_perm) {
  __shared__ float sh_m[4], sh_l[4];
  __shared__ float sh_acc[4][32 * EPL];
  pdl_wait();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int h = blockIdx.x % Hq;
  const int tok = blockIdx.x / Hq;  // b*L + l
  const int b = tok / L, l = tok - b * L;
  const int hk = h / (Hq / Hkv);
  int klen = lengths[b] + l + 1;
  if (klen > Tcap) klen = (int)Tcap;
  const int d0 = lane * EPL;
  const bool act = d0 < D;
  float qv[EPL];
  load_vec<T, EPL>(q + ((int64_t)tok * Hq + h) * D + d0, qv, act);
#pragma unroll
  for (int j = 0; j < EPL; ++j) qv[j] *= scale;
  const T* kb = kc + ((int64_t)b * Hkv + hk) * Tcap * D + d0;
  const T* vb = vc + ((int64_t)b * Hkv + hk) * Tcap * D + d0;
  float m = -INFINITY, lsum = 0.f, acc[EPL];
#pragma unroll
  for (int j = 0; j < EPL; ++j) acc[j] = 0.f;
  for (int t = warp; t < klen; t += 4) {
    float kv[EPL], vv[EPL];
    load_vec<T, EPL>(kb + (int64_t)t * D, kv, act);
    load_vec<T, EPL>(vb + (int64_t)t * D, vv, act);
    float dot = 0.f;
#pragma unroll
    for (int j = 0; j < EPL; ++j) dot = fmaf(qv[j], kv[j], dot);
    dot = warp_sum(dot);
    float mn = fmaxf(m, dot);
    float corr = __expf(m - mn);  // exp(-inf) = 0 on the first key
    float p = __expf(dot - mn);
    lsum = lsum * corr + p;
#pragma unroll
    for (int j = 0; j < EPL; ++j) acc[j] = acc[j] * corr + p * vv[j];
    m = mn;
  }
  if (lane == 0) { sh_m[warp] = m; sh_l[warp] = lsum; }
#pragma unroll
  for (int j = 0; j < EPL; ++j) sh_acc[warp][d0 + j] = acc[j];
  __syncthreads();
  pdl_launch_dependents();
  if (warp == 0) {
    float M = fmaxf(fmaxf(sh_m[0], sh_m[1]), fmaxf(sh_m[2], sh_m[3]));
    float lt = 0.f, o[EPL];
#pragma unroll
    for (int j = 0; j < EPL; ++j) o[j] = 0.f;
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      float f = (sh_m[w] == -INFINITY) ? 0.f : __expf(sh_m[w] - M);
      lt += sh_l[w] * f;
#pragma unroll
      for (int j = 0; j < EPL; ++j) o[j] += sh_acc[w][d0 + j] * f;
    }
    if (act) {
      T* dst = out + ((int64_t)tok * Hq + h) * D;
      float inv = 1.0f / lt;
#pragma unroll
      for (int j = 0; j < EPL; ++j) dst[out_perm ? perm8_pos(d0 + j) : d0 + j] = Tr<T>::from_f(o[j] * inv);
    }
  }
}

int attn_generic_launch(const void* q, const void* kc, const void* vc, const int32_t* lengths, void* out,
                        int64_t B, int64_t L, int Hq, int Hkv, int D, int64_t Tcap, float scale, int dtype,
                        int out_perm, cudaStream_t st) {
  B200P_REQUIRE(!out_perm || D % 8 == 0, B200P_ERR_ARG, "B200P_OUT_PERM8 needs head_dim %% 8 == 0");
  B200P_REQUIRE(D >= 1 && D <= 256, B200P_ERR_ARG, "generic attention supports head_dim <= 256, got %d", D);
  int epl = (D + 31) / 32;
  if (epl == 3) epl = 4;
  if (epl > 4) epl = 8;
  B200P_REQUIRE(D % epl == 0, B200P_ERR_ARG, "head_dim %d not divisible by per-lane width %d", D, epl);
  int64_t nblk = B * L * Hq;
  B200P_REQUIRE(nblk < (1ll << 31), B200P_ERR_ARG, "grid too large");
  if (nblk == 0) return B200P_OK;
#define GO(E_)                                                                                              \
  B200P_DISPATCH_DTYPE(dtype, T, (launch_pdl(attn_generic_kernel<T, E_>, (unsigned)nblk, 128, 0, st,                 \
                                     (const T*)q, (const T*)kc, (const T*)vc, lengths, (T*)out, (int)L, Hq, \
                                     Hkv, D, Tcap, scale, out_perm)))
  switch (epl) {
    case 1: GO(1); break;
    case 2: GO(2); break;
    case 4: GO(4); break;
    default: GO(8); break;
  }
#undef GO
  B200P_LAUNCH_CHECK("attn_generic_kernel");
  return B200P_OK;
}

}  // namespace b200p
