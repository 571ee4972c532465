// Causal flash-style attention for prefill chunks (L > 1 query tokens per row) over the 16-bit KV cache, head_dim 128.
//
//   out[b, l, h, :] = softmax_j( scale * q[b,l,h,:] . K[b, h/G, j, :] ) V[b, h/G, j, :]   for keys j < lengths[b] + l + 1
//
// (the causal rule of create_additive_causal_mask, models/base.py:17-40, with the per-row cache offset of
// models/llama.py:214-223, applied as a bound instead of an additive -1e9 mask; the chunk's own keys were appended by
// b200p_rope_kv_append before this call).
//
// One CTA = 64 consecutive queries of one (row b, query head h): 4 warps x 16 query rows.  K/V tiles of 64 keys stream
// global -> shared with cp.async (two stages, XOR-swizzled 16-byte chunks, shared by the four warps); S = Q K^T and O += P V
// run on mma.sync m16n8k16 with an fp32 online softmax in registers -- the same fragment plumbing as the decode kernel
// (attn_decode.cu).  Tensor-pipe bound in principle (4 L^2 D flop per head), but at prefill sizes of a few hundred tokens per
// row the kernel is short; what matters is that it replaces the CUDA-core generic kernel, which took 40 % of the prefill.
// Legacy mma.sync rather than tcgen05: 64 x 64 x 128 tiles per CTA with a softmax between the two GEMMs fit registers; a
// TMEM pipeline pays off only for much longer sequences.
#include "common.cuh"
#include "mma_util.cuh"

namespace b200p {

constexpr int AP_D = 128;
constexpr int AP_WARPS = 4;
constexpr int AP_THREADS = AP_WARPS * 32;
constexpr int AP_QT = 16 * AP_WARPS;            // queries per CTA
constexpr int AP_KT = 64;                       // keys per tile
constexpr int AP_TILE_BYTES = AP_KT * AP_D * 2; // 16 KB
constexpr int AP_STAGES = 2;
constexpr int AP_SMEM = AP_STAGES * 2 * AP_TILE_BYTES;   // 64 KB

template <typename T>
__global__ void __launch_bounds__(AP_THREADS)
attn_prefill_kernel(const T* __restrict__ q, const T* __restrict__ kc, const T* __restrict__ vc,
                    const int32_t* __restrict__ lengths, T* __restrict__ out, int L, int Hq, int Hkv, int64_t Tcap,
                    float scale_log2e, int out_perm) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int quad_row = lane >> 2, qc = lane & 3;
  const int q0 = blockIdx.x * AP_QT, hq = blockIdx.y, b = blockIdx.z;
  const int kvh = hq / (Hq / Hkv);
  pdl_wait();
  const int len_b = lengths[b];
  // this thread's two query rows (clamped: rows past L compute a valid row and are not stored)
  const int l0 = q0 + warp * 16 + quad_row, l1 = l0 + 8;
  const int lc0 = min(l0, L - 1), lc1 = min(l1, L - 1);
  const int bound0 = (int)min((int64_t)len_b + lc0 + 1, Tcap), bound1 = (int)min((int64_t)len_b + lc1 + 1, Tcap);
  const int kend = (int)min((int64_t)len_b + min(L, q0 + AP_QT), Tcap);     // keys any row of this CTA may see
  const int n_tiles = (kend + AP_KT - 1) / AP_KT;

  uint32_t qf[AP_D / 16][4];
  {
    const uint32_t* p0 = reinterpret_cast<const uint32_t*>(q + (((int64_t)b * L + lc0) * Hq + hq) * AP_D);
    const uint32_t* p1 = reinterpret_cast<const uint32_t*>(q + (((int64_t)b * L + lc1) * Hq + hq) * AP_D);
#pragma unroll
    for (int ks = 0; ks < AP_D / 16; ++ks) {
      qf[ks][0] = p0[ks * 8 + qc];
      qf[ks][1] = p1[ks * 8 + qc];
      qf[ks][2] = p0[ks * 8 + 4 + qc];
      qf[ks][3] = p1[ks * 8 + 4 + qc];
    }
  }

  const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(smem);
  const T* kbase = kc + ((int64_t)b * Hkv + kvh) * Tcap * AP_D;
  const T* vbase = vc + ((int64_t)b * Hkv + kvh) * Tcap * AP_D;
  auto issue_tile = [&](int tile, int stage) {
    const uint32_t ks = sbase + stage * 2 * AP_TILE_BYTES;
    const uint32_t vs = ks + AP_TILE_BYTES;
    const int key0 = tile * AP_KT;
#pragma unroll
    for (int i = 0; i < AP_KT * 16 / AP_THREADS; ++i) {
      const int idx = i * AP_THREADS + threadIdx.x;
      const int key = idx >> 4, ch = idx & 15;
      const bool ok = key0 + key < kend;
      const int64_t goff = (int64_t)(key0 + (ok ? key : 0)) * AP_D + ch * 8;
      const uint32_t soff = (uint32_t)(key * 256 + ((ch ^ (key & 7)) << 4));
      cp_async16(ks + soff, kbase + goff, ok);
      cp_async16(vs + soff, vbase + goff, ok);
    }
  };

  float o[AP_D / 8][4];
#pragma unroll
  for (int j = 0; j < AP_D / 8; ++j) o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f;
  float m0 = -INFINITY, m1 = -INFINITY, s0 = 0.f, s1 = 0.f;     // running max / sum of rows l0, l1

  if (n_tiles > 0) issue_tile(0, 0);
  cp_async_commit();
  for (int it = 0; it < n_tiles; ++it) {
    const int stage = it & 1;
    if (it + 1 < n_tiles) issue_tile(it + 1, stage ^ 1);
    cp_async_commit();
    cp_async_wait<1>();
    __syncthreads();
    const uint32_t ks = sbase + stage * 2 * AP_TILE_BYTES;
    const uint32_t vs = ks + AP_TILE_BYTES;
    const int key0 = it * AP_KT;
    const int mat = lane >> 3, r = lane & 7;

    // ---- S = Q K^T : 8 n-tiles of 8 keys ------------------------------------------------------------
    float sc[AP_KT / 8][4];
#pragma unroll
    for (int nt = 0; nt < AP_KT / 8; ++nt) sc[nt][0] = sc[nt][1] = sc[nt][2] = sc[nt][3] = 0.f;
#pragma unroll
    for (int np = 0; np < AP_KT / 16; ++np) {
      const int key = 16 * np + r + 8 * (mat >> 1);
#pragma unroll
      for (int kstep = 0; kstep < AP_D / 16; ++kstep) {
        const int ch = 2 * kstep + (mat & 1);
        uint32_t bf[4];
        ldsm_x4(ks + (uint32_t)(key * 256 + ((ch ^ (key & 7)) << 4)), bf);
        mma16816<T>(sc[2 * np], qf[kstep], bf[0], bf[1]);
        mma16816<T>(sc[2 * np + 1], qf[kstep], bf[2], bf[3]);
      }
    }
    // ---- scale, causal bound, online softmax ----------------------------------------------------------
    float t0 = -INFINITY, t1 = -INFINITY;
#pragma unroll
    for (int nt = 0; nt < AP_KT / 8; ++nt) {
      const int kcol = key0 + 8 * nt + 2 * qc;
      sc[nt][0] = kcol < bound0 ? sc[nt][0] * scale_log2e : -INFINITY;
      sc[nt][1] = kcol + 1 < bound0 ? sc[nt][1] * scale_log2e : -INFINITY;
      sc[nt][2] = kcol < bound1 ? sc[nt][2] * scale_log2e : -INFINITY;
      sc[nt][3] = kcol + 1 < bound1 ? sc[nt][3] * scale_log2e : -INFINITY;
      t0 = fmaxf(t0, fmaxf(sc[nt][0], sc[nt][1]));
      t1 = fmaxf(t1, fmaxf(sc[nt][2], sc[nt][3]));
    }
    t0 = fmaxf(t0, __shfl_xor_sync(0xffffffffu, t0, 1)); t0 = fmaxf(t0, __shfl_xor_sync(0xffffffffu, t0, 2));
    t1 = fmaxf(t1, __shfl_xor_sync(0xffffffffu, t1, 1)); t1 = fmaxf(t1, __shfl_xor_sync(0xffffffffu, t1, 2));
    const float mn0 = fmaxf(m0, t0), mn1 = fmaxf(m1, t1);      // finite from tile 0 on: key 0 is inside every bound
    const float corr0 = exp2f(m0 - mn0), corr1 = exp2f(m1 - mn1);
    m0 = mn0; m1 = mn1;
    float ps0 = 0.f, ps1 = 0.f;
#pragma unroll
    for (int nt = 0; nt < AP_KT / 8; ++nt) {
      sc[nt][0] = exp2f(sc[nt][0] - mn0); sc[nt][1] = exp2f(sc[nt][1] - mn0);
      sc[nt][2] = exp2f(sc[nt][2] - mn1); sc[nt][3] = exp2f(sc[nt][3] - mn1);
      ps0 += sc[nt][0] + sc[nt][1];
      ps1 += sc[nt][2] + sc[nt][3];
    }
    s0 = s0 * corr0 + ps0;
    s1 = s1 * corr1 + ps1;
#pragma unroll
    for (int j = 0; j < AP_D / 8; ++j) { o[j][0] *= corr0; o[j][1] *= corr0; o[j][2] *= corr1; o[j][3] *= corr1; }
    // ---- O += P V : 4 k-steps of 16 keys x 16 n-tiles of 8 head dims ------------------------------------
#pragma unroll
    for (int kk = 0; kk < AP_KT / 16; ++kk) {
      uint32_t pf[4];
      pf[0] = pack2<T>(sc[2 * kk][0], sc[2 * kk][1]);
      pf[1] = pack2<T>(sc[2 * kk][2], sc[2 * kk][3]);
      pf[2] = pack2<T>(sc[2 * kk + 1][0], sc[2 * kk + 1][1]);
      pf[3] = pack2<T>(sc[2 * kk + 1][2], sc[2 * kk + 1][3]);
      const int key = 16 * kk + r + 8 * (mat & 1);
#pragma unroll
      for (int jp = 0; jp < AP_D / 16; ++jp) {
        const int ch = 2 * jp + (mat >> 1);
        uint32_t bf[4];
        ldsm_x4_trans(vs + (uint32_t)(key * 256 + ((ch ^ (key & 7)) << 4)), bf);
        mma16816<T>(o[2 * jp], pf, bf[0], bf[1]);
        mma16816<T>(o[2 * jp + 1], pf, bf[2], bf[3]);
      }
    }
    __syncthreads();                                 // this stage is refilled by the next iteration's prefetch
  }
  cp_async_wait<0>();
  pdl_launch_dependents();
  s0 += __shfl_xor_sync(0xffffffffu, s0, 1); s0 += __shfl_xor_sync(0xffffffffu, s0, 2);
  s1 += __shfl_xor_sync(0xffffffffu, s1, 1); s1 += __shfl_xor_sync(0xffffffffu, s1, 2);
  const float inv0 = 1.0f / s0, inv1 = 1.0f / s1;
  T* o0 = out + (((int64_t)b * L + lc0) * Hq + hq) * AP_D;
  T* o1 = out + (((int64_t)b * L + lc1) * Hq + hq) * AP_D;
#pragma unroll
  for (int j = 0; j < AP_D / 8; ++j) {
    const int c = 8 * j + 2 * qc;
    if (out_perm) {
      if (l0 < L) { o0[perm8_pos(c)] = Tr<T>::from_f(o[j][0] * inv0); o0[perm8_pos(c + 1)] = Tr<T>::from_f(o[j][1] * inv0); }
      if (l1 < L) { o1[perm8_pos(c)] = Tr<T>::from_f(o[j][2] * inv1); o1[perm8_pos(c + 1)] = Tr<T>::from_f(o[j][3] * inv1); }
    } else {
      if (l0 < L) *reinterpret_cast<uint32_t*>(o0 + c) = pack2<T>(o[j][0] * inv0, o[j][1] * inv0);
      if (l1 < L) *reinterpret_cast<uint32_t*>(o1 + c) = pack2<T>(o[j][2] * inv1, o[j][3] * inv1);
    }
  }
}

bool attn_prefill_supported(int64_t L, int Hq, int Hkv, int D, int dtype) {
  return L > 1 && D == AP_D && (dtype == B200P_F16 || dtype == B200P_BF16) && Hq % Hkv == 0;
}

template <typename T>
static int launch_attn_prefill(const void* q, const void* kc, const void* vc, const int32_t* lengths, void* out, int64_t B,
                               int64_t L, int Hq, int Hkv, int64_t Tcap, float scale, int out_perm, cudaStream_t st) {
  auto kern = attn_prefill_kernel<T>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, AP_SMEM);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(attn_prefill_kernel)");
    attr_set = true;
  }
  dim3 grid((unsigned)((L + AP_QT - 1) / AP_QT), (unsigned)Hq, (unsigned)B);
  launch_pdl(kern, grid, AP_THREADS, AP_SMEM, st, (const T*)q, (const T*)kc, (const T*)vc, lengths, (T*)out, (int)L, Hq, Hkv,
             Tcap, scale * 1.4426950408889634f, out_perm);
  B200P_LAUNCH_CHECK("attn_prefill_kernel");
  return B200P_OK;
}

int attn_prefill_launch(const void* q, const void* kc, const void* vc, const int32_t* lengths, void* out, int64_t B, int64_t L,
                        int Hq, int Hkv, int D, int64_t Tcap, float scale, int dtype, int out_perm, cudaStream_t st) {
  B200P_REQUIRE(attn_prefill_supported(L, Hq, Hkv, D, dtype), B200P_ERR_ARG, "tensor prefill attention: unsupported shape");
  B200P_REQUIRE(B < 65536 && Hq < 65536 && L < (1ll << 30), B200P_ERR_ARG, "grid too large");
  B200P_REQUIRE(aligned(q, 16) && aligned(kc, 16) && aligned(vc, 16) && aligned(out, 4), B200P_ERR_ALIGN,
                "q / caches must be 16-byte aligned");
  if (dtype == B200P_F16) return launch_attn_prefill<__half>(q, kc, vc, lengths, out, B, L, Hq, Hkv, Tcap, scale, out_perm, st);
  return launch_attn_prefill<__nv_bfloat16>(q, kc, vc, lengths, out, B, L, Hq, Hkv, Tcap, scale, out_perm, st);
}

}  // namespace b200p
