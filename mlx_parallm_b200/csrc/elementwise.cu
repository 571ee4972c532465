// dequantize / embedding gather / RMSNorm / RoPE + KV append / argmax.
// All HBM- or latency-bound byte work: one pass, 128-bit accesses, no tensor cores.
#include <type_traits>

#include <cooperative_groups.h>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace b200p {

// ------------------------------------------------------------------------------------------------
// dequantize: out[n,k] = q*s + b under one of three rounding recipes (include/b200p.h B200P_DQ_*)
// ------------------------------------------------------------------------------------------------
template <typename T, int MODE>
__device__ __forceinline__ float dq_one(uint32_t q, float s, float b) {
  float qf = u8_to_float(q);
  if (MODE == B200P_DQ_F32_MULADD) {
    return __fadd_rn(__fmul_rn(qf, s), b);
  } else if (MODE == B200P_DQ_FMA) {
    return __fmaf_rn(qf, s, b);
  } else {
    float p = Tr<T>::to_f(Tr<T>::from_f(__fmul_rn(qf, s)));
    return __fadd_rn(p, b);
  }
}

// one thread per FOUR packed uint32 words (one 16-byte load, 32 / 16 outputs = 64 bytes of 16-bit stores); blockIdx.y = row,
// so there is no 64-bit division on the index path; the row index comes from `rows` (gather) or is the identity.
template <typename T, int BITS, int MODE, typename OUT>
__global__ void __launch_bounds__(256)
dequant_kernel(const uint32_t* __restrict__ w, const T* __restrict__ scales,
               const T* __restrict__ biases, OUT* __restrict__ out,
               const int32_t* __restrict__ rows, int64_t row0, int64_t V, int64_t K, int gs) {
  constexpr int PER = 32 / BITS;
  const int words_per_row = (int)(K / PER);
  const int w0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;          // first of this thread's 4 words
  if (w0 >= words_per_row) return;
  pdl_wait();   // `rows` (token ids) and `out` belong to the previous kernels
  const int64_t r = row0 + blockIdx.y;
  int64_t src = rows ? (int64_t)rows[r] : r;
  if (src < 0 || src >= V) src = 0;  // out-of-vocab ids read row 0 (MLX would index modulo/UB)
  const uint32_t* wp = w + src * words_per_row + w0;
  uint32_t wd[4];
  const int nw = min(4, words_per_row - w0);
  if (nw == 4 && ((reinterpret_cast<uintptr_t>(wp) & 15) == 0)) {
    const uint4 u = ldg_stream(wp);
    wd[0] = u.x; wd[1] = u.y; wd[2] = u.z; wd[3] = u.w;
  } else {
#pragma unroll
    for (int i = 0; i < 4; ++i) wd[i] = i < nw ? wp[i] : 0u;
  }
  const int ng = (int)(K / gs);
  OUT* dst_row = out + r * K;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    if (i >= nw) break;
    const int k0 = (w0 + i) * PER;
    const int g = k0 / gs;
    const float s = Tr<T>::to_f(scales[src * ng + g]);
    const float b = Tr<T>::to_f(biases[src * ng + g]);
    __align__(16) OUT vals[PER];
#pragma unroll
    for (int j = 0; j < PER; ++j) {
      const uint32_t q = (wd[i] >> (j * BITS)) & ((1u << BITS) - 1u);
      const float v = dq_one<T, MODE>(q, s, b);
      const T vt = Tr<T>::from_f(v);  // storage rounding always happens
      vals[j] = Tr<OUT>::from_f(Tr<T>::to_f(vt));
    }
    OUT* dst = dst_row + k0;
    constexpr int BYTES = PER * sizeof(OUT);
    if (BYTES % 16 == 0) {
#pragma unroll
      for (int j = 0; j < BYTES / 16; ++j) reinterpret_cast<uint4*>(dst)[j] = reinterpret_cast<uint4*>(vals)[j];
    } else {
#pragma unroll
      for (int j = 0; j < BYTES / 8; ++j) reinterpret_cast<uint2*>(dst)[j] = reinterpret_cast<uint2*>(vals)[j];
    }
  }
}

template <typename T, typename OUT>
static int launch_dequant(const uint32_t* w, const void* scales, const void* biases, void* out,
                          const int32_t* rows, int64_t n_rows, int64_t V, int64_t K, int gs, int bits,
                          int mode, cudaStream_t st) {
  const int per = 32 / bits;
  const int64_t words_per_row = K / per;
  if (n_rows * words_per_row == 0) return B200P_OK;
  B200P_REQUIRE(n_rows < 65536 * 1024ll && words_per_row < (1ll << 30), B200P_ERR_ARG, "matrix too large for one dequantize launch");
  // rows go to blockIdx.y (<= 65535 per launch): larger matrices are cut into row slabs
  const int threads = words_per_row / 4 >= 256 ? 256 : (int)(((words_per_row + 3) / 4 + 31) / 32 * 32);
  dim3 block((unsigned)threads);
  const unsigned gx = (unsigned)(((words_per_row + 3) / 4 + threads - 1) / threads);
#define L(B_, M_)                                                                                   \
  launch_pdl(dequant_kernel<T, B_, M_, OUT>, grid, block, 0, st, w, (const T*)scales, (const T*)biases,      \
                                                         (OUT*)out, rows, row0, V, K, gs)
  for (int64_t row0 = 0; row0 < n_rows; row0 += 65535) {
    const int64_t nr = n_rows - row0 < 65535 ? n_rows - row0 : 65535;
    dim3 grid(gx, (unsigned)nr);
    if (bits == 4) {
      if (mode == 0) L(4, 0); else if (mode == 1) L(4, 1); else L(4, 2);
    } else {
      if (mode == 0) L(8, 0); else if (mode == 1) L(8, 1); else L(8, 2);
    }
  }
#undef L
  B200P_LAUNCH_CHECK("dequant_kernel");
  return B200P_OK;
}

static int check_quant_args(const void* w, const void* s, const void* b, int64_t K, int gs, int bits, int mode) {
  B200P_REQUIRE(w && s && b, B200P_ERR_ARG, "null quantized-weight pointer");
  B200P_REQUIRE(bits == 4 || bits == 8, B200P_ERR_ARG, "bits must be 4 or 8, got %d", bits);
  B200P_REQUIRE(gs >= 32 && gs % 32 == 0 && K % gs == 0, B200P_ERR_ARG,
                "group_size %d must be a multiple of 32 dividing K=%lld", gs, (long long)K);
  B200P_REQUIRE(mode >= 0 && mode <= 2, B200P_ERR_ARG, "bad dequantize mode %d", mode);
  B200P_REQUIRE(aligned(w, 16), B200P_ERR_ALIGN, "w must be 16-byte aligned");
  return B200P_OK;
}

}  // namespace b200p

using namespace b200p;

extern "C" int b200p_dequantize(const uint32_t* w, const void* scales, const void* biases, void* out,
                                int64_t N, int64_t K, int group_size, int bits, int dtype, int mode,
                                void* stream) {
  PdlClass pdl_cls(32);
  B200P_REQUIRE(N >= 0, B200P_ERR_ARG, "N < 0");
  if (N == 0) return B200P_OK;                      // empty input: nothing to launch (pointers may be null)
  int rc = check_quant_args(w, scales, biases, K, group_size, bits, mode);
  if (rc) return rc;
  B200P_REQUIRE(out && aligned(out, 16), B200P_ERR_ALIGN, "out must be non-null and 16-byte aligned");
  B200P_DISPATCH_DTYPE(dtype, T, return (launch_dequant<T, T>(w, scales, biases, out, nullptr, N, N, K,
                                                              group_size, bits, mode, (cudaStream_t)stream)));
  return B200P_OK;
}

extern "C" int b200p_embed(const int32_t* ids, const uint32_t* w, const void* scales, const void* biases,
                           void* out, int64_t M, int64_t V, int64_t K, int group_size, int bits, int dtype,
                           int mode, int out_f32, void* stream) {
  PdlClass pdl_cls(32);
  if (M == 0) return B200P_OK;
  int rc = check_quant_args(w, scales, biases, K, group_size, bits, mode);
  if (rc) return rc;
  B200P_REQUIRE(ids && out && aligned(out, 16), B200P_ERR_ALIGN, "ids/out must be non-null, out 16-byte aligned");
  if (out_f32) {
    B200P_DISPATCH_DTYPE(dtype, T, return (launch_dequant<T, float>(w, scales, biases, out, ids, M, V, K,
                                                                    group_size, bits, mode, (cudaStream_t)stream)));
  } else {
    B200P_DISPATCH_DTYPE(dtype, T, return (launch_dequant<T, T>(w, scales, biases, out, ids, M, V, K,
                                                                group_size, bits, mode, (cudaStream_t)stream)));
  }
  return B200P_OK;
}

// ------------------------------------------------------------------------------------------------
// RMSNorm: fp32 residual row -> dtype row.  One CTA per row.
// ------------------------------------------------------------------------------------------------
namespace b200p {

__device__ __forceinline__ float block_sum(float v, float* sh) {
  v = warp_sum(v);
  int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  if (lane == 0) sh[wid] = v;
  __syncthreads();
  float t = (threadIdx.x < nw) ? sh[threadIdx.x] : 0.f;
  if (wid == 0) {
    t = warp_sum(t);
    if (lane == 0) sh[32] = t;
  }
  __syncthreads();
  float r = sh[32];
  __syncthreads();
  return r;
}

// NT = 128 is the "early" variant (B200P_RMS_EARLY): at most 32 registers per thread, so that a CTA fits next to a resident
// tensor-path GEMM CTA (768 threads x 80 registers), launched with the PDL attribute, and it releases ITS dependents before
// waiting for its own producer -- the GEMM that follows the norm is then scheduled while the GEMM before the norm drains.
template <typename T, int NT>
__global__ void __launch_bounds__(NT, NT == 128 ? 16 : 1)
rmsnorm_kernel(const float* __restrict__ x, const T* __restrict__ w, T* __restrict__ y, int64_t H, float eps, int perm) {
  __shared__ float sh[33];
  if constexpr (NT == 128) pdl_launch_dependents();
  pdl_wait();
  const float* xr = x + (int64_t)blockIdx.x * H;
  T* yr = y + (int64_t)blockIdx.x * H;
  float ss = 0.f;
  const int64_t H4 = H / 4;
  for (int64_t i = threadIdx.x; i < H4; i += blockDim.x) {
    float4 v = reinterpret_cast<const float4*>(xr)[i];
    ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  for (int64_t i = H4 * 4 + threadIdx.x; i < H; i += blockDim.x) ss += xr[i] * xr[i];
  ss = block_sum(ss, sh);
  float inv = 1.0f / sqrtf(ss / (float)H + eps);
  pdl_launch_dependents();
  if (perm) {   // 16-bit output in the PERM8 order: one thread per aligned group of 8 (H % 8 == 0 checked by the host)
    if constexpr (sizeof(T) == 2) {
      for (int64_t i = threadIdx.x; i < H / 8; i += blockDim.x) {
        const float4 a = reinterpret_cast<const float4*>(xr)[2 * i], b = reinterpret_cast<const float4*>(xr)[2 * i + 1];
        const uint4 wv = reinterpret_cast<const uint4*>(w)[i];
        const T* wp = reinterpret_cast<const T*>(&wv);
        __align__(16) T o[8];
        o[0] = Tr<T>::from_f(a.x * inv * Tr<T>::to_f(wp[0]));
        o[2] = Tr<T>::from_f(a.y * inv * Tr<T>::to_f(wp[1]));
        o[4] = Tr<T>::from_f(a.z * inv * Tr<T>::to_f(wp[2]));
        o[6] = Tr<T>::from_f(a.w * inv * Tr<T>::to_f(wp[3]));
        o[1] = Tr<T>::from_f(b.x * inv * Tr<T>::to_f(wp[4]));
        o[3] = Tr<T>::from_f(b.y * inv * Tr<T>::to_f(wp[5]));
        o[5] = Tr<T>::from_f(b.z * inv * Tr<T>::to_f(wp[6]));
        o[7] = Tr<T>::from_f(b.w * inv * Tr<T>::to_f(wp[7]));
        reinterpret_cast<uint4*>(yr)[i] = *reinterpret_cast<uint4*>(o);
      }
    }
    return;
  }
  for (int64_t i = threadIdx.x; i < H4; i += blockDim.x) {
    float4 v = reinterpret_cast<const float4*>(xr)[i];
    __align__(16) T o[4];
    o[0] = Tr<T>::from_f(v.x * inv * Tr<T>::to_f(w[4 * i + 0]));
    o[1] = Tr<T>::from_f(v.y * inv * Tr<T>::to_f(w[4 * i + 1]));
    o[2] = Tr<T>::from_f(v.z * inv * Tr<T>::to_f(w[4 * i + 2]));
    o[3] = Tr<T>::from_f(v.w * inv * Tr<T>::to_f(w[4 * i + 3]));
    if (sizeof(T) == 2) reinterpret_cast<uint2*>(yr)[i] = *reinterpret_cast<uint2*>(o);
    else reinterpret_cast<uint4*>(yr)[i] = *reinterpret_cast<uint4*>(o);
  }
  for (int64_t i = H4 * 4 + threadIdx.x; i < H; i += blockDim.x)
    yr[i] = Tr<T>::from_f(xr[i] * inv * Tr<T>::to_f(w[i]));
}

// "early" RMSNorm in ONE pass over the row: the row is split over a thread-block cluster of C = H / 2048 CTAs (2 at H = 4096,
// 4 at 8192), every thread keeps its 16 elements in registers between the sum of squares and the output, and the C partial
// sums meet through distributed shared memory (each CTA writes its sum into slot [rank] of every CTA, one barrier.cluster).
// The generic kernel above reads the row twice, and with 32 registers per thread its second read is four dependent round
// trips to L2 (a resident GEMM CTA leaves no L1).  Same launch contract as the NT = 128 variant: 128 threads x 32 registers,
// PDL attribute, dependents released before its own wait.  The norm weights do not depend on the previous kernel and are
// loaded before the wait.  The sum is taken in (warp, CTA rank) order: reproducible, not bit-identical to the generic kernel's.
template <typename T>
__global__ void __launch_bounds__(128, 16)
rmsnorm_cluster_kernel(const float* __restrict__ x, const T* __restrict__ w, T* __restrict__ y, int64_t H, float eps, int perm) {
  static_assert(sizeof(T) == 2, "16-bit outputs only");
  constexpr int NG = 2;                            // groups of 8 consecutive elements per thread
  __shared__ float s_warp[4];
  __shared__ float s_cta[8];
  cg::cluster_group cluster = cg::this_cluster();
  const unsigned rank = cluster.block_rank(), C = cluster.num_blocks();
  const int tid = threadIdx.x;
  pdl_launch_dependents();
  const int64_t row = blockIdx.x / C;
  uint4 wv[NG];
#pragma unroll
  for (int j = 0; j < NG; ++j) wv[j] = reinterpret_cast<const uint4*>(w)[(j * C + rank) * 128 + tid];
  pdl_wait();
  const float4* xr = reinterpret_cast<const float4*>(x + row * H);
  float4 a[NG], b[NG];
#pragma unroll
  for (int j = 0; j < NG; ++j) {
    const int64_t g = (int64_t)(j * C + rank) * 128 + tid;
    a[j] = xr[2 * g];
    b[j] = xr[2 * g + 1];
  }
  float ss = 0.f;
#pragma unroll
  for (int j = 0; j < NG; ++j)
    ss += a[j].x * a[j].x + a[j].y * a[j].y + a[j].z * a[j].z + a[j].w * a[j].w + b[j].x * b[j].x + b[j].y * b[j].y +
          b[j].z * b[j].z + b[j].w * b[j].w;
  ss = warp_sum(ss);
  if ((tid & 31) == 0) s_warp[tid >> 5] = ss;
  __syncthreads();
  if (tid < (int)C) *cluster.map_shared_rank(&s_cta[rank], tid) = (s_warp[0] + s_warp[1]) + (s_warp[2] + s_warp[3]);
  cluster.sync();
  float tot = 0.f;
  for (unsigned r = 0; r < C; ++r) tot += s_cta[r];
  const float inv = 1.0f / sqrtf(tot / (float)H + eps);
  uint4* yr = reinterpret_cast<uint4*>(y + row * H);
#pragma unroll
  for (int j = 0; j < NG; ++j) {
    const T* wp = reinterpret_cast<const T*>(&wv[j]);
    const float v[8] = {a[j].x, a[j].y, a[j].z, a[j].w, b[j].x, b[j].y, b[j].z, b[j].w};
    __align__(16) T o[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const T r = Tr<T>::from_f(v[e] * inv * Tr<T>::to_f(wp[e]));
      // PERM8: element e of the group sits at position [0,4,1,5,2,6,3,7]^-1 = perm8_pos(e) (compile-time after unrolling)
      if (perm) o[perm8_pos(e)] = r; else o[e] = r;
    }
    yr[(int64_t)(j * C + rank) * 128 + tid] = *reinterpret_cast<uint4*>(o);
  }
}

bool pdl_enabled_gemm();      // the global PDL switch (b200p_set_pdl / B200P_PDL != 0): the early norms follow the GEMM launches

template <typename T>
static cudaError_t launch_rmsnorm_cluster(int C, int64_t M, cudaStream_t st, const float* x, const void* w, void* y, int64_t H,
                                          float eps, int perm) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(M * C));
  cfg.blockDim = dim3(128);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)C;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled_gemm() ? 2 : 1;       // without the attribute the griddepcontrol instructions are no-ops
  return cudaLaunchKernelEx(&cfg, rmsnorm_cluster_kernel<T>, x, (const T*)w, (T*)y, H, eps, perm);
}
int sm_count();

}  // namespace b200p

extern "C" int b200p_rmsnorm(const float* x, const void* w, void* y, int64_t M, int64_t H, float eps,
                             int dtype, int flags, void* stream) {
  PdlClass pdl_cls(2);
  B200P_REQUIRE(x && w && y, B200P_ERR_ARG, "null pointer");
  B200P_REQUIRE((flags & ~B200P_OUT_PERM8) == 0, B200P_ERR_ARG, "unknown flags 0x%x", flags);
  const int perm = (flags & B200P_OUT_PERM8) ? 1 : 0;
  B200P_REQUIRE(!perm || (dtype != B200P_F32 && H % 8 == 0 && aligned(w, 16)), B200P_ERR_ARG,
                "B200P_OUT_PERM8 needs a 16-bit dtype, H %% 8 == 0 and a 16-byte aligned weight");
  B200P_REQUIRE(aligned(x, 16) && aligned(y, 16) && H % 4 == 0, B200P_ERR_ALIGN,
                "rmsnorm needs 16-byte aligned rows (H %% 4 == 0)");
  if (M == 0) return B200P_OK;
  // Which norm.  The one-pass cluster form (16-bit rows of 2048 * C elements, C in {1, 2, 4, 8}, at most one CTA per SM) is an
  // "early" kernel and wins at every row count (B=1 T=8192 2.29 -> 2.22 ms, B=8 3.46 -> 3.41, B=64 T=256 2.88 -> 2.86:
  // profiles/r2_run33_*, r2_run37_*).  Where it does not apply, the two-pass early kernel pays from 16 rows on (the next kernel
  // is then a tile GEMM that streams weights while it waits: B=64 T=256 2.91 -> 2.87 ms) and loses below (a 128-thread CTA is
  // the slower norm and the GEMV after it gains nothing: B=1 T=8192 2.38 -> 2.43 ms, profiles/r2_run29_*).
  static const int early_env = [] { const char* e = getenv("B200P_RMS_EARLY"); return e ? atoi(e) : -1; }();
  static const int cluster_env = [] { const char* e = getenv("B200P_RMS_CLUSTER"); return e ? atoi(e) : 1; }();
  const int64_t C = H / 2048;
  const bool cluster_ok = cluster_env && dtype != B200P_F32 && H % 2048 == 0 && (C == 1 || C == 2 || C == 4 || C == 8) &&
                          M * C <= sm_count() && aligned(w, 16);
  const bool early = early_env >= 0 ? early_env != 0 : (M >= 16 || cluster_ok);
  if (early && cluster_ok) {
    cudaError_t e = dtype == B200P_F16
        ? launch_rmsnorm_cluster<__half>((int)C, M, (cudaStream_t)stream, x, w, y, H, eps, perm)
        : launch_rmsnorm_cluster<__nv_bfloat16>((int)C, M, (cudaStream_t)stream, x, w, y, H, eps, perm);
    if (e != cudaSuccess) return cuda_fail(e, "rmsnorm_cluster_kernel");
    return B200P_OK;
  }
  if (early) {
    B200P_DISPATCH_DTYPE(dtype, T, (launch_pdl_opt(b200p::pdl_enabled_gemm(), rmsnorm_kernel<T, 128>, (unsigned)M, 128, 0, (cudaStream_t)stream,
                                       x, (const T*)w, (T*)y, H, eps, perm)));
  } else {
    B200P_DISPATCH_DTYPE(dtype, T, (launch_pdl(rmsnorm_kernel<T, 256>, (unsigned)M, 256, 0, (cudaStream_t)stream,
                                       x, (const T*)w, (T*)y, H, eps, perm)));
  }
  B200P_LAUNCH_CHECK("rmsnorm_kernel");
  return B200P_OK;
}

// ------------------------------------------------------------------------------------------------
// (qwen3 q/k head RMSNorm) + RoPE (non-traditional pairing i, i+D/2) + KV-cache append.
// grid = (B*L, Hq + 2*Hkv); block = D/2 threads, thread i owns the pair (i, i + D/2).
// ------------------------------------------------------------------------------------------------
namespace b200p {

// Rows whose position reached the cache capacity: their K / V are NOT written (there is no room) and the step's attention
// would read a stale slot -- plausible but wrong tokens.  The Python surface grows the cache before every step and cannot get
// here; a direct C-ABI caller can.  Every such (token, V head) bumps this per-device counter, read with
// b200p_kv_overflow_count at generation boundaries.
__device__ unsigned int g_kv_overflow = 0;

// TIN = T: qkv is the projection's 16-bit output.  TIN = float: qkv is the fp32 split-K accumulation buffer of the
// projection (b200p_qmm with ADD_RESIDUAL into a zeroed buffer): values are rounded to T first (the rounding the
// reference's projection output has) and the buffer is re-zeroed behind the read, ready for the next step.
template <typename T, typename TIN>
__global__ void rope_kv_append_kernel(TIN* __restrict__ qkv, T* __restrict__ q_out, T* __restrict__ kc,
                                      T* __restrict__ vc, const int32_t* __restrict__ lengths,
                                      const T* __restrict__ qn_w, const T* __restrict__ kn_w, int L, int Hq,
                                      int Hkv, int D, int64_t Tcap, float log2_base, float rope_scale,
                                      float eps) {
  __shared__ float sh[33];
  pdl_wait();
  const int tok = blockIdx.x;  // b*L + l
  const int b = tok / L, l = tok - b * L;
  const int head = blockIdx.y;
  const int half = D >> 1;
  const int i = threadIdx.x;
  const int64_t row_stride = (int64_t)(Hq + 2 * Hkv) * D;
  TIN* src = qkv + (int64_t)tok * row_stride + (int64_t)head * D;
  const int pos_i = lengths[b] + l;
  const bool act = i < half;  // block is padded to whole warps
  constexpr bool F32IN = sizeof(TIN) != sizeof(T) || std::is_same<TIN, float>::value;
  float x1 = 0.f, x2 = 0.f;
  if (act) {
    if constexpr (std::is_same<TIN, float>::value && !std::is_same<T, float>::value) {
      x1 = Tr<T>::to_f(Tr<T>::from_f(src[i]));
      x2 = Tr<T>::to_f(Tr<T>::from_f(src[i + half]));
      src[i] = 0.f;
      src[i + half] = 0.f;
    } else {
      x1 = Tr<TIN>::to_f(src[i]);
      x2 = Tr<TIN>::to_f(src[i + half]);
    }
  }
  (void)F32IN;
  if (head >= Hq + Hkv) {  // V head: plain append
    int hv = head - Hq - Hkv;
    if (pos_i >= Tcap && i == 0) atomicAdd(&g_kv_overflow, 1u);
    if (pos_i < Tcap && act) {
      T* dst = vc + (((int64_t)b * Hkv + hv) * Tcap + pos_i) * D;
      dst[i] = Tr<T>::from_f(x1);
      dst[i + half] = Tr<T>::from_f(x2);
    }
    return;
  }
  const bool is_q = head < Hq;
  const T* nw = is_q ? qn_w : kn_w;
  if (nw != nullptr) {  // mx.fast.rms_norm over head_dim (qwen3.py:65-70), rounded to storage type
    float ss = block_sum(x1 * x1 + x2 * x2, sh);
    float inv = 1.0f / sqrtf(ss / (float)D + eps);
    if (act) {
      x1 = Tr<T>::to_f(Tr<T>::from_f(x1 * inv * Tr<T>::to_f(nw[i])));
      x2 = Tr<T>::to_f(Tr<T>::from_f(x2 * inv * Tr<T>::to_f(nw[i + half])));
    }
  }
  if (!act) return;
  // theta = pos * scale * base^(-i/half)
  float inv_freq = exp2f(-(float)i * (log2_base / (float)half));
  float ang = (float)pos_i * rope_scale * inv_freq;
  float sn, cs;
  sincosf(ang, &sn, &cs);
  float o1 = x1 * cs - x2 * sn;
  float o2 = x1 * sn + x2 * cs;
  if (is_q) {
    T* dst = q_out + ((int64_t)tok * Hq + head) * D;
    dst[i] = Tr<T>::from_f(o1);
    dst[i + half] = Tr<T>::from_f(o2);
  } else if (pos_i < Tcap) {
    int hk = head - Hq;
    T* dst = kc + (((int64_t)b * Hkv + hk) * Tcap + pos_i) * D;
    dst[i] = Tr<T>::from_f(o1);
    dst[i + half] = Tr<T>::from_f(o2);
  }
}

// head_dim 128 (every BASELINE model): a WARP per (token, head), 8 heads of a token per CTA; lane = pairs (2*lane, 2*lane + 1)
// of the first half and their partners 64 further.  The kernel is pure latency (4 MB of traffic): every warp issues its
// position load and its four 8-byte value loads at once and finishes after one round trip; the per-head RMSNorm of Qwen3 is a
// warp reduction (no block barrier); 8-byte (fp32 input) / 4-byte (16-bit) accesses per lane, whole lines per warp.
// (A warp looping over 6 heads to reuse its sincos was tried first: 6 dependent round trips, slower -- 18 us cold in ncu
// against 14.6 us for the generic kernel above, profiles/r1_run37_small_kernels_ncu.txt.)
template <typename T, typename TIN>
__global__ void __launch_bounds__(256)
rope_kv_append_d128_kernel(TIN* __restrict__ qkv, T* __restrict__ q_out, T* __restrict__ kc, T* __restrict__ vc,
                           const int32_t* __restrict__ lengths, const T* __restrict__ qn_w, const T* __restrict__ kn_w,
                           int L, int Hq, int Hkv, int64_t Tcap, float log2_base, float rope_scale, float eps) {
  constexpr int D = 128, HALF = 64;
  constexpr bool ACC_IN = std::is_same<TIN, float>::value && !std::is_same<T, float>::value;
  pdl_wait();
  const int tok = blockIdx.x;  // b*L + l
  const int b = tok / L, l = tok - b * L;
  const int lane = threadIdx.x & 31;
  const int H = Hq + 2 * Hkv;
  const int head = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (head >= H) return;
  const int i0 = 2 * lane;
  const int pos_i = lengths[b] + l;
  TIN* row = qkv + (int64_t)tok * H * D;
  {
    TIN* src = row + (int64_t)head * D;
    float x1[2], x2[2];
    if constexpr (std::is_same<TIN, float>::value) {
      const float2 a = *reinterpret_cast<const float2*>(src + i0), c = *reinterpret_cast<const float2*>(src + HALF + i0);
      x1[0] = a.x; x1[1] = a.y; x2[0] = c.x; x2[1] = c.y;
      if constexpr (ACC_IN) {    // the projection's fp32 accumulator: round as its 16-bit output would be, re-zero behind the read
#pragma unroll
        for (int j = 0; j < 2; ++j) { x1[j] = Tr<T>::to_f(Tr<T>::from_f(x1[j])); x2[j] = Tr<T>::to_f(Tr<T>::from_f(x2[j])); }
        *reinterpret_cast<float2*>(src + i0) = make_float2(0.f, 0.f);
        *reinterpret_cast<float2*>(src + HALF + i0) = make_float2(0.f, 0.f);
      }
    } else {
      x1[0] = Tr<TIN>::to_f(src[i0]); x1[1] = Tr<TIN>::to_f(src[i0 + 1]);
      x2[0] = Tr<TIN>::to_f(src[HALF + i0]); x2[1] = Tr<TIN>::to_f(src[HALF + i0 + 1]);
    }
    T* dst;
    float o1[2], o2[2];
    if (head >= Hq + Hkv) {      // V head: plain append
      if (pos_i >= Tcap) {
        if (lane == 0) atomicAdd(&g_kv_overflow, 1u);
        return;
      }
      dst = vc + (((int64_t)b * Hkv + (head - Hq - Hkv)) * Tcap + pos_i) * D;
      o1[0] = x1[0]; o1[1] = x1[1]; o2[0] = x2[0]; o2[1] = x2[1];
    } else {
      const bool is_q = head < Hq;
      const T* nw = is_q ? qn_w : kn_w;
      if (nw != nullptr) {       // mx.fast.rms_norm over head_dim (qwen3.py:65-70), rounded to the storage type
        const float ss = warp_sum(x1[0] * x1[0] + x1[1] * x1[1] + x2[0] * x2[0] + x2[1] * x2[1]);
        const float inv = 1.0f / sqrtf(ss / (float)D + eps);
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          x1[j] = Tr<T>::to_f(Tr<T>::from_f(x1[j] * inv * Tr<T>::to_f(nw[i0 + j])));
          x2[j] = Tr<T>::to_f(Tr<T>::from_f(x2[j] * inv * Tr<T>::to_f(nw[HALF + i0 + j])));
        }
      }
#pragma unroll
      for (int j = 0; j < 2; ++j) {  // theta = pos * scale * base^(-i/half)
        const float inv_freq = exp2f(-(float)(i0 + j) * (log2_base / (float)HALF));
        float sn, cs;
        sincosf((float)pos_i * rope_scale * inv_freq, &sn, &cs);
        o1[j] = x1[j] * cs - x2[j] * sn;
        o2[j] = x1[j] * sn + x2[j] * cs;
      }
      if (is_q) {
        dst = q_out + ((int64_t)tok * Hq + head) * D;
      } else {
        if (pos_i >= Tcap) return;
        dst = kc + (((int64_t)b * Hkv + (head - Hq)) * Tcap + pos_i) * D;
      }
    }
    if constexpr (sizeof(T) == 2) {     // one 32-bit store per pair: 128 contiguous bytes per warp
      __align__(4) T p1[2] = {Tr<T>::from_f(o1[0]), Tr<T>::from_f(o1[1])};
      __align__(4) T p2[2] = {Tr<T>::from_f(o2[0]), Tr<T>::from_f(o2[1])};
      *reinterpret_cast<uint32_t*>(dst + i0) = *reinterpret_cast<const uint32_t*>(p1);
      *reinterpret_cast<uint32_t*>(dst + HALF + i0) = *reinterpret_cast<const uint32_t*>(p2);
    } else {
      dst[i0] = Tr<T>::from_f(o1[0]);
      dst[i0 + 1] = Tr<T>::from_f(o1[1]);
      dst[HALF + i0] = Tr<T>::from_f(o2[0]);
      dst[HALF + i0 + 1] = Tr<T>::from_f(o2[1]);
    }
  }
}

// act[m, i] = silu(gu[m, i]) * gu[m, I + i] from the fp32 accumulation buffer of the fused gate|up projection
// (models/llama.py:165); the buffer is re-zeroed behind the read.  One thread per aligned group of 8 outputs.
template <typename T>
__global__ void __launch_bounds__(256)
silu_mul_kernel(float* __restrict__ gu, T* __restrict__ out, int64_t M, int64_t I, int64_t ldg, int64_t ldo, int perm) {
  pdl_wait();
  const int64_t groups = I / 8;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= M * groups) return;
  const int64_t m = idx / groups, g8 = idx - m * groups;
  float4* gp = reinterpret_cast<float4*>(gu + m * ldg + g8 * 8);
  float4* up = reinterpret_cast<float4*>(gu + m * ldg + I + g8 * 8);
  const float4 g0 = gp[0], g1 = gp[1], u0 = up[0], u1 = up[1];
  const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
  gp[0] = z; gp[1] = z; up[0] = z; up[1] = z;
  pdl_launch_dependents();
  const float gv[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
  const float uv[8] = {u0.x, u0.y, u0.z, u0.w, u1.x, u1.y, u1.z, u1.w};
  __align__(16) T o[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float a = gv[j] / (1.0f + __expf(-gv[j])) * uv[j];
    o[perm ? perm8_pos(j) : j] = Tr<T>::from_f(a);
  }
  *reinterpret_cast<uint4*>(out + m * ldo + g8 * 8) = *reinterpret_cast<uint4*>(o);
}

__global__ void advance_lengths_kernel(int32_t* lengths, int64_t B, int32_t delta) {
  pdl_wait();
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < B) lengths[i] += delta;
}

}  // namespace b200p

extern "C" int b200p_rope_kv_append(void* qkv, void* q_out, void* kcache, void* vcache,
                                    const int32_t* lengths, const void* q_norm_w, const void* k_norm_w,
                                    int64_t B, int64_t L, int Hq, int Hkv, int D, int64_t Tcap,
                                    float rope_base, float rope_scale, float norm_eps, int dtype,
                                    int qkv_f32, void* stream) {
  PdlClass pdl_cls(4);
  B200P_REQUIRE(qkv && q_out && kcache && vcache && lengths, B200P_ERR_ARG, "null pointer");
  B200P_REQUIRE(!qkv_f32 || dtype != B200P_F32, B200P_ERR_ARG, "qkv_f32 is the accumulation-buffer input of 16-bit models");
  B200P_REQUIRE(D >= 2 && D % 2 == 0 && D <= 2048, B200P_ERR_ARG, "head_dim %d unsupported", D);
  B200P_REQUIRE(Hq > 0 && Hkv > 0 && Hq % Hkv == 0, B200P_ERR_ARG, "bad head counts %d/%d", Hq, Hkv);
  B200P_REQUIRE(B * L < (1ll << 31) && Hq + 2 * Hkv < 65536, B200P_ERR_ARG, "grid too large");
  if (B * L == 0) return B200P_OK;
  if (D == 128) {
    const dim3 g128((unsigned)(B * L), (unsigned)((Hq + 2 * Hkv + 7) / 8)), b128(256);
    if (qkv_f32) {
      B200P_DISPATCH_DTYPE(dtype, T, (launch_pdl(rope_kv_append_d128_kernel<T, float>, g128, b128, 0, (cudaStream_t)stream,
                                         (float*)qkv, (T*)q_out, (T*)kcache, (T*)vcache, lengths, (const T*)q_norm_w,
                                         (const T*)k_norm_w, (int)L, Hq, Hkv, Tcap, log2f(rope_base), rope_scale, norm_eps)));
    } else {
      B200P_DISPATCH_DTYPE(dtype, T, (launch_pdl(rope_kv_append_d128_kernel<T, T>, g128, b128, 0, (cudaStream_t)stream,
                                         (T*)qkv, (T*)q_out, (T*)kcache, (T*)vcache, lengths, (const T*)q_norm_w,
                                         (const T*)k_norm_w, (int)L, Hq, Hkv, Tcap, log2f(rope_base), rope_scale, norm_eps)));
    }
    B200P_LAUNCH_CHECK("rope_kv_append_d128_kernel");
    return B200P_OK;
  }
  dim3 grid((unsigned)(B * L), (unsigned)(Hq + 2 * Hkv)), block(((D / 2 + 31) / 32) * 32);
  if (qkv_f32) {
    B200P_DISPATCH_DTYPE(dtype, T, (launch_pdl(rope_kv_append_kernel<T, float>, grid, block, 0, (cudaStream_t)stream, 
                                       (float*)qkv, (T*)q_out, (T*)kcache, (T*)vcache, lengths,
                                       (const T*)q_norm_w, (const T*)k_norm_w, (int)L, Hq, Hkv, D, Tcap,
                                       log2f(rope_base), rope_scale, norm_eps)));
  } else {
    B200P_DISPATCH_DTYPE(dtype, T, (launch_pdl(rope_kv_append_kernel<T, T>, grid, block, 0, (cudaStream_t)stream, 
                                       (T*)qkv, (T*)q_out, (T*)kcache, (T*)vcache, lengths,
                                       (const T*)q_norm_w, (const T*)k_norm_w, (int)L, Hq, Hkv, D, Tcap,
                                       log2f(rope_base), rope_scale, norm_eps)));
  }
  B200P_LAUNCH_CHECK("rope_kv_append_kernel");
  return B200P_OK;
}

extern "C" int b200p_silu_mul(float* gu, void* out, int64_t M, int64_t I, int64_t ldg, int64_t ldo, int dtype, int flags,
                              void* stream) {
  PdlClass pdl_cls(16);
  B200P_REQUIRE(gu && out, B200P_ERR_ARG, "null pointer");
  B200P_REQUIRE(dtype == B200P_F16 || dtype == B200P_BF16, B200P_ERR_ARG, "silu_mul writes a 16-bit activation");
  B200P_REQUIRE((flags & ~B200P_OUT_PERM8) == 0, B200P_ERR_ARG, "unknown flags 0x%x", flags);
  B200P_REQUIRE(I % 8 == 0 && ldg % 4 == 0 && ldo % 8 == 0 && ldg >= 2 * I && ldo >= I && aligned(gu, 16) && aligned(out, 16),
                B200P_ERR_ALIGN, "silu_mul needs I %% 8 == 0 and 16-byte aligned rows");
  if (M * I == 0) return B200P_OK;
  const int64_t n = M * (I / 8);
  const int perm = (flags & B200P_OUT_PERM8) ? 1 : 0;
  if (dtype == B200P_F16)
    launch_pdl(silu_mul_kernel<__half>, (unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream, gu, (__half*)out, M, I, ldg, ldo, perm);
  else
    launch_pdl(silu_mul_kernel<__nv_bfloat16>, (unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream, gu, (__nv_bfloat16*)out, M, I, ldg, ldo, perm);
  B200P_LAUNCH_CHECK("silu_mul_kernel");
  return B200P_OK;
}

extern "C" int b200p_advance_lengths(int32_t* lengths, int64_t B, int32_t delta, void* stream) {
  PdlClass pdl_cls(32);
  B200P_REQUIRE(lengths, B200P_ERR_ARG, "null lengths");
  if (B == 0) return B200P_OK;
  launch_pdl(advance_lengths_kernel, (unsigned)((B + 255) / 256), 256, 0, (cudaStream_t)stream, lengths, B, delta);
  B200P_LAUNCH_CHECK("advance_lengths_kernel");
  return B200P_OK;
}

extern "C" int b200p_kv_overflow_count(unsigned int* count, int reset, void* stream) {
  B200P_REQUIRE(count, B200P_ERR_ARG, "null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t e = cudaMemcpyFromSymbolAsync(count, b200p::g_kv_overflow, sizeof(unsigned int), 0, cudaMemcpyDeviceToHost, st);
  if (e != cudaSuccess) return b200p::cuda_fail(e, "cudaMemcpyFromSymbolAsync(g_kv_overflow)");
  if (reset) {
    const unsigned int zero = 0;
    e = cudaMemcpyToSymbolAsync(b200p::g_kv_overflow, &zero, sizeof(unsigned int), 0, cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return b200p::cuda_fail(e, "cudaMemcpyToSymbolAsync(g_kv_overflow)");
  }
  e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) return b200p::cuda_fail(e, "cudaStreamSynchronize");
  return B200P_OK;
}
