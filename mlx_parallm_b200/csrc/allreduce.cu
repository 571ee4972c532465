// Tensor-parallel exchange over NVLink peer memory (new: the reference is single-device).
//
// b200p_allreduce_residual_rmsnorm fuses what follows every row-parallel projection (o_proj, down_proj) under TP:
//
//     h[m, :]  +=  sum over ranks p = 0..W-1 of  partial_p[m, :]          (all-reduce of the fp32 partial sums + residual add)
//     y[m, :]   =  rms_norm(h[m, :], w, eps)   in the storage type         (the NEXT RMSNorm of the graph, optional)
//     other[m, :] = 0                                                     (re-arms the buffer the next projection accumulates into)
//
// in ONE kernel per rank, instead of  memset + ncclAllReduce + rmsnorm.  It is a one-shot all-reduce: the projection's
// split-K epilogue (TMA reduce-add) accumulates straight into this rank's partial buffer, which lives in memory every
// peer has mapped (cudaIpc); after a flag handshake each rank READS all W partial buffers over NVLink/NVSwitch
// (W-1 of them remote, 16-byte ld.relaxed.sys) and sums them in rank order, so every rank obtains the bit-identical
// residual stream and no broadcast is needed.  One CTA per row: the row's RMS needs the whole row anyway, and B rows
// x W x H x 4 bytes of reads per CTA keep enough loads in flight (1 MiB per peer at B=64, H=4096).
//
// Handshake (per CTA, no grid-wide sync): CTA c of rank r bumps its epoch e, stores e into slot [c][r] of EVERY peer's flag
// array and spins until its own slots [c][p] have all reached e (relaxed system-scope accesses; see peer_handshake for
// where a fence is and is not needed -- a MEMBAR.SYS per handshake doubled the exchange time).
// A peer's CTA c signalling means the peer's kernel has started, i.e. (stream order, griddepcontrol.wait) its projection
// has finished writing its partial buffer AND its previous exchange has finished reading ours -- which is what makes the
// two-buffer rotation safe: partial buffers alternate A (o_proj) / B (down_proj); the exchange on A zeroes the local B
// behind the handshake and vice versa, and a buffer is next written only after the following exchange's handshake.
// A spin that sees no progress for ~4 s (peer died) sets status[0] = 1 and falls through instead of hanging the GPU.
// Because CTA c waits for CTA c of every peer, a grid never exceeds what the device can hold resident at once.
//
// Algorithmic bytes per launch: W * M * H * 4 read (of which (W-1)/W over NVLink) + M * H * (4 + 4 + 2) local.
#include "common.cuh"

namespace b200p {

int sm_count();

constexpr int AR_MAX_WORLD = 8;
constexpr int AR_THREADS = 512;

struct ArPeers {
  const float* partial[AR_MAX_WORLD];   // partial[p]: rank p's partial-sum buffer as mapped in THIS process
  uint32_t* flags[AR_MAX_WORLD];        // flags[p]:   rank p's flag array [max_ctas][AR_MAX_WORLD] (+ status word at the end)
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_relaxed_sys_f4(const float* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

__device__ __forceinline__ float block_sum_ar(float v, float* sh) {
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  if (lane == 0) sh[wid] = v;
  __syncthreads();
  float t = (threadIdx.x < nw) ? sh[threadIdx.x] : 0.f;
  if (wid == 0) {
    t = warp_sum(t);
    if (lane == 0) sh[32] = t;
  }
  __syncthreads();
  const float r = sh[32];
  __syncthreads();
  return r;
}

__device__ __forceinline__ uint32_t ld_relaxed_sys_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ void st_relaxed_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// cross-rank handshake of CTA `cta`: returns when every peer's CTA `cta` has entered the same exchange.
//
// Fences are the expensive part (a MEMBAR.SYS costs microseconds, the flag's flight ~1 us), so they are used only where the
// hardware needs them:
//  * RELEASE = false (start of an exchange): what the peers will read -- our partial buffer -- was written by the PREVIOUS
//    kernel, and a kernel boundary already makes its writes (TMA reduce-adds included) performed at the L2 that serves the
//    peers' NVLink reads; the flag is a plain st.relaxed.sys.
//  * RELEASE = true (after this CTA pushed rows into peer memory): one fence.acq_rel.sys by the converged warp orders the
//    whole CTA's remote stores (made visible to it by the preceding __syncthreads) before the flag.
//  * no acquire fence on the way in: everything read afterwards from peer-written memory is read with ld.relaxed.sys
//    (performed at the coherence point, never served from a stale L1 line) after the __syncthreads below, or by a later kernel.
template <bool RELEASE>
__device__ __forceinline__ void peer_handshake(const ArPeers& pe, int W, int rank, int cta, uint32_t* status) {
  if (threadIdx.x < 32) {
    const int t = threadIdx.x;
    uint32_t* mine = pe.flags[rank] + (size_t)cta * AR_MAX_WORLD;
    uint32_t e = 0;
    if (t < W) e = ld_relaxed_sys_u32(mine + rank) + 1u;   // my own slot holds my last epoch
    __syncwarp();                                          // every lane has read it before lane `rank` overwrites it
    if (RELEASE) asm volatile("fence.acq_rel.sys;" ::: "memory");
    if (t < W) {
      st_relaxed_sys(pe.flags[t] + (size_t)cta * AR_MAX_WORLD + rank, e);
      unsigned long long t0 = 0;
      uint32_t spins = 0;
      while ((int32_t)(ld_relaxed_sys_u32(mine + t) - e) < 0) {
        if ((++spins & 0xfffu) == 0) {
          const unsigned long long now = globaltimer_ns();
          if (t0 == 0) t0 = now;
          else if (now - t0 > 4000000000ull) { atomicExch(status, 1u); break; }
        }
      }
    }
    __syncwarp();
  }
  __syncthreads();
}

// one CTA per row (grid-stride over rows); dynamic shared memory: H floats (the summed row) + 33 reduction slots
template <typename T, int W>
__global__ void __launch_bounds__(AR_THREADS)
allreduce_residual_rmsnorm_kernel(const ArPeers pe, int rank, uint32_t* status, float* __restrict__ h, const T* __restrict__ nw,
                                  T* __restrict__ y, float* __restrict__ zero_buf, int64_t M, int64_t H, int64_t ldp,
                                  float eps, int perm) {
  extern __shared__ __align__(16) float ar_sm[];
  float* row = ar_sm;
  float* sh = ar_sm + H;
  pdl_wait();                      // our partial buffer is the previous kernel's output
  peer_handshake<false>(pe, W, rank, (int)blockIdx.x, status);
  pdl_launch_dependents();
  const int64_t H4 = H / 4;
  for (int64_t m = blockIdx.x; m < M; m += gridDim.x) {
    float ss = 0.f;
    for (int64_t i = threadIdx.x; i < H4; i += AR_THREADS) {
      float4 v[W];
#pragma unroll
      for (int p = 0; p < W; ++p) v[p] = ld_relaxed_sys_f4(pe.partial[p] + m * ldp + 4 * i);
      float4 a = reinterpret_cast<const float4*>(h + m * H)[i];
#pragma unroll
      for (int p = 0; p < W; ++p) {   // rank order on every rank: bit-identical residual streams
        a.x += v[p].x; a.y += v[p].y; a.z += v[p].z; a.w += v[p].w;
      }
      reinterpret_cast<float4*>(h + m * H)[i] = a;
      reinterpret_cast<float4*>(row)[i] = a;
      ss += a.x * a.x + a.y * a.y + a.z * a.z + a.w * a.w;
      if (zero_buf) reinterpret_cast<float4*>(zero_buf + m * ldp)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    if (y == nullptr) continue;     // uniform per launch
    ss = block_sum_ar(ss, sh);      // also orders the row[] writes before the reads below
    const float inv = 1.0f / sqrtf(ss / (float)H + eps);
    T* yr = y + m * H;
    if (perm) {
      if constexpr (sizeof(T) == 2) {
        for (int64_t i = threadIdx.x; i < H / 8; i += AR_THREADS) {
          const float4 a = reinterpret_cast<const float4*>(row)[2 * i], b = reinterpret_cast<const float4*>(row)[2 * i + 1];
          const uint4 wv = reinterpret_cast<const uint4*>(nw)[i];
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
    } else {
      for (int64_t i = threadIdx.x; i < H4; i += AR_THREADS) {
        const float4 v = reinterpret_cast<const float4*>(row)[i];
        __align__(16) T o[4];
        o[0] = Tr<T>::from_f(v.x * inv * Tr<T>::to_f(nw[4 * i + 0]));
        o[1] = Tr<T>::from_f(v.y * inv * Tr<T>::to_f(nw[4 * i + 1]));
        o[2] = Tr<T>::from_f(v.z * inv * Tr<T>::to_f(nw[4 * i + 2]));
        o[3] = Tr<T>::from_f(v.w * inv * Tr<T>::to_f(nw[4 * i + 3]));
        if (sizeof(T) == 2) reinterpret_cast<uint2*>(yr)[i] = *reinterpret_cast<uint2*>(o);
        else reinterpret_cast<uint4*>(yr)[i] = *reinterpret_cast<uint4*>(o);
      }
    }
    __syncthreads();                // row[] is rewritten by the next row of this CTA
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// Two-shot variant ("rows"): the residual stream is ROW-SHARDED across the ranks for the whole layer stack.  Rank r owns rows
// [r*rpr, (r+1)*rpr) (rpr = ceil(M / W)): it pulls only ITS rows of the W partial buffers (reduce-scatter by rows, 1/W of the
// one-shot traffic), updates its rows of h, normalises them and PUSHES the 16-bit normalised rows into every rank's x buffer
// (all-gather of the GEMM input, 16-byte remote stores), then handshakes a second time so that no rank's kernel retires
// before all rows of its x have landed.  Per rank and exchange: (W-1)/W * M*H*4 bytes in, (W-1)/W * M*H*2 bytes out --
// at W = 8, M = 64, H = 4096: 0.9 MiB + 0.45 MiB instead of 7 MiB.  h rows a rank does not own go stale; only x is global.
// One CTA per owned row (grid-stride when rpr exceeds the resident capacity); every rank launches the same grid so the
// same-index handshake stays symmetric even for ranks whose rows lie beyond M.
struct ArPeersX {
  void* x[AR_MAX_WORLD];                // x[p]: rank p's GEMM-input buffer [M, H] storage type, as mapped in this process
};

template <typename T, int W>
__global__ void __launch_bounds__(AR_THREADS)
allreduce_rows_rmsnorm_kernel(const ArPeers pe, const ArPeersX px, int rank, uint32_t* status, float* __restrict__ h,
                              const T* __restrict__ nw, float* __restrict__ zero_buf, int64_t M, int64_t H, int64_t ldp,
                              int64_t rpr, float eps, int perm) {
  extern __shared__ __align__(16) float ar_sm[];
  float* row = ar_sm;
  float* sh = ar_sm + H;
  pdl_wait();
  peer_handshake<false>(pe, W, rank, (int)blockIdx.x, status);   // every peer's partial buffer is complete
  pdl_launch_dependents();
  const int64_t H8 = H / 8;
  const int64_t row_end = min(M, (int64_t)(rank + 1) * rpr);
  for (int64_t m = (int64_t)rank * rpr + blockIdx.x; m < row_end; m += gridDim.x) {
    float ss = 0.f;
    for (int64_t i = threadIdx.x; i < H8; i += AR_THREADS) {
      float4 v0[W], v1[W];
#pragma unroll
      for (int p = 0; p < W; ++p) {
        v0[p] = ld_relaxed_sys_f4(pe.partial[p] + m * ldp + 8 * i);
        v1[p] = ld_relaxed_sys_f4(pe.partial[p] + m * ldp + 8 * i + 4);
      }
      float4 a = reinterpret_cast<const float4*>(h + m * H)[2 * i], b = reinterpret_cast<const float4*>(h + m * H)[2 * i + 1];
#pragma unroll
      for (int p = 0; p < W; ++p) {   // rank order: the same sum whoever owns the row
        a.x += v0[p].x; a.y += v0[p].y; a.z += v0[p].z; a.w += v0[p].w;
        b.x += v1[p].x; b.y += v1[p].y; b.z += v1[p].z; b.w += v1[p].w;
      }
      reinterpret_cast<float4*>(h + m * H)[2 * i] = a;
      reinterpret_cast<float4*>(h + m * H)[2 * i + 1] = b;
      reinterpret_cast<float4*>(row)[2 * i] = a;
      reinterpret_cast<float4*>(row)[2 * i + 1] = b;
      ss += a.x * a.x + a.y * a.y + a.z * a.z + a.w * a.w + b.x * b.x + b.y * b.y + b.z * b.z + b.w * b.w;
    }
    ss = block_sum_ar(ss, sh);
    const float inv = 1.0f / sqrtf(ss / (float)H + eps);
    for (int64_t i = threadIdx.x; i < H8; i += AR_THREADS) {
      const float4 a = reinterpret_cast<const float4*>(row)[2 * i], b = reinterpret_cast<const float4*>(row)[2 * i + 1];
      const float xv[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
      if constexpr (sizeof(T) == 2) {
        const uint4 wv = reinterpret_cast<const uint4*>(nw)[i];
        const T* wp = reinterpret_cast<const T*>(&wv);
        __align__(16) T o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[perm ? perm8_pos(j) : j] = Tr<T>::from_f(xv[j] * inv * Tr<T>::to_f(wp[j]));
        const uint4 ov = *reinterpret_cast<uint4*>(o);
#pragma unroll
        for (int p = 0; p < W; ++p) reinterpret_cast<uint4*>(reinterpret_cast<T*>(px.x[p]) + m * H)[i] = ov;
      } else {
        float4 o0, o1;
        o0.x = xv[0] * inv * nw[8 * i + 0]; o0.y = xv[1] * inv * nw[8 * i + 1]; o0.z = xv[2] * inv * nw[8 * i + 2];
        o0.w = xv[3] * inv * nw[8 * i + 3]; o1.x = xv[4] * inv * nw[8 * i + 4]; o1.y = xv[5] * inv * nw[8 * i + 5];
        o1.z = xv[6] * inv * nw[8 * i + 6]; o1.w = xv[7] * inv * nw[8 * i + 7];
#pragma unroll
        for (int p = 0; p < W; ++p) {
          float4* dst = reinterpret_cast<float4*>(reinterpret_cast<float*>(px.x[p]) + m * H);
          dst[2 * i] = o0;
          dst[2 * i + 1] = o1;
        }
      }
    }
    __syncthreads();                // row[] is rewritten by the next row of this CTA
  }
  // re-arm the other partial buffer: rows of every rank's slice that this CTA index covers (all peers have passed the first
  // handshake of THIS exchange, so their previous exchange has finished reading it)
  if (zero_buf) {
    for (int q = 0; q < W; ++q)
      for (int64_t m = (int64_t)q * rpr + blockIdx.x; m < min(M, (int64_t)(q + 1) * rpr); m += gridDim.x)
        for (int64_t i = threadIdx.x; i < H / 4; i += AR_THREADS)
          reinterpret_cast<float4*>(zero_buf + m * ldp)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  __syncthreads();                  // all of this CTA's remote stores are issued ...
  peer_handshake<true>(pe, W, rank, (int)blockIdx.x, status);    // ... released by one system fence; on return every peer's CTA of this index has pushed its rows to us
}

// All-gather of small per-rank blocks through peer memory (the vocab-parallel greedy head gathers the W per-rank records
// {max, sum exp, argmax, -} of every row with it): dst[p * n16 + i] = peer p's src[i].  One CTA; flag slot `cta_slot`.
__global__ void __launch_bounds__(256)
peer_allgather_kernel(const ArPeers pe, int W, int rank, int cta_slot, uint32_t* status, float4* __restrict__ dst, int64_t n16) {
  pdl_wait();
  peer_handshake<false>(pe, W, rank, cta_slot, status);
  for (int64_t i = threadIdx.x; i < n16 * W; i += blockDim.x) {
    const int p = (int)(i / n16);
    dst[i] = ld_relaxed_sys_f4(pe.partial[p] + 4 * (i - (int64_t)p * n16));
  }
}

}  // namespace b200p

using namespace b200p;

// ---- peer-memory setup (host side; NOT stream-ordered, NOT graph-capturable: call once at attach time) --------------------
extern "C" int b200p_symm_alloc(void** ptr, size_t bytes) {
  B200P_REQUIRE(ptr && bytes > 0, B200P_ERR_ARG, "bad symm_alloc arguments");
  cudaError_t e = cudaMalloc(ptr, bytes);
  if (e != cudaSuccess) return cuda_fail(e, "cudaMalloc (peer-visible buffer)");
  e = cudaMemset(*ptr, 0, bytes);
  if (e != cudaSuccess) return cuda_fail(e, "cudaMemset");
  e = cudaDeviceSynchronize();
  if (e != cudaSuccess) return cuda_fail(e, "cudaDeviceSynchronize");
  return B200P_OK;
}
extern "C" int b200p_symm_free(void* ptr) {
  if (!ptr) return B200P_OK;
  cudaError_t e = cudaFree(ptr);
  if (e != cudaSuccess) return cuda_fail(e, "cudaFree");
  return B200P_OK;
}
extern "C" int b200p_ipc_export(void* ptr, void* handle64) {
  B200P_REQUIRE(ptr && handle64, B200P_ERR_ARG, "null pointer");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
  cudaError_t e = cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t*>(handle64), ptr);
  if (e != cudaSuccess) return cuda_fail(e, "cudaIpcGetMemHandle");
  return B200P_OK;
}
extern "C" int b200p_ipc_import(const void* handle64, void** ptr) {
  B200P_REQUIRE(ptr && handle64, B200P_ERR_ARG, "null pointer");
  cudaIpcMemHandle_t hdl;
  memcpy(&hdl, handle64, sizeof(hdl));
  cudaError_t e = cudaIpcOpenMemHandle(ptr, hdl, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) return cuda_fail(e, "cudaIpcOpenMemHandle (is NVLink P2P available between the ranks' devices?)");
  return B200P_OK;
}
extern "C" int b200p_ipc_close(void* ptr) {
  if (!ptr) return B200P_OK;
  cudaError_t e = cudaIpcCloseMemHandle(ptr);
  if (e != cudaSuccess) return cuda_fail(e, "cudaIpcCloseMemHandle");
  return B200P_OK;
}

static int fill_peers(ArPeers& pe, const void* const* partial, void* const* flags, int world, int rank) {
  B200P_REQUIRE(world >= 1 && world <= AR_MAX_WORLD && rank >= 0 && rank < world, B200P_ERR_ARG,
                "world must be 1..%d and 0 <= rank < world (got world %d rank %d)", AR_MAX_WORLD, world, rank);
  B200P_REQUIRE(partial && flags, B200P_ERR_ARG, "null peer table");
  for (int p = 0; p < AR_MAX_WORLD; ++p) {
    pe.partial[p] = p < world ? reinterpret_cast<const float*>(partial[p]) : nullptr;
    pe.flags[p] = p < world ? reinterpret_cast<uint32_t*>(flags[p]) : nullptr;
    B200P_REQUIRE(p >= world || (pe.partial[p] && pe.flags[p] && aligned(pe.partial[p], 16)), B200P_ERR_ALIGN,
                  "peer %d: buffers must be non-null and 16-byte aligned", p);
  }
  return B200P_OK;
}

static int g_ar_max_ctas = 0;     // 0 = auto (co-resident capacity of the device)
// Programmatic dependent launch of the exchange kernel (B200P_AR_PDL=1).  Off by default: measured on 2 x B200 it makes the
// decode step SLOWER (3.43 -> 3.70 ms at B=64, T=256) -- the early-launched exchange CTAs spin in the handshake and keep the
// next projection's persistent one-CTA-per-SM GEMM off their SMs.
static int g_ar_pdl = [] { const char* e = getenv("B200P_AR_PDL"); return e ? atoi(e) : 0; }();
extern "C" int b200p_allreduce_set_max_ctas(int n) {
  B200P_REQUIRE(n >= 0 && n < B200P_AR_MAX_CTAS, B200P_ERR_ARG, "max ctas must be in [0, %d)", B200P_AR_MAX_CTAS);
  g_ar_max_ctas = n;
  return B200P_OK;
}

extern "C" size_t b200p_allreduce_flag_bytes(void) { return (size_t)(B200P_AR_MAX_CTAS + 1) * AR_MAX_WORLD * sizeof(uint32_t); }

extern "C" int b200p_allreduce_residual_rmsnorm(const void* const* peer_partial, void* const* peer_flags, int world, int rank,
                                                float* h, const void* norm_w, void* y, float* zero_buf, int64_t M, int64_t H,
                                                int64_t ld_partial, float eps, int dtype, int flags, void* stream) {
  ArPeers pe;
  int rc = fill_peers(pe, peer_partial, peer_flags, world, rank);
  if (rc) return rc;
  B200P_REQUIRE(h && aligned(h, 16) && H % 4 == 0 && ld_partial % 4 == 0 && ld_partial >= H, B200P_ERR_ALIGN,
                "h must be 16-byte aligned, H and ld_partial multiples of 4, ld_partial >= H");
  B200P_REQUIRE((y == nullptr) == (norm_w == nullptr), B200P_ERR_ARG, "y and norm_w go together");
  B200P_REQUIRE((flags & ~B200P_OUT_PERM8) == 0, B200P_ERR_ARG, "unknown flags 0x%x", flags);
  const int perm = (flags & B200P_OUT_PERM8) ? 1 : 0;
  B200P_REQUIRE(!perm || (y && dtype != B200P_F32 && H % 8 == 0 && aligned(norm_w, 16)), B200P_ERR_ARG,
                "B200P_OUT_PERM8 needs y, a 16-bit dtype, H %% 8 == 0 and a 16-byte aligned weight");
  B200P_REQUIRE(!y || aligned(y, 16), B200P_ERR_ALIGN, "y must be 16-byte aligned");
  B200P_REQUIRE(!zero_buf || aligned(zero_buf, 16), B200P_ERR_ALIGN, "zero_buf must be 16-byte aligned");
  B200P_REQUIRE(world == 1 || world == 2 || world == 4 || world == 8, B200P_ERR_ARG, "world must be 1, 2, 4 or 8");
  if (M == 0) return B200P_OK;     // every rank skips together: the handshake stays matched
  const size_t smem = (size_t)(H + 40) * sizeof(float);
  int64_t cap = B200P_AR_MAX_CTAS - 1;
  uint32_t* status = pe.flags[rank] + (size_t)B200P_AR_MAX_CTAS * AR_MAX_WORLD;
  cudaStream_t st = (cudaStream_t)stream;
#define AR_LAUNCH(T, W_)                                                                                             \
  do {                                                                                                               \
    if (smem > 48 * 1024) {                                                                                          \
      cudaError_t e_ = cudaFuncSetAttribute(allreduce_residual_rmsnorm_kernel<T, W_>,                                \
                                            cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);                 \
      if (e_ != cudaSuccess) return cuda_fail(e_, "cudaFuncSetAttribute(allreduce)");                                \
    }                                                                                                                \
    /* CTA c waits for CTA c of every peer: all CTAs of the grid must be able to be resident at once */             \
    int occ_ = 0;                                                                                                    \
    cudaError_t eo_ = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_, allreduce_residual_rmsnorm_kernel<T, W_>, \
                                                                    AR_THREADS, smem);                               \
    if (eo_ != cudaSuccess || occ_ < 1) return cuda_fail(eo_, "occupancy query (allreduce): row too long for shared memory?"); \
    if ((int64_t)occ_ * sm_count() < cap) cap = (int64_t)occ_ * sm_count();                                          \
    if (g_ar_max_ctas > 0 && g_ar_max_ctas < cap) cap = g_ar_max_ctas;                                               \
    const unsigned grid = (unsigned)(M < cap ? M : cap);                                                             \
    launch_pdl_opt(g_ar_pdl != 0, allreduce_residual_rmsnorm_kernel<T, W_>, grid, AR_THREADS, smem, st, pe, rank, status, h, \
               (const T*)norm_w, (T*)y, zero_buf, M, H, ld_partial, eps, perm);                                      \
  } while (0)
#define AR_DISPATCH_W(T)                                                                          \
  switch (world) {                                                                                \
    case 1: AR_LAUNCH(T, 1); break;                                                               \
    case 2: AR_LAUNCH(T, 2); break;                                                               \
    case 4: AR_LAUNCH(T, 4); break;                                                               \
    default: AR_LAUNCH(T, 8); break;                                                              \
  }
  B200P_DISPATCH_DTYPE(dtype, T, AR_DISPATCH_W(T));
#undef AR_DISPATCH_W
#undef AR_LAUNCH
  B200P_LAUNCH_CHECK("allreduce_residual_rmsnorm_kernel");
  return B200P_OK;
}

extern "C" int b200p_allreduce_rows_rmsnorm(const void* const* peer_partial, void* const* peer_x, void* const* peer_flags,
                                            int world, int rank, float* h, const void* norm_w, float* zero_buf, int64_t M,
                                            int64_t H, int64_t ld_partial, float eps, int dtype, int flags, void* stream) {
  ArPeers pe;
  int rc = fill_peers(pe, peer_partial, peer_flags, world, rank);
  if (rc) return rc;
  B200P_REQUIRE(peer_x && norm_w && h, B200P_ERR_ARG, "null pointer");
  ArPeersX px;
  for (int p = 0; p < AR_MAX_WORLD; ++p) {
    px.x[p] = p < world ? peer_x[p] : nullptr;
    B200P_REQUIRE(p >= world || (px.x[p] && aligned(px.x[p], 16)), B200P_ERR_ALIGN, "peer %d: x must be non-null and 16-byte aligned", p);
  }
  B200P_REQUIRE(aligned(h, 16) && H % 8 == 0 && ld_partial % 4 == 0 && ld_partial >= H && aligned(norm_w, 16), B200P_ERR_ALIGN,
                "h / norm_w must be 16-byte aligned, H a multiple of 8, ld_partial a multiple of 4 and >= H");
  B200P_REQUIRE((flags & ~B200P_OUT_PERM8) == 0, B200P_ERR_ARG, "unknown flags 0x%x", flags);
  const int perm = (flags & B200P_OUT_PERM8) ? 1 : 0;
  B200P_REQUIRE(!perm || dtype != B200P_F32, B200P_ERR_ARG, "B200P_OUT_PERM8 is a 16-bit layout");
  B200P_REQUIRE(!zero_buf || aligned(zero_buf, 16), B200P_ERR_ALIGN, "zero_buf must be 16-byte aligned");
  B200P_REQUIRE(world == 1 || world == 2 || world == 4 || world == 8, B200P_ERR_ARG, "world must be 1, 2, 4 or 8");
  if (M == 0) return B200P_OK;
  const int64_t rpr = (M + world - 1) / world;
  const size_t smem = (size_t)(H + 40) * sizeof(float);
  int64_t cap = B200P_AR_MAX_CTAS - 1;
  uint32_t* status = pe.flags[rank] + (size_t)B200P_AR_MAX_CTAS * AR_MAX_WORLD;
  cudaStream_t st = (cudaStream_t)stream;
#define AR2_LAUNCH(T, W_)                                                                                            \
  do {                                                                                                               \
    if (smem > 48 * 1024) {                                                                                          \
      cudaError_t e_ = cudaFuncSetAttribute(allreduce_rows_rmsnorm_kernel<T, W_>,                                    \
                                            cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);                 \
      if (e_ != cudaSuccess) return cuda_fail(e_, "cudaFuncSetAttribute(allreduce rows)");                           \
    }                                                                                                                \
    int occ_ = 0;                                                                                                    \
    cudaError_t eo_ = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_, allreduce_rows_rmsnorm_kernel<T, W_>,     \
                                                                    AR_THREADS, smem);                               \
    if (eo_ != cudaSuccess || occ_ < 1) return cuda_fail(eo_, "occupancy query (allreduce rows)");                   \
    if ((int64_t)occ_ * sm_count() < cap) cap = (int64_t)occ_ * sm_count();                                          \
    if (g_ar_max_ctas > 0 && g_ar_max_ctas < cap) cap = g_ar_max_ctas;                                               \
    const unsigned grid = (unsigned)(rpr < cap ? rpr : cap);                                                         \
    launch_pdl_opt(g_ar_pdl != 0, allreduce_rows_rmsnorm_kernel<T, W_>, grid, AR_THREADS, smem, st, pe, px, rank,    \
                   status, h, (const T*)norm_w, zero_buf, M, H, ld_partial, rpr, eps, perm);                         \
  } while (0)
#define AR2_DISPATCH_W(T)                                                                         \
  switch (world) {                                                                                \
    case 1: AR2_LAUNCH(T, 1); break;                                                              \
    case 2: AR2_LAUNCH(T, 2); break;                                                              \
    case 4: AR2_LAUNCH(T, 4); break;                                                              \
    default: AR2_LAUNCH(T, 8); break;                                                             \
  }
  B200P_DISPATCH_DTYPE(dtype, T, AR2_DISPATCH_W(T));
#undef AR2_DISPATCH_W
#undef AR2_LAUNCH
  B200P_LAUNCH_CHECK("allreduce_rows_rmsnorm_kernel");
  return B200P_OK;
}

extern "C" int b200p_peer_allgather(const void* const* peer_src, void* const* peer_flags, int world, int rank, void* dst,
                                    int64_t bytes_per_rank, void* stream) {
  ArPeers pe;
  int rc = fill_peers(pe, peer_src, peer_flags, world, rank);
  if (rc) return rc;
  B200P_REQUIRE(dst && aligned(dst, 16) && bytes_per_rank % 16 == 0, B200P_ERR_ALIGN, "dst / bytes_per_rank must be 16-byte multiples");
  if (bytes_per_rank == 0) return B200P_OK;
  uint32_t* status = pe.flags[rank] + (size_t)B200P_AR_MAX_CTAS * AR_MAX_WORLD;
  // the last flag slot is reserved for this kernel (the all-reduce uses slots [0, B200P_AR_MAX_CTAS - 1))
  launch_pdl(peer_allgather_kernel, 1, 256, 0, (cudaStream_t)stream, pe, world, rank, B200P_AR_MAX_CTAS - 1, status,
             reinterpret_cast<float4*>(dst), bytes_per_rank / 16);
  B200P_LAUNCH_CHECK("peer_allgather_kernel");
  return B200P_OK;
}

extern "C" int b200p_peer_status(const void* flags_local, uint32_t* host_out, void* stream) {
  B200P_REQUIRE(flags_local && host_out, B200P_ERR_ARG, "null pointer");
  const uint32_t* status = reinterpret_cast<const uint32_t*>(flags_local) + (size_t)B200P_AR_MAX_CTAS * AR_MAX_WORLD;
  cudaError_t e = cudaMemcpyAsync(host_out, status, sizeof(uint32_t), cudaMemcpyDeviceToHost, (cudaStream_t)stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize((cudaStream_t)stream);
  if (e != cudaSuccess) return cuda_fail(e, "peer status read");
  return B200P_OK;
}
