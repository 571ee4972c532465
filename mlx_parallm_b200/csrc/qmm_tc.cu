// Tensor-core path of b200p_qmm: W4A16 / W8A16 GEMM on tcgen05 with accumulators in TMEM.
//
//   Y[tok, n] = sum_k X[tok, k] * (q[n,k] * s[n, k/64] + b[n, k/64])
//
// "swap-AB": the 128 rows of a weight block are the M side of the MMA (A operand, K-major, dequantised
// into shared memory in the canonical SWIZZLE_128B layout), the tokens are the N side (B operand = X tile,
// loaded by TMA with SWIZZLE_128B), D[128 x TN] fp32 lives in TMEM.  A CTA owns a PAIR of 128-row weight
// blocks (they share the X tile; for SILU_MUL the pair is gate block p / up block p) and a K range
// (split-K); partial accumulators are merged by the last-arriving CTA of a tile (deterministic order).
//
// warp roles (320 threads):  warp 0 = TMA producer, warp 1 = TMEM owner + single-thread MMA issuer,
// warps 2..9 = dequantisers (one weight row per thread per k-block) and, after the main loop, the epilogue
// (TMEM -> registers -> epilogue -> global, lanes along n so stores are coalesced).
//
// rings:  IN  (packed weights 2x128 rows x 64 k + X tile TN x 64)  filled by TMA, freed by MMA completion
//         A   (2 x [128 x 64] 16-bit, SW128)                       filled by dequantisers, freed by MMA completion
//
// Algorithmic HBM bytes per launch: N*K*bits/8 + 2*N*(K/64)*2 (weights, scales, biases) -- X, Y and the
// split-K partials are L2 traffic.
#include <cuda.h>

#include <stdlib.h>
#include <type_traits>

#include "common.cuh"

namespace b200p {

int sm_count();

// ------------------------------------------------------------------------------------------------ PTX
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra WAIT_DONE;\n"
      "bra WAIT_LOOP;\n"
      "WAIT_DONE:\n"
      "}\n" ::"r"(bar),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(dst),
      "l"(map), "r"(c0), "r"(c1), "r"(bar)
      : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}

__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc]^T ; kind::f16 covers f16 and bf16 inputs with fp32 accumulation
__device__ __forceinline__ void umma_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accum) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accum)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float (&v)[8]) {
  uint32_t r[8];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr));
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
  uint32_t r;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
  return r;
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(addr));
  return r;
}
__device__ __forceinline__ void sts128(uint32_t addr, uint4 v) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// K-major, SWIZZLE_128B shared-memory matrix descriptor (cute::UMMA::SmemDescriptor):
//   [0,14) start>>4 | [16,30) LBO>>4 (=1, unused for swizzled K-major) | [32,46) SBO>>4 (8 rows * 128 B = 1024)
//   | [46,48) version = 1 | [61,64) layout = 2 (SWIZZLE_128B)
__device__ __forceinline__ uint64_t make_desc_sw128(uint32_t smem_addr) {
  return (uint64_t)((smem_addr >> 4) & 0x3FFFu) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}

// ------------------------------------------------------------------------------------------------ dequant
template <typename T> struct Magic;
template <> struct Magic<__half> {
  static constexpr uint32_t BYTES = 0x64646464u;      // 0x64nn = 1024 + nn
  static __device__ __forceinline__ uint32_t fix(uint32_t h2, uint32_t s2, uint32_t b2) {
    __half2 q = __hsub2(*reinterpret_cast<__half2*>(&h2), __half2half2(__ushort_as_half(0x6400)));
    __half2 r = __hfma2(q, *reinterpret_cast<__half2*>(&s2), *reinterpret_cast<__half2*>(&b2));
    return *reinterpret_cast<uint32_t*>(&r);
  }
};
template <> struct Magic<__nv_bfloat16> {
  static constexpr uint32_t BYTES = 0x43434343u;      // 0x43nn = 128 + nn (nn < 128)
  static __device__ __forceinline__ uint32_t fix(uint32_t h2, uint32_t s2, uint32_t b2) {
    __nv_bfloat162 q = __hsub2(*reinterpret_cast<__nv_bfloat162*>(&h2), __bfloat162bfloat162(__ushort_as_bfloat16(0x4300)));
    __nv_bfloat162 r = __hfma2(q, *reinterpret_cast<__nv_bfloat162*>(&s2), *reinterpret_cast<__nv_bfloat162*>(&b2));
    return *reinterpret_cast<uint32_t*>(&r);
  }
};

// 8 consecutive 4-bit codes of one word -> 8 dequantised 16-bit values in k order
template <typename T>
__device__ __forceinline__ uint4 dequant_word4(uint32_t w, uint32_t s2, uint32_t b2) {
  uint32_t e = w & 0x0f0f0f0fu;          // bytes n0 n2 n4 n6
  uint32_t o = (w >> 4) & 0x0f0f0f0fu;   // bytes n1 n3 n5 n7
  uint32_t lo = prmt(e, o, 0x5140u);     // n0 n1 n2 n3
  uint32_t hi = prmt(e, o, 0x7362u);     // n4 n5 n6 n7
  uint4 r;
  r.x = Magic<T>::fix(prmt(lo, Magic<T>::BYTES, 0x4140u), s2, b2);
  r.y = Magic<T>::fix(prmt(lo, Magic<T>::BYTES, 0x4342u), s2, b2);
  r.z = Magic<T>::fix(prmt(hi, Magic<T>::BYTES, 0x4140u), s2, b2);
  r.w = Magic<T>::fix(prmt(hi, Magic<T>::BYTES, 0x4342u), s2, b2);
  return r;
}
// 8 consecutive 8-bit codes (two words) -> 8 dequantised values (f16 only: 1024 + q is exact for q < 1024)
template <typename T>
__device__ __forceinline__ uint4 dequant_word8(uint32_t w0, uint32_t w1, uint32_t s2, uint32_t b2) {
  uint4 r;
  r.x = Magic<T>::fix(prmt(w0, Magic<T>::BYTES, 0x4140u), s2, b2);
  r.y = Magic<T>::fix(prmt(w0, Magic<T>::BYTES, 0x4342u), s2, b2);
  r.z = Magic<T>::fix(prmt(w1, Magic<T>::BYTES, 0x4140u), s2, b2);
  r.w = Magic<T>::fix(prmt(w1, Magic<T>::BYTES, 0x4342u), s2, b2);
  return r;
}

template <typename T> __device__ __forceinline__ uint32_t dup16(T v) {
  uint32_t u = (uint32_t)(*reinterpret_cast<uint16_t*>(&v));
  return u | (u << 16);
}

__device__ __forceinline__ unsigned long long gtimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#define TL(slot)                                                                        \
  do {                                                                                  \
    if (p.tl != nullptr && blockIdx.y == 0 && blockIdx.z == 0 && blockIdx.x < 16)       \
      p.tl[blockIdx.x * 16 + (slot)] = gtimer();                                        \
  } while (0)

#define TC(i, k)                                                                                      \
  do {                                                                                                \
    if (p.tl != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && (i) >= 16 && (i) < 20) \
      p.tl[256 + ((i) - 16) * 16 + (k)] = (unsigned long long)clock64();                               \
  } while (0)

__device__ __forceinline__ float silu_tc(float g) { return g / (1.0f + __expf(-g)); }

// ------------------------------------------------------------------------------------------------ kernel
constexpr int TC_THREADS = 320;
constexpr int DQ_THREADS = 256;
constexpr int A_STAGES = 3;
constexpr int A_STAGE_BYTES = 2 * 128 * 128;  // two [128 x 64] 16-bit tiles

struct TcParams {
  const void* scales;
  const void* biases;
  void* y;
  float* ws_partial;
  int* ws_counter;
  int M_tok, N, K, TN;
  int64_t ldy;
  int epi;
  int num_kb;      // K / 64
  int splits;
  int kb_unit;     // 8 when scales/biases rows allow 128-bit loads (num_kb % 8 == 0), else 1
  int in_stages;
  int in_stage_bytes;
  int tmem_cols;
  unsigned long long* tl;   // optional per-CTA timeline (ns stamps), profiling only
  int debug;        // bit0: skip dequant math+stores, bit1: skip MMA issue, bit2: skip epilogue stores (profiling only)
};

template <typename T, int BITS>
__global__ void __launch_bounds__(TC_THREADS, 1)
qmm_tc_kernel(const __grid_constant__ CUtensorMap tm_w, const __grid_constant__ CUtensorMap tm_x, const TcParams p) {
  constexpr int PKB = 64 * BITS / 8;            // packed bytes per row per k-block
  constexpr int W_STAGE_BYTES = 2 * 128 * PKB;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // carve: [A ring][IN ring][barriers]
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t a_ring = smem_base;
  const uint32_t in_ring = a_ring + A_STAGES * A_STAGE_BYTES;
  const uint32_t bar_base = in_ring + p.in_stages * p.in_stage_bytes;
  // barrier layout (8 bytes each)
  auto full_in = [&](int s) { return bar_base + 8u * s; };
  auto empty_in = [&](int s) { return bar_base + 8u * (p.in_stages + s); };
  auto full_a = [&](int s) { return bar_base + 8u * (2 * p.in_stages + s); };
  auto empty_a = [&](int s) { return bar_base + 8u * (2 * p.in_stages + A_STAGES + s); };
  const uint32_t accum_full = bar_base + 8u * (2 * p.in_stages + 2 * A_STAGES);
  const uint32_t tmem_slot = accum_full + 8u;
  const uint32_t flag_slot = tmem_slot + 4u;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int pair = blockIdx.x, tile = blockIdx.y, split = blockIdx.z;
  const int TN = p.TN;
  const bool silu = p.epi == B200P_EPI_SILU_MUL;
  const int rowA0 = silu ? pair * 128 : pair * 256;
  const int rowB0 = silu ? p.N / 2 + pair * 128 : pair * 256 + 128;
  const int tok0 = tile * TN;
  // K range of this split
  const int n_units = p.num_kb / p.kb_unit;           // split boundaries fall on kb_unit multiples
  const int kb_begin = p.kb_unit * (int)(((int64_t)n_units * split) / p.splits);
  const int kb_end = (split == p.splits - 1) ? p.num_kb : p.kb_unit * (int)(((int64_t)n_units * (split + 1)) / p.splits);
  const int nkb = kb_end - kb_begin;

  if (threadIdx.x == 0) TL(0);
  if (threadIdx.x == 0) {
    for (int s = 0; s < p.in_stages; ++s) { mbar_init(full_in(s), 1); mbar_init(empty_in(s), 1); }
    for (int s = 0; s < A_STAGES; ++s) { mbar_init(full_a(s), DQ_THREADS / 32); mbar_init(empty_a(s), 1); }
    mbar_init(accum_full, 1);
    fence_barrier_init();
    tma_prefetch_desc(&tm_w);
    tma_prefetch_desc(&tm_x);
  }
  if (warp == 1) tmem_alloc(tmem_slot, (uint32_t)p.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  pdl_wait();   // X (and the residual) come from the previous kernel
  if (threadIdx.x == 0) TL(1);

  if (warp == 0) {
    // ===================================== TMA producer =====================================
    if (lane == 0) {
      const bool do_w = !(p.debug & 8), do_x = !(p.debug & 16);
      const uint32_t tx = (uint32_t)((do_w ? W_STAGE_BYTES : 0) + (do_x ? TN * 128 : 0));
      for (int i = 0; i < nkb; ++i) {
        const int s = i % p.in_stages, ph = (i / p.in_stages) & 1;
        TC(i, 12);
        mbar_wait(empty_in(s), ph ^ 1);
        TC(i, 13);
        const uint32_t dst = in_ring + s * p.in_stage_bytes;
        if (tx) mbar_arrive_expect_tx(full_in(s), tx); else mbar_arrive(full_in(s));
        const int kb = kb_begin + i;
        if (do_w) {
          tma_load_2d(dst, &tm_w, kb * PKB, rowA0, full_in(s));
          tma_load_2d(dst + 128 * PKB, &tm_w, kb * PKB, rowB0, full_in(s));
        }
        if (do_x) tma_load_2d(dst + W_STAGE_BYTES, &tm_x, kb * 64, tok0, full_in(s));
        if (i == 0) TL(2);
        if (i == nkb - 1) TL(3);
      }
    }
  } else if (warp == 1) {
    // ===================================== MMA issuer =======================================
    // instruction descriptor (cute::UMMA::InstrDescriptor): c_format F32 (1<<4), a/b format (0 f16, 1 bf16) at
    // bits 7/10, K-major A and B, N>>3 at bit 17, M>>4 at bit 24
    const uint32_t fmt = (sizeof(T) == 2 && std::is_same<T, __nv_bfloat16>::value) ? 1u : 0u;
    const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(TN >> 3) << 17) | ((128u >> 4) << 24);
    for (int i = 0; i < nkb; ++i) {
      const int si = i % p.in_stages, phi = (i / p.in_stages) & 1;
      const int sa = i % A_STAGES, pha = (i / A_STAGES) & 1;
      if (lane == 0) TC(i, 8);
      mbar_wait(full_in(si), phi);
      if (lane == 0) TC(i, 9);
      mbar_wait(full_a(sa), pha);
      tc_fence_after();
      if (lane == 0) TC(i, 10);
      if (lane == 0) {
        if (!(p.debug & 2)) {
        const uint32_t xs = in_ring + si * p.in_stage_bytes + W_STAGE_BYTES;
        const uint32_t as = a_ring + sa * A_STAGE_BYTES;
#pragma unroll
        for (int blk = 0; blk < 2; ++blk) {
#pragma unroll
          for (int k4 = 0; k4 < 4; ++k4) {
            const uint64_t ad = make_desc_sw128(as + blk * (128 * 128) + k4 * 32);
            const uint64_t bd = make_desc_sw128(xs + k4 * 32);
            umma_f16(tmem_base + (uint32_t)(blk * TN), ad, bd, idesc, (i > 0 || k4 > 0) ? 1u : 0u);
          }
        }
        }
        if (p.debug & 64) {
          mbar_arrive(empty_in(si)); mbar_arrive(empty_a(sa));
          if (i == nkb - 1) mbar_arrive(accum_full);
        } else {
        umma_commit(empty_in(si));
        umma_commit(empty_a(sa));
        if (i == nkb - 1) umma_commit(accum_full);
        }
        TC(i, 11);
      }
      __syncwarp();
      if (lane == 0 && i == 0) TL(4);
      if (lane == 0 && i == nkb - 1) TL(5);
    }
  } else {
    // ===================================== dequantisers =====================================
    const int dt = threadIdx.x - 64;                 // 0..255 : one weight row of the pair each
    const int blk = dt >> 7, r = dt & 127;
    const int grow = (blk ? rowB0 : rowA0) + r;      // global weight row
    const bool row_ok = grow < p.N;
    const T* sp = reinterpret_cast<const T*>(p.scales) + (int64_t)(row_ok ? grow : 0) * p.num_kb;
    const T* bp = reinterpret_cast<const T*>(p.biases) + (int64_t)(row_ok ? grow : 0) * p.num_kb;
    // swizzled positions
    const uint32_t a_row_off = (uint32_t)(blk * (128 * 128) + (r >> 3) * 1024 + (r & 7) * 128);
    const uint32_t a_xor = (uint32_t)(r & 7);
    // TMA SWIZZLE_32B / SWIZZLE_64B of the packed tile: 16-byte chunk index ^= address bits [7, 7+log2(PKB/16))
    const uint32_t w_sw = (BITS == 4) ? ((uint32_t)(dt >> 2) & 1u) : ((uint32_t)(dt >> 1) & 3u);

    auto process = [&](int i, uint32_t s2, uint32_t b2) {
      const int si = i % p.in_stages, phi = (i / p.in_stages) & 1;
      const int sa = i % A_STAGES, pha = (i / A_STAGES) & 1;
      if (dt == 0) TC(i, 0);
      mbar_wait(full_in(si), phi);
      if (dt == 0) TC(i, 1);
      if (dt == 0 && i == 0) TL(6);
      if (dt == 0 && i == 8) TL(12);
      if (dt == 0 && i == 16) TL(13);
      const uint32_t wsrc = in_ring + si * p.in_stage_bytes + (uint32_t)(dt * PKB);
      uint4 pk[PKB / 16];
#pragma unroll
      for (int c = 0; c < PKB / 16; ++c) pk[c] = lds128(wsrc + (((uint32_t)c ^ w_sw) << 4));
      if (dt == 0) TC(i, 2);
      mbar_wait(empty_a(sa), pha ^ 1);
      if (dt == 0) TC(i, 3);
      const uint32_t adst = a_ring + sa * A_STAGE_BYTES + a_row_off;
      const uint32_t* pw = reinterpret_cast<const uint32_t*>(pk);
      if (!(p.debug & 1)) {
#pragma unroll
        for (int c = 0; c < 8; ++c) {                // 8 chunks of 8 values = 64 k
          uint4 v = (BITS == 4) ? dequant_word4<T>(pw[c], s2, b2) : dequant_word8<T>(pw[2 * c], pw[2 * c + 1], s2, b2);
          sts128(adst + (((uint32_t)c ^ a_xor) << 4), v);
        }
      }
      if (dt == 0) TC(i, 4);
      if (!(p.debug & 32)) fence_proxy_async();      // generic-proxy stores -> visible to the tensor core (async proxy)
      __syncwarp();
      if (lane == 0) mbar_arrive(full_a(sa));
      if (dt == 0) TC(i, 5);
      if (dt == 0 && i == 0) TL(7);
      if (dt == 0 && i == nkb - 1) TL(8);
    };

    if (p.kb_unit == 8) {
      // scales / biases of 8 consecutive k-blocks per 128-bit load, prefetched one group (8 k-blocks) ahead
      const uint4 z4 = make_uint4(0u, 0u, 0u, 0u);
      const uint4* sp4 = reinterpret_cast<const uint4*>(sp + kb_begin);
      const uint4* bp4 = reinterpret_cast<const uint4*>(bp + kb_begin);
      uint4 sv_n = (row_ok && nkb > 0) ? ldg_stream(sp4) : z4;
      uint4 bv_n = (row_ok && nkb > 0) ? ldg_stream(bp4) : z4;
      for (int i0 = 0; i0 < nkb; i0 += 8) {
        const uint4 sv = sv_n, bv = bv_n;
        if (i0 + 8 < nkb && row_ok) { sv_n = ldg_stream(sp4 + (i0 >> 3) + 1); bv_n = ldg_stream(bp4 + (i0 >> 3) + 1); }
        const uint32_t sw[4] = {sv.x, sv.y, sv.z, sv.w}, bw[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          if (i0 + j < nkb) {
            const uint32_t s2 = (j & 1) ? prmt(sw[j >> 1], 0u, 0x3232u) : prmt(sw[j >> 1], 0u, 0x1010u);
            const uint32_t b2 = (j & 1) ? prmt(bw[j >> 1], 0u, 0x3232u) : prmt(bw[j >> 1], 0u, 0x1010u);
            process(i0 + j, s2, b2);
          }
        }
      }
    } else {
      const T zero = Tr<T>::from_f(0.f);
      T s_next = row_ok && nkb > 0 ? sp[kb_begin] : zero;
      T b_next = row_ok && nkb > 0 ? bp[kb_begin] : zero;
      for (int i = 0; i < nkb; ++i) {
        const uint32_t s2 = dup16(s_next), b2 = dup16(b_next);
        if (i + 1 < nkb && row_ok) { s_next = sp[kb_begin + i + 1]; b_next = bp[kb_begin + i + 1]; }
        process(i, s2, b2);
      }
    }

    // ===================================== epilogue =========================================
    mbar_wait(accum_full, 0);
    tc_fence_after();
    if (dt == 0) TL(9);
    const int q = warp & 3;                          // TMEM lane quarter this warp may read
    const int th = (warp - 2) >> 2;                  // token half
    const int er = q * 32 + lane;                    // row within the 128-row block
    const int n0 = rowA0 + er, n1 = rowB0 + er;
    const int t_lo = th * (TN / 2), t_hi = t_lo + TN / 2;
    const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16);
    const bool multi = p.splits > 1;
    float* part = nullptr;
    const int tile_id = blockIdx.y * gridDim.x + blockIdx.x;
    if (multi) part = p.ws_partial + ((int64_t)tile_id * p.splits + split) * (2 * (int64_t)TN * 128);

    auto finish = [&](int t, float a0, float a1) {   // final epilogue for token t (tile-local), rows n0 / n1
      const int tok = tok0 + t;
      if (tok >= p.M_tok) return;
      if (p.epi == B200P_EPI_STORE) {
        T* y = reinterpret_cast<T*>(p.y) + (int64_t)tok * p.ldy;
        if (n0 < p.N) y[n0] = Tr<T>::from_f(a0);
        if (n1 < p.N) y[n1] = Tr<T>::from_f(a1);
      } else if (p.epi == B200P_EPI_STORE_F32) {
        float* y = reinterpret_cast<float*>(p.y) + (int64_t)tok * p.ldy;
        if (n0 < p.N) y[n0] = a0;
        if (n1 < p.N) y[n1] = a1;
      } else if (p.epi == B200P_EPI_ADD_RESIDUAL) {
        float* y = reinterpret_cast<float*>(p.y) + (int64_t)tok * p.ldy;
        if (n0 < p.N) y[n0] += a0;
        if (n1 < p.N) y[n1] += a1;
      } else {
        T* y = reinterpret_cast<T*>(p.y) + (int64_t)tok * p.ldy;
        if (n0 < p.N / 2) y[n0] = Tr<T>::from_f(silu_tc(a0) * a1);
      }
    };

    for (int t = t_lo; t < t_hi && !(p.debug & 4); t += 8) {
      float v0[8], v1[8];
      tmem_ld8(lane_base + (uint32_t)t, v0);
      tmem_ld8(lane_base + (uint32_t)(TN + t), v1);
      tmem_ld_wait();
      if (!multi) {
#pragma unroll
        for (int j = 0; j < 8; ++j) finish(t + j, v0[j], v1[j]);
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          part[(int64_t)(t + j) * 128 + er] = v0[j];
          part[((int64_t)TN + t + j) * 128 + er] = v1[j];
        }
      }
    }
    if (multi) {
      __threadfence();
      named_bar_sync(1, DQ_THREADS);
      if (dt == 0) {
        int old = atomicAdd(p.ws_counter + tile_id, 1);
        int last = (old == p.splits - 1);
        if (last) p.ws_counter[tile_id] = 0;          // self-cleaning for the next launch
        asm volatile("st.shared.u32 [%0], %1;" ::"r"(flag_slot), "r"(last) : "memory");
      }
      named_bar_sync(1, DQ_THREADS);
      int last;
      asm volatile("ld.shared.u32 %0, [%1];" : "=r"(last) : "r"(flag_slot));
      if (last) {
        __threadfence();
        const float* base = p.ws_partial + (int64_t)tile_id * p.splits * (2 * (int64_t)TN * 128);
        for (int t = t_lo; t < t_hi; ++t) {
          if (tok0 + t >= p.M_tok) break;
          float a0 = 0.f, a1 = 0.f;
          for (int s = 0; s < p.splits; ++s) {        // fixed order -> deterministic
            const float* ps = base + (int64_t)s * (2 * (int64_t)TN * 128);
            a0 += __ldcg(ps + (int64_t)t * 128 + er);
            a1 += __ldcg(ps + ((int64_t)TN + t) * 128 + er);
          }
          finish(t, a0, a1);
        }
      }
    }
    tc_fence_before();
    if (dt == 0) TL(10);
  }
  pdl_launch_dependents();
  __syncthreads();
  if (threadIdx.x == 0) TL(11);
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, (uint32_t)p.tmem_cols);
  }
}

// ------------------------------------------------------------------------------------------------ host
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static unsigned long long* g_timeline = nullptr;
extern "C" void b200p_debug_set_timeline(void* dev_ptr) { g_timeline = (unsigned long long*)dev_ptr; }

static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  }
  return fn;
}

constexpr int SMEM_BUDGET = 227 * 1024;
constexpr int MIN_TC_M = 5;      // below this the CUDA-core GEMV wins (weights read once either way)

struct TcPlan {
  int kb_unit;
  int TN, tiles, pairs, splits, in_stages, in_stage_bytes, tmem_cols, smem_bytes;
};

static TcPlan make_plan(int64_t M, int64_t N, int64_t K, int bits, int epi) {
  TcPlan pl;
  int tn = (int)(M >= 256 ? 256 : ((M + 15) / 16) * 16);
  pl.TN = tn;
  pl.tiles = (int)((M + tn - 1) / tn);
  pl.pairs = (int)(epi == B200P_EPI_SILU_MUL ? (N / 2 + 127) / 128 : (N + 255) / 256);
  const int num_kb = (int)(K / 64);
  const int ctas = pl.pairs * pl.tiles;
  const int sms = sm_count();
  pl.kb_unit = (num_kb % 8 == 0) ? 8 : 1;
  int splits = 1;
  if (ctas < (sms * 3) / 4) {
    splits = sms / ctas;
    int max_by_k = pl.kb_unit == 8 ? num_kb / 8 : num_kb / 4;
    if (splits > max_by_k) splits = max_by_k;
    if (splits < 1) splits = 1;
  }
  pl.splits = splits;
  const int pkb = 64 * bits / 8;
  pl.in_stage_bytes = 2 * 128 * pkb + tn * 128;
  int avail = SMEM_BUDGET - 1024 /*align*/ - A_STAGES * A_STAGE_BYTES - 512 /*barriers*/;
  int st = avail / pl.in_stage_bytes;
  if (st > 12) st = 12;
  pl.in_stages = st;
  int cols = 32;
  while (cols < 2 * tn) cols <<= 1;
  pl.tmem_cols = cols;
  pl.smem_bytes = 1024 + A_STAGES * A_STAGE_BYTES + st * pl.in_stage_bytes + 512;
  return pl;
}

bool qmm_tc_supported(int64_t M, int64_t N, int64_t K, int gs, int bits, int dtype, int epi) {
  if (gs != 64) return false;
  if (dtype != B200P_F16 && dtype != B200P_BF16) return false;
  if (bits == 8 && dtype == B200P_BF16) return false;        // 128 + q is not exact in bf16 for q >= 128
  if (M < MIN_TC_M) return false;
  if (K % 64 != 0 || (K * bits / 8) % 16 != 0) return false;
  if (N % 8 != 0) return false;
  if (epi == B200P_EPI_SILU_MUL && (N / 2) % 8 != 0) return false;
  if (N >= (1ll << 31) || M >= (1ll << 31)) return false;
  TcPlan pl = make_plan(M, N, K, bits, epi);
  return pl.in_stages >= 2;
}

size_t qmm_tc_workspace_bytes(int64_t M, int64_t N, int64_t K) {
  // worst case over epilogues / bit widths for this shape: counters + split-K partials
  size_t best = 0;
  for (int epi = 0; epi < 4; epi += 3) {
    for (int bits = 4; bits <= 8; bits += 4) {
      if (M < MIN_TC_M || K % 64) continue;
      TcPlan pl = make_plan(M, N, K, bits, epi);
      size_t counters = ((size_t)pl.pairs * pl.tiles * sizeof(int) + 255) & ~(size_t)255;
      size_t partial = pl.splits > 1 ? (size_t)pl.pairs * pl.tiles * pl.splits * 2 * pl.TN * 128 * sizeof(float) : 0;
      if (counters + partial > best) best = counters + partial;
    }
  }
  return best;
}

template <typename T, int BITS>
static int launch_tc(const CUtensorMap& tmw, const CUtensorMap& tmx, const TcParams& p, const TcPlan& pl, cudaStream_t st) {
  auto kern = qmm_tc_kernel<T, BITS>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BUDGET);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(qmm_tc_kernel)");
    attr_set = true;
  }
  dim3 grid((unsigned)pl.pairs, (unsigned)pl.tiles, (unsigned)pl.splits);
  kern<<<grid, TC_THREADS, pl.smem_bytes, st>>>(tmw, tmx, p);
  B200P_LAUNCH_CHECK("qmm_tc_kernel");
  return B200P_OK;
}

int qmm_tc_launch(const void* x, const uint32_t* w, const void* s, const void* b, void* y, int64_t M, int64_t N,
                  int64_t K, int64_t ldx, int64_t ldy, int bits, int dtype, int epi, void* ws, size_t ws_bytes,
                  cudaStream_t st) {
  EncodeTiledFn enc = get_encode();
  B200P_REQUIRE(enc != nullptr, B200P_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
  TcPlan pl = make_plan(M, N, K, bits, epi);
  const int pkb = 64 * bits / 8;
  CUtensorMap tmw, tmx;
  {
    cuuint64_t dims[2] = {(cuuint64_t)(K * bits / 8), (cuuint64_t)N};
    cuuint64_t strides[1] = {(cuuint64_t)(K * bits / 8)};
    cuuint32_t box[2] = {(cuuint32_t)pkb, 128u};
    cuuint32_t estr[2] = {1u, 1u};
    CUresult r = enc(&tmw, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<uint32_t*>(w), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, bits == 4 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_64B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    B200P_REQUIRE(r == CUDA_SUCCESS, B200P_ERR_CUDA, "cuTensorMapEncodeTiled(weights) failed: %d", (int)r);
  }
  {
    cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)M};
    cuuint64_t strides[1] = {(cuuint64_t)(ldx * 2)};
    cuuint32_t box[2] = {64u, (cuuint32_t)pl.TN};
    cuuint32_t estr[2] = {1u, 1u};
    CUresult r = enc(&tmx, dtype == B200P_F16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                     const_cast<void*>(x), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    B200P_REQUIRE(r == CUDA_SUCCESS, B200P_ERR_CUDA, "cuTensorMapEncodeTiled(x) failed: %d", (int)r);
  }
  TcParams p;
  p.scales = s; p.biases = b; p.y = y;
  size_t counters = ((size_t)pl.pairs * pl.tiles * sizeof(int) + 255) & ~(size_t)255;
  p.ws_counter = reinterpret_cast<int*>(ws);
  p.ws_partial = reinterpret_cast<float*>(reinterpret_cast<char*>(ws) + counters);
  size_t need = counters + (pl.splits > 1 ? (size_t)pl.pairs * pl.tiles * pl.splits * 2 * pl.TN * 128 * sizeof(float) : 0);
  B200P_REQUIRE(ws_bytes >= need && aligned(ws, 16), B200P_ERR_ARG, "qmm workspace too small (%zu < %zu)", ws_bytes, need);
  p.M_tok = (int)M; p.N = (int)N; p.K = (int)K; p.TN = pl.TN; p.ldy = ldy; p.epi = epi;
  p.num_kb = (int)(K / 64); p.splits = pl.splits; p.kb_unit = pl.kb_unit; p.in_stages = pl.in_stages; p.in_stage_bytes = pl.in_stage_bytes;
  p.tmem_cols = pl.tmem_cols;
  p.tl = g_timeline;
  { const char* d = getenv("B200P_TC_DEBUG"); p.debug = d ? atoi(d) : 0; }
  if (dtype == B200P_F16) {
    return bits == 4 ? launch_tc<__half, 4>(tmw, tmx, p, pl, st) : launch_tc<__half, 8>(tmw, tmx, p, pl, st);
  } else {
    B200P_REQUIRE(bits == 4, B200P_ERR_ARG, "bf16 8-bit is not on the tensor path");
    return launch_tc<__nv_bfloat16, 4>(tmw, tmx, p, pl, st);
  }
}

}  // namespace b200p
