// Register-level dequantisation of MLX affine 4-/8-bit codes to packed 16-bit pairs (shared by the tcgen05 GEMM and the
// mma.sync GEMV).
#pragma once
#include <type_traits>

#include "common.cuh"

namespace b200p {

__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
  uint32_t r;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
  return r;
}
__device__ __forceinline__ uint32_t lop3_and_or(uint32_t a, uint32_t mask, uint32_t orv) {
  uint32_t r;
  asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(r) : "r"(a), "r"(mask), "r"(orv));   // (a & mask) | orv
  return r;
}

// ------------------------------------------------------------------------------------------------ dequant
// Integer code -> 16-bit float through the exponent-magic trick (no I2F): OR the code into the mantissa of a
// power of two, subtract the power of two (exact), then ONE fused q*s+b per pair (half2 / bf162 FMA).
template <typename T> struct H2;
template <> struct H2<__half> {
  typedef __half2 V;
  static constexpr uint32_t MAGIC = 0x64006400u;        // 1024.0 | q        -> 1024 + q   (q < 1024)
  static constexpr uint32_t MAGIC_BYTES = 0x64646464u;
  static __device__ __forceinline__ V as(uint32_t u) { return *reinterpret_cast<V*>(&u); }
  static __device__ __forceinline__ uint32_t bits(V v) { return *reinterpret_cast<uint32_t*>(&v); }
  static __device__ __forceinline__ V q_lo(uint32_t m) { return __hsub2(as(m), as(0x64006400u)); }              // (1024+q) - 1024
  static __device__ __forceinline__ V q_hi(uint32_t m) { return __hfma2(as(m), as(0x2C002C00u), as(0xD400D400u)); }  // (1024+16q)/16 - 64
  static __device__ __forceinline__ uint32_t fma(V q, uint32_t s2, uint32_t b2) { return bits(__hfma2(q, as(s2), as(b2))); }
};
template <> struct H2<__nv_bfloat16> {
  typedef __nv_bfloat162 V;
  static constexpr uint32_t MAGIC = 0x43004300u;        // 128.0 | q         -> 128 + q    (q < 128)
  static constexpr uint32_t MAGIC_BYTES = 0x43434343u;
  static __device__ __forceinline__ V as(uint32_t u) { return *reinterpret_cast<V*>(&u); }
  static __device__ __forceinline__ uint32_t bits(V v) { return *reinterpret_cast<uint32_t*>(&v); }
  static __device__ __forceinline__ V q_lo(uint32_t m) { return __hsub2(as(m), as(0x43004300u)); }
  static __device__ __forceinline__ uint32_t fma(V q, uint32_t s2, uint32_t b2) { return bits(__hfma2(q, as(s2), as(b2))); }
};

// 8 consecutive codes (one word at 4 bits, two words at 8 bits) -> 8 dequantised 16-bit values, k order,
// as four packed pairs (the layout of four consecutive TMEM columns of an A row).
// With XPERM (4 bits) the pairs are emitted as they fall out of the masks -- k order [0,4,1,5,2,6,3,7] -- and the
// activations carry the same permutation (B200P_X_PERM8).
template <typename T, int BITS, bool XPERM>
__device__ __forceinline__ void dequant8(const uint32_t* pw, uint32_t s2, uint32_t b2, uint32_t* out) {
  typedef H2<T> H;
  if constexpr (BITS == 4) {
    const uint32_t w = pw[0];
    uint32_t a04, a15, a26, a37;
    if constexpr (std::is_same<T, __half>::value) {
      const uint32_t t = w >> 8;
      a04 = H::fma(H::q_lo(lop3_and_or(w, 0x000f000fu, H::MAGIC)), s2, b2);     // codes 0 | 4
      a15 = H::fma(H::q_hi(lop3_and_or(w, 0x00f000f0u, H::MAGIC)), s2, b2);     // codes 1 | 5 (x16 in the mantissa)
      a26 = H::fma(H::q_lo(lop3_and_or(t, 0x000f000fu, H::MAGIC)), s2, b2);     // codes 2 | 6
      a37 = H::fma(H::q_hi(lop3_and_or(t, 0x00f000f0u, H::MAGIC)), s2, b2);     // codes 3 | 7
    } else {
      a04 = H::fma(H::q_lo(lop3_and_or(w, 0x000f000fu, H::MAGIC)), s2, b2);
      a15 = H::fma(H::q_lo(lop3_and_or(w >> 4, 0x000f000fu, H::MAGIC)), s2, b2);
      a26 = H::fma(H::q_lo(lop3_and_or(w >> 8, 0x000f000fu, H::MAGIC)), s2, b2);
      a37 = H::fma(H::q_lo(lop3_and_or(w >> 12, 0x000f000fu, H::MAGIC)), s2, b2);
    }
    if constexpr (XPERM) {
      out[0] = a04; out[1] = a15; out[2] = a26; out[3] = a37;
    } else {
      out[0] = prmt(a04, a15, 0x5410u);   // k0 k1
      out[1] = prmt(a26, a37, 0x5410u);   // k2 k3
      out[2] = prmt(a04, a15, 0x7632u);   // k4 k5
      out[3] = prmt(a26, a37, 0x7632u);   // k6 k7
    }
  } else if constexpr (std::is_same<T, __half>::value) {
    out[0] = H::fma(H::q_lo(prmt(pw[0], H::MAGIC_BYTES, 0x4140u)), s2, b2);
    out[1] = H::fma(H::q_lo(prmt(pw[0], H::MAGIC_BYTES, 0x4342u)), s2, b2);
    out[2] = H::fma(H::q_lo(prmt(pw[1], H::MAGIC_BYTES, 0x4140u)), s2, b2);
    out[3] = H::fma(H::q_lo(prmt(pw[1], H::MAGIC_BYTES, 0x4342u)), s2, b2);
  } else {
    // bf16 holds 128 + q exactly only for q < 128 (7 mantissa bits), so an 8-bit code goes in as two nibbles:
    //   q*s + b = lo*s + (hi*(16 s) + b)      (16 s is exact; two bf16 FMAs, i.e. two roundings instead of one --
    // one more half-ulp of bf16 on the A operand, inside the tolerance bf16 models are held to)
    const typename H::V s16 = __hmul2(H::as(s2), H::as(0x41804180u));          // 16.0 | 16.0
    const uint32_t s16u = H::bits(s16);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const uint32_t x = prmt(pw[i >> 1], 0u, (i & 1) ? 0x4342u : 0x4140u);    // two codes, one per 16-bit lane
      const uint32_t t = H::fma(H::q_lo(lop3_and_or(x >> 4, 0x000f000fu, H::MAGIC)), s16u, b2);
      out[i] = H::fma(H::q_lo(lop3_and_or(x, 0x000f000fu, H::MAGIC)), s2, t);
    }
  }
}

// ------------------------------------------------------------------------------------------------ small-scale f16 path
// A 4-bit code sitting in mantissa bits [0,4) of a half IS the subnormal q * 2^-24 (bits [4,8): q * 2^-20; an 8-bit code in
// bits [0,8): q * 2^-24), and half2 FMAs take subnormal inputs at full rate.  So the masked word is already the operand:
// q*s + b is ONE half2 FMA per pair with the scale pre-multiplied by a power of two -- 2 ops per pair instead of 3, and
// half the load on the FMA pipe, which is the pipe the dequantisers saturate first (B300_MICROARCH: HFMA2 and LOP3 issue on
// different pipes at one warp instruction per 2 cycles each).  To keep s * 2^20 finite the operand is produced at 1/16 of
// its value, a' = (q*s + b) / 16, and the consumer multiplies the fp32 accumulator by 16 (exact).  The single rounding of
// q*s + b is the same as in the 3-op form, so results are bit-identical as long as |s| < 2^-4 (checked by the caller:
// B200P_W_SMALLSCALE) and a' stays a normal half (|a| >= 2^-10; below that the absolute error is at most 2^-21).
struct SmallScale {
  uint32_t m_lo, m_hi, b16;     // s * 2^20, s * 2^16, b / 16 (half2, both halves equal)
};
__device__ __forceinline__ SmallScale small_scale_prep(uint32_t s2, uint32_t b2) {
  typedef H2<__half> H;
  SmallScale r;
  const __half2 t = __hmul2(__hmul2(H::as(s2), H::as(0x64006400u)), H::as(0x54005400u));   // * 1024 * 64
  r.m_hi = H::bits(t);
  r.m_lo = H::bits(__hmul2(t, H::as(0x4C004C00u)));                                         // * 16
  r.b16 = H::bits(__hmul2(H::as(b2), H::as(0x2C002C00u)));                                  // / 16
  return r;
}
template <int BITS, bool XPERM>
__device__ __forceinline__ void dequant8_small(const uint32_t* pw, const SmallScale& c, uint32_t* out) {
  typedef H2<__half> H;
  if constexpr (BITS == 4) {
    const uint32_t w = pw[0], t = w >> 8;
    const uint32_t a04 = H::fma(H::as(w & 0x000f000fu), c.m_lo, c.b16);     // codes 0 | 4
    const uint32_t a15 = H::fma(H::as(w & 0x00f000f0u), c.m_hi, c.b16);     // codes 1 | 5 (x16 in the mantissa)
    const uint32_t a26 = H::fma(H::as(t & 0x000f000fu), c.m_lo, c.b16);     // codes 2 | 6
    const uint32_t a37 = H::fma(H::as(t & 0x00f000f0u), c.m_hi, c.b16);     // codes 3 | 7
    if constexpr (XPERM) {
      out[0] = a04; out[1] = a15; out[2] = a26; out[3] = a37;
    } else {
      out[0] = prmt(a04, a15, 0x5410u);
      out[1] = prmt(a26, a37, 0x5410u);
      out[2] = prmt(a04, a15, 0x7632u);
      out[3] = prmt(a26, a37, 0x7632u);
    }
  } else {
    out[0] = H::fma(H::as(prmt(pw[0], 0u, 0x4140u)), c.m_lo, c.b16);        // bytes 0 | 1, zero-extended
    out[1] = H::fma(H::as(prmt(pw[0], 0u, 0x4342u)), c.m_lo, c.b16);
    out[2] = H::fma(H::as(prmt(pw[1], 0u, 0x4140u)), c.m_lo, c.b16);
    out[3] = H::fma(H::as(prmt(pw[1], 0u, 0x4342u)), c.m_lo, c.b16);
  }
}

template <typename T> __device__ __forceinline__ uint32_t dup16(T v) {
  uint32_t u = (uint32_t)(*reinterpret_cast<uint16_t*>(&v));
  return u | (u << 16);
}


}  // namespace b200p
