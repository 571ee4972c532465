// PTX wrappers shared by the tcgen05 GEMM kernels (qmm_tc.cu: whole output tiles, fused register epilogues; qmm_sk.cu:
// stream-K with TMA reduce-add into fp32): mbarriers, TMA loads / stores / reduce-adds, TMEM allocation, tcgen05.mma with
// the A operand in TMEM, tcgen05.ld / .st, shared-memory matrix descriptors.
#pragma once
#include <cuda.h>

#include <stdint.h>

namespace b200p {

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
// one lane of a CONVERGED warp; unlike `lane == 0` the compiler knows the guarded code is warp-uniform, so tcgen05 /
// TMA operands go straight to uniform registers (no per-instruction uniformisation loop)
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "elect.sync _|P1, 0xffffffff;\n"
      "selp.b32 %0, 1, 0, P1;\n"
      "}\n"
      : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
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
// D[tmem] (+)= A[tmem] * B[smem desc]^T ; kind::f16 covers f16 and bf16 inputs with fp32 accumulation
__device__ __forceinline__ void umma_f16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accum) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n"
      "}\n" ::"r"(d_tmem),
      "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accum)
      : "memory");
}
// All eight MMAs of one k-block (64 k of 2 x 128 weight rows) in one block: the issuing warp is ONE serial instruction
// stream for the whole CTA (measured: ~96 SASS instructions per k-block, mostly descriptor arithmetic and R2UR moves,
// were what bounded the kernel at ~3.7 TB/s), so the per-MMA operands are formed with the fewest adds possible:
//   A slices at a, a+8, ..., a+56 TMEM columns (block 0: 0..24, block 1: 32..56), B descriptors at b, b+2, b+4, b+6
//   (16-byte units: 32 bytes = 16 k per MMA).  acc = 0 clears D on the first slice of each block.
// The two blocks' MMAs ALTERNATE: consecutive tcgen05.mma into the SAME accumulator are a dependent chain (one every
// 64 cycles at M=128, K=16 whatever N <= 128), two independent accumulators interleaved retire one MMA every 32 cycles at
// N=64 = the full rate of the pipe (scripts/probe_mma2.cu, profiles/r2_run22_probe_mma2.txt); four in a row per block, the
// order used before, averaged 51.
__device__ __forceinline__ void umma_kblock_f16_ts(uint32_t d0, uint32_t d1, uint32_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n"
      ".reg .pred p, pt;\n"
      ".reg .b32 a1, a2, a3, a4, a5, a6, a7;\n"
      ".reg .b64 b1, b2, b3;\n"
      "setp.ne.b32 p, %5, 0;\n"
      "setp.eq.b32 pt, %4, %4;\n"
      "add.u32 a1, %2, 8;\n"
      "add.u32 a2, %2, 16;\n"
      "add.u32 a3, %2, 24;\n"
      "add.u32 a4, %2, 32;\n"
      "add.u32 a5, %2, 40;\n"
      "add.u32 a6, %2, 48;\n"
      "add.u32 a7, %2, 56;\n"
      "add.u64 b1, %3, 2;\n"
      "add.u64 b2, %3, 4;\n"
      "add.u64 b3, %3, 6;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%2], %3, %4, p;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%1], [a4], %3, %4, p;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [a1], b1, %4, pt;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%1], [a5], b1, %4, pt;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [a2], b2, %4, pt;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%1], [a6], b2, %4, pt;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [a3], b3, %4, pt;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%1], [a7], b3, %4, pt;\n"
      "}\n" ::"r"(d0),
      "r"(d1), "r"(a), "l"(b), "r"(idesc), "r"(acc)
      : "memory");
}
// The four MMAs of ONE block of a k-block (A slices a, a+8, a+16, a+24; B descriptors b, b+2, b+4, b+6).  The GEMM kernels run
// TWO issuing warps, one per 128-row block: a single thread issues one tcgen05.mma (M=128, K=16, N <= 64) every 51-64 cycles
// whatever the order of its accumulators, two threads of different warps together retire one every 32 cycles = the full rate of
// the pipe at N=64 (scripts/probe_mma2.cu, profiles/r2_run22_probe_mma2.txt).
__device__ __forceinline__ void umma_half_kblock_f16_ts(uint32_t d, uint32_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n"
      ".reg .pred p, pt;\n"
      ".reg .b32 a1, a2, a3;\n"
      ".reg .b64 b1, b2, b3;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "setp.eq.b32 pt, %3, %3;\n"
      "add.u32 a1, %1, 8;\n"
      "add.u32 a2, %1, 16;\n"
      "add.u32 a3, %1, 24;\n"
      "add.u64 b1, %2, 2;\n"
      "add.u64 b2, %2, 4;\n"
      "add.u64 b3, %2, 6;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [a1], b1, %3, pt;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [a2], b2, %3, pt;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [a3], b3, %3, pt;\n"
      "}\n" ::"r"(d),
      "r"(a), "l"(b), "r"(idesc), "r"(acc)
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
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
  uint32_t r[16];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                 "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
               : "r"(taddr));
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void sts32(uint32_t addr, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory"); }
template <typename T> __device__ __forceinline__ void sts16(uint32_t addr, T v) {
  asm volatile("st.shared.u16 [%0], %1;" ::"r"(addr), "h"(*reinterpret_cast<uint16_t*>(&v)) : "memory");
}
// generic-proxy writes to shared memory -> visible to the async proxy (TMA store / reduce reads them)
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
// shared-memory tile -> global tensor (rows / columns outside the tensor are clipped by the TMA unit)
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, int c0, int c1, uint32_t src) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];" ::"l"(map), "r"(c0), "r"(c1), "r"(src)
               : "memory");
}
// the same with an element-wise atomic add performed at the destination (fp32 tensor map)
__device__ __forceinline__ void tma_reduce_add_2d(const CUtensorMap* map, int c0, int c1, uint32_t src) {
  asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%1, %2}], [%3];" ::"l"(map), "r"(c0),
               "r"(c1), "r"(src)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
// each lane writes 32 consecutive 32-bit columns of ITS TMEM lane (= 64 consecutive 16-bit k of one A row)
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,"
      "%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]),
      "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]),
      "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(addr));
  return r;
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

__device__ __forceinline__ float silu_tc(float g) { return g / (1.0f + __expf(-g)); }


// ---- host helpers shared by the two kernels' launchers (defined in qmm_tc.cu)
typedef CUresult (*TcEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                               const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                               CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
TcEncodeFn tc_get_encode();      // cuTensorMapEncodeTiled through the runtime's driver entry point (no -lcuda dependency on it)
// 2-d tiled tensor map, memoised per process: a map is a pure function of (base, dims, stride, box, type, swizzle)
int tc_encode_2d(TcEncodeFn enc, CUtensorMap* tm, CUtensorMapDataType dt, const void* base, uint64_t d0, uint64_t d1,
                 uint64_t stride_bytes, uint32_t b0, uint32_t b1, CUtensorMapSwizzle sw, const char* what);

}  // namespace b200p
