// Tensor-core path of b200p_qmm: W4A16 / W8A16 GEMM on tcgen05, A operand AND accumulators in TMEM.
//
//   Y[tok, n] = sum_k X[tok, k] * (q[n,k] * s[n, k/64] + b[n, k/64])
//
// "swap-AB": 128 rows of the weight matrix are the M side of the MMA, the tokens are the N side.
//   A  = dequantised weights [128 x 64] 16-bit, written by the dequant warps STRAIGHT INTO TMEM with tcgen05.st
//        (row m in lane m, k packed two per 32-bit column) -- no shared-memory round trip, no proxy fence;
//   B  = X tile [TN x 64] 16-bit, TMA-loaded into shared memory with SWIZZLE_128B (K-major);
//   D  = [128 x TN] fp32 in TMEM, double-buffered when TN <= 64 so the epilogue of one segment overlaps the
//        main loop of the next.
// A CTA owns a PAIR of 128-row weight blocks per output tile (they share the X tile; for SILU_MUL the pair is
// gate block p / up block p).
//
// Scheduling is persistent "stream-K": the work of a launch is the linear sequence of units
// (output tile, K stage of 128 packed bytes per row = 256 k at 4 bits); CTA c of G takes the contiguous range
// [c*U/G, (c+1)*U/G), so every SM streams the same number of weight bytes whatever N, K are.  A tile whose K
// range is shared by several CTAs is finished in two steps, with plain stores and loads only (one SM retires
// fp32 reds at ~1.3 cycles per lane: a 64 KB partial tile summed with atomics cost ~10 us, more than the main loop):
//   1. every contributor writes its fp32 partial tile to its own slot of the workspace (coalesced st.global.cg);
//   2. at the end of the kernel the contributors of a tile meet on an arrival counter (all CTAs of the grid are
//      co-resident: grid <= #SMs, one CTA per SM) and SHARE the reduction: contributor j sums token groups
//      j, j+S, ... of all S slots in CTA order and applies the fused epilogue to them.
// Sums are therefore taken in a fixed order: results are bit-reproducible run to run.
//
// warp roles (736 threads):
//   warps 0..15  dequantisers: 4 groups of 4 warps, group g handles weight block g&1 on k-blocks of parity g>>1;
//                a thread owns one weight row: LDS packed codes -> 16-bit q*s+b -> tcgen05.st into the A ring;
//   warps 16..19 epilogue: TMEM -> registers -> global (thread = output column n: every warp store is one coalesced
//                line per token), stream-K partial tiles and their final reduction;
//   warp 20      TMA producer of weights and scale/bias tiles (independent of the previous kernel: starts at once);
//   warp 21      TMA producer of X tiles (after griddepcontrol.wait);
//   warp 22      TMEM owner + MMA issuer (elect.sync, operands in uniform registers).
//
// rings:  W   (packed weights, 2 x [128 rows x 128 B] per stage) in shared memory, SWIZZLE_128B, filled by TMA and
//             handed back as soon as the dequant warps hold the codes in registers, so the HBM stream never waits
//             for dequant math or MMAs;
//         X   (one [TN x 64] activation tile per k-block, 8 slots) SWIZZLE_128B, filled by TMA, freed by MMA
//             completion (tcgen05.commit);
//         SB  (scales + biases of 8 k-blocks for both blocks, [128 rows x 16 B] tiles) filled by TMA when the
//             rows of the scale matrix are 16-byte multiples (K % 512 == 0), else the dequantisers read them
//             from global memory; freed by the dequant warps as soon as the values are in registers;
//         A   (4 slots x 2 blocks x 32 TMEM columns), filled by the dequantisers, freed by MMA completion.
//
// X layout: B200P_X_PERM8 says the 16-bit activations arrive with every aligned group of 8 k permuted to
// [0,4,1,5,2,6,3,7] -- the order in which a 4-bit word's nibbles fall out of two masks as packed pairs -- which
// removes the four PRMT per word from the dequant loop (the ALU pipe is the scarce resource there).  Our own
// producer kernels (rmsnorm, attention, the SILU_MUL epilogue) write that layout on request.
//
// Algorithmic HBM bytes per launch: N*K*bits/8 + 2*N*(K/64)*2 (weights, scales, biases) -- X, Y and the
// stream-K partials are L2 traffic.
#include <cuda.h>

#include <stdlib.h>
#include <type_traits>

#include "common.cuh"
#include "dequant.cuh"

namespace b200p {

int sm_count();
bool pdl_enabled_gemm();

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

// ------------------------------------------------------------------------------------------------ kernel
constexpr int DQ_WARPS = 16;
constexpr int DQ_THREADS = DQ_WARPS * 32;
constexpr int EPI_WARPS = 4;
constexpr int EPI_THREADS = EPI_WARPS * 32;
constexpr int WARP_W = DQ_WARPS + EPI_WARPS;     // TMA producer: weights + scale/bias tiles (never waits for the previous grid)
constexpr int WARP_X = WARP_W + 1;               // TMA producer: activation tiles (after griddepcontrol.wait)
constexpr int WARP_MMA = WARP_X + 1;
constexpr int TC_THREADS = (WARP_MMA + 1) * 32;
constexpr int TMEM_COLS = 512;
constexpr int MAX_TN = 128;
constexpr int W_TILE_BYTES = 128 * 128;          // one weight block of one stage: 128 rows x 128 packed bytes
constexpr int W_STAGE_BYTES = 2 * W_TILE_BYTES;
constexpr int NSLOT = 4;                         // A ring slots (each: 2 blocks x 32 TMEM columns)
constexpr uint32_t A_COL0 = 256;                 // TMEM columns [0,256) hold D, [256,512) the A ring
constexpr int NSB = 2;                           // scale/bias ring slots
constexpr int SB_TILE_BYTES = 128 * 16;          // 128 rows x 8 k-blocks x 2 bytes
constexpr int SB_SLOT_BYTES = 4 * SB_TILE_BYTES; // block 0 scales, block 0 biases, block 1 scales, block 1 biases
constexpr int MAX_STAGES = 8;                    // weight ring stages / X ring slots (upper bounds)
constexpr int X_RING_BYTES = 64 * 1024;
constexpr int EPI_TOK = 8;                       // tokens per epilogue step = work item of the shared split-tile reduction
// Profiling build only (-DB200P_TC_DEBUG=mask): 1 no X loads, 2 no MMA issue, 4 no epilogue, 8 no A stores, 32 no
// scale/bias loads.  The shipped kernel is compiled with 0: every branch below folds away.
#ifndef B200P_TC_DEBUG
#define B200P_TC_DEBUG 0
#endif
constexpr int kDebug = B200P_TC_DEBUG;

struct TcParams {
  const void* scales;
  const void* biases;
  void* y;
  float* ws_part;      // stream-K partial tiles: [grid][2 slots][2 blocks][TN tokens][128 rows] fp32 (no initial state)
  int* ws_counter;     // {arrived, done} per output tile (zero on entry, left zero)
  int M_tok, N, K, TN;
  int64_t ldy;
  int epi;
  int y_perm;          // 1: 16-bit outputs (STORE / SILU_MUL) are written in the PERM8 column order
  int num_kb;          // K / 64
  int num_st;          // stages along K = ceil(num_kb / KPS)
  int pairs, tiles_m;  // output tile t = pair * tiles_m + token tile
  int U;               // total units = pairs * tiles_m * num_st (the host checks U * (grid + 1) < 2^31: 32-bit unit arithmetic)
  int sb_tma;          // 1: scale/bias tiles come through TMA (num_kb % 8 == 0)
  int w_stages;        // weight ring depth
  int x_slots;         // X ring depth (4 or 8; one slot = the [TN x 64] tile of one k-block)
};

// CTA that owns linear unit x when U units are dealt to G CTAs as [c*U/G, (c+1)*U/G)
__device__ __forceinline__ int owner_of(int x, int U, int G) { return ((x + 1) * G - 1) / U; }

// FAST: the small-scale dequant of dequant.cuh (f16, caller promised B200P_W_SMALLSCALE): A and D are produced at 1/16.
template <typename T, int BITS, bool XPERM, bool FAST>
__global__ void __launch_bounds__(TC_THREADS, 1)
qmm_tc_kernel(const __grid_constant__ CUtensorMap tm_w, const __grid_constant__ CUtensorMap tm_x,
              const __grid_constant__ CUtensorMap tm_s, const __grid_constant__ CUtensorMap tm_b, const TcParams p) {
  static_assert(!FAST || std::is_same<T, __half>::value, "the subnormal-operand dequant is an f16 trick");
  constexpr int PKB = 64 * BITS / 8;            // packed bytes per row per k-block
  constexpr int KPS = 128 / PKB;                // k-blocks per stage (a stage row is one 128-byte line)
  constexpr int SPG = 8 / KPS;                  // stages per scale/bias group of 8 k-blocks
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t w_ring = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int S = p.w_stages;
  const int TN = p.TN;
  const uint32_t x_ring = w_ring + S * W_STAGE_BYTES;
  const uint32_t x_slot_bytes = (uint32_t)(TN * 128);
  const uint32_t x_mask = (uint32_t)(p.x_slots - 1);
  const int x_shift = p.x_slots == 8 ? 3 : 2;
  const uint32_t sb_ring = x_ring + p.x_slots * x_slot_bytes;
  const uint32_t bar_base = sb_ring + NSB * SB_SLOT_BYTES;
  const int NBUF = TN > 64 ? 1 : 2;              // D buffers: buffer d, block b at columns d*128 + b*64 (TN <= 64) / b*128
  auto full_w = [&](int s) { return bar_base + 8u * s; };
  auto empty_w = [&](int s) { return bar_base + 8u * (MAX_STAGES + s); };
  auto full_x = [&](int s) { return bar_base + 8u * (2 * MAX_STAGES + s); };
  auto empty_x = [&](int s) { return bar_base + 8u * (3 * MAX_STAGES + s); };
  auto full_a = [&](int s) { return bar_base + 8u * (4 * MAX_STAGES + s); };
  auto empty_a = [&](int s) { return bar_base + 8u * (4 * MAX_STAGES + NSLOT + s); };
  auto sb_full = [&](int s) { return bar_base + 8u * (4 * MAX_STAGES + 2 * NSLOT + s); };
  auto sb_empty = [&](int s) { return bar_base + 8u * (4 * MAX_STAGES + 2 * NSLOT + NSB + s); };
  auto accum_full = [&](int d) { return bar_base + 8u * (4 * MAX_STAGES + 2 * NSLOT + 2 * NSB + d); };
  auto accum_empty = [&](int d) { return bar_base + 8u * (4 * MAX_STAGES + 2 * NSLOT + 2 * NSB + 2 + d); };
  const uint32_t tmem_slot = bar_base + 8u * (4 * MAX_STAGES + 2 * NSLOT + 2 * NSB + 4);
  auto d_col = [&](int d, int b) { return (uint32_t)(TN > 64 ? b * 128 : d * 128 + b * 64); };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int G = (int)gridDim.x, cta = (int)blockIdx.x;
  const int u_begin = (int)(((long long)p.U * cta) / G), u_end = (int)(((long long)p.U * (cta + 1)) / G);
  const bool silu = p.epi == B200P_EPI_SILU_MUL;

  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) { mbar_init(full_w(s), 1); mbar_init(empty_w(s), DQ_WARPS); }
    for (int s = 0; s < p.x_slots; ++s) { mbar_init(full_x(s), 1); mbar_init(empty_x(s), 1); }
    for (int s = 0; s < NSLOT; ++s) { mbar_init(full_a(s), 8); mbar_init(empty_a(s), 1); }
    for (int s = 0; s < NSB; ++s) { mbar_init(sb_full(s), 1); mbar_init(sb_empty(s), DQ_WARPS); }
    for (int d = 0; d < 2; ++d) { mbar_init(accum_full(d), 1); mbar_init(accum_empty(d), EPI_WARPS); }
    fence_barrier_init();
    tma_prefetch_desc(&tm_w);
    tma_prefetch_desc(&tm_x);
    if (p.sb_tma) { tma_prefetch_desc(&tm_s); tma_prefetch_desc(&tm_b); }
  }
  if (warp == WARP_MMA) tmem_alloc(tmem_slot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
  pdl_launch_dependents();   // the next kernel may be scheduled as SMs free up; it waits for our completion itself

  if (warp == WARP_W) {
    // ===================================== weight producer ==================================
    // Weights and scale/bias tiles do not depend on the previous kernel: no griddepcontrol.wait here, so under
    // programmatic dependent launch the HBM stream of this kernel starts while the previous kernel drains.
    int si = 0, ph = 0, sbi = 0, sbph = 0;
    for (int u = u_begin; u < u_end; ++u) {
      const int tile = u / p.num_st, st = u % p.num_st;
      const int pair = tile / p.tiles_m;
      const int rowA0 = silu ? pair * 128 : pair * 256;
      const int rowB0 = silu ? p.N / 2 + pair * 128 : pair * 256 + 128;
      if (p.sb_tma && (u == u_begin || (st % SPG) == 0)) {     // a new group of 8 k-blocks starts with this stage
        mbar_wait(sb_empty(sbi), sbph ^ 1);
        if (elect_one()) {
          const uint32_t sdst = sb_ring + sbi * SB_SLOT_BYTES;
          const int kb0 = (st / SPG) * 8;
          if (kDebug & 32) {
            mbar_arrive(sb_full(sbi));
          } else {
            mbar_arrive_expect_tx(sb_full(sbi), (uint32_t)SB_SLOT_BYTES);
            tma_load_2d(sdst, &tm_s, kb0, rowA0, sb_full(sbi));
            tma_load_2d(sdst + SB_TILE_BYTES, &tm_b, kb0, rowA0, sb_full(sbi));
            tma_load_2d(sdst + 2 * SB_TILE_BYTES, &tm_s, kb0, rowB0, sb_full(sbi));
            tma_load_2d(sdst + 3 * SB_TILE_BYTES, &tm_b, kb0, rowB0, sb_full(sbi));
          }
        }
        __syncwarp();
        if (++sbi == NSB) { sbi = 0; sbph ^= 1; }
      }
      mbar_wait(empty_w(si), ph ^ 1);
      if (elect_one()) {
        const uint32_t dst = w_ring + si * W_STAGE_BYTES;
        mbar_arrive_expect_tx(full_w(si), (uint32_t)W_STAGE_BYTES);
        tma_load_2d(dst, &tm_w, st * 128, rowA0, full_w(si));
        tma_load_2d(dst + W_TILE_BYTES, &tm_w, st * 128, rowB0, full_w(si));
      }
      __syncwarp();
      if (++si == S) { si = 0; ph ^= 1; }
    }
  } else if (warp == WARP_X) {
    // ===================================== activation producer ==============================
    pdl_wait();                                      // X is the previous kernel's output
    uint32_t it = 0;
    for (int u = u_begin; u < u_end; ++u) {
      const int tile = u / p.num_st, st = u % p.num_st;
      const int tm = tile % p.tiles_m;
#pragma unroll
      for (int jj = 0; jj < KPS; ++jj, ++it) {
        const uint32_t xs = it & x_mask;
        mbar_wait(empty_x(xs), ((it >> x_shift) & 1u) ^ 1u);
        if (elect_one()) {
          if (kDebug & 1) {
            mbar_arrive(full_x(xs));
          } else {
            mbar_arrive_expect_tx(full_x(xs), x_slot_bytes);
            tma_load_2d(x_ring + xs * x_slot_bytes, &tm_x, (st * KPS + jj) * 64, tm * TN, full_x(xs));
          }
        }
        __syncwarp();
      }
    }
  } else if (warp == WARP_MMA) {
    // ===================================== MMA issuer =======================================
    // instruction descriptor (cute::UMMA::InstrDescriptor): c_format F32 (1<<4), a/b format (0 f16, 1 bf16) at
    // bits 7/10, K-major A and B, N>>3 at bit 17, M>>4 at bit 24
    const uint32_t fmt = std::is_same<T, __nv_bfloat16>::value ? 1u : 0u;
    const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(TN >> 3) << 17) | ((128u >> 4) << 24);
    int seg = 0;
    uint32_t it = 0;
    for (int u = u_begin; u < u_end; ++seg) {
      const int s0 = u % p.num_st;
      const int ns = min(p.num_st - s0, u_end - u);
      const int buf = NBUF == 2 ? (seg & 1) : 0;
      if (seg >= NBUF) {                               // the epilogue of the segment that last used this D buffer has drained it
        mbar_wait(accum_empty(buf), (uint32_t)(((seg / NBUF) - 1) & 1));
        tc_fence_after();
      }
      const uint32_t d0 = tmem_base + d_col(buf, 0), d1 = tmem_base + d_col(buf, 1);
      const int nkb = ns * KPS;
      for (int j = 0; j < nkb; ++j, ++it) {
        const uint32_t slot = it & (NSLOT - 1), xs = it & x_mask;
        mbar_wait(full_x(xs), (it >> x_shift) & 1u);
        mbar_wait(full_a(slot), (it >> 2) & 1u);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t xsm = x_ring + xs * x_slot_bytes;
          if (!(kDebug & 2))
#pragma unroll
          for (int blk = 0; blk < 2; ++blk) {
#pragma unroll
            for (int k4 = 0; k4 < 4; ++k4) {
              const uint32_t at = tmem_base + A_COL0 + slot * 64 + (uint32_t)(blk * 32 + k4 * 8);
              umma_f16_ts(blk ? d1 : d0, at, make_desc_sw128(xsm + k4 * 32), idesc, (j > 0 || k4 > 0) ? 1u : 0u);
            }
          }
          umma_commit(empty_a(slot));
          umma_commit(empty_x(xs));
          if (j == nkb - 1) umma_commit(accum_full(buf));
        }
        __syncwarp();
      }
      u += ns;
    }
  } else if (warp < DQ_WARPS) {
    // ===================================== dequantisers =====================================
    const int grp = warp >> 2;                         // 0..3
    const int blk = grp & 1, kpar = grp >> 1;
    const int q = warp & 3;                          // TMEM lane quarter this warp may touch
    const int r = q * 32 + lane;                     // weight row within the 128-row block
    const uint32_t a_dst = tmem_base + ((uint32_t)(q * 32) << 16) + A_COL0 + (uint32_t)(blk * 32);
    // SWIZZLE_128B tile [128 rows x 128 B]: row r at (r/8)*1024 + (r%8)*128, 16-byte chunk c at (c ^ (r%8))
    const uint32_t w_row_off = (uint32_t)(blk * W_TILE_BYTES + (r >> 3) * 1024 + (r & 7) * 128);
    const uint32_t w_xor = (uint32_t)(r & 7);
    const uint32_t sb_off = (uint32_t)(blk * 2 * SB_TILE_BYTES + r * 16);
    const uint32_t half_sel = kpar ? 0x3232u : 0x1010u;   // which 16-bit half of a scale word is this thread's k-block

    int si = 0, ph = 0, sbi = 0, sbph = 0;
    uint32_t it = (uint32_t)kpar;                    // global k-block counter of this thread (slot = it % 4)
    uint4 s8 = make_uint4(0u, 0u, 0u, 0u), b8 = s8;

    constexpr int CPK = PKB / 16;                    // 16-byte chunks of packed codes per k-block
    constexpr int KPT = KPS / 2;                     // k-blocks per thread per stage
    uint4 pk[KPT * CPK];
    // wait for the stage's weights, pull this thread's codes into registers and hand the slot straight back to the
    // producer: the HBM stream then only waits for LDS, not for the dequant math or the MMAs
    auto fetch_stage = [&]() {
      mbar_wait(full_w(si), (uint32_t)ph);
      const uint32_t wsrc = w_ring + si * W_STAGE_BYTES + w_row_off;
#pragma unroll
      for (int j2 = 0; j2 < KPT; ++j2)
#pragma unroll
        for (int c = 0; c < CPK; ++c)
          pk[j2 * CPK + c] = lds128(wsrc + ((((uint32_t)((kpar + 2 * j2) * CPK + c)) ^ w_xor) << 4));
      __syncwarp();
      if (lane == 0) mbar_arrive(empty_w(si));
      if (++si == S) { si = 0; ph ^= 1; }
    };
    // one k-block: 64 codes of row r -> 32 packed 16-bit pairs -> A ring slot
    auto kblock = [&](int j2, uint32_t s2, uint32_t b2) {
      uint32_t o[32];
      const uint32_t* pw = reinterpret_cast<const uint32_t*>(pk + j2 * CPK);
      if constexpr (FAST) {
        const SmallScale sc = small_scale_prep(s2, b2);
#pragma unroll
        for (int c = 0; c < 8; ++c) dequant8_small<BITS, XPERM>(pw + c * (BITS / 4), sc, o + 4 * c);
      } else {
#pragma unroll
        for (int c = 0; c < 8; ++c) dequant8<T, BITS, XPERM>(pw + c * (BITS / 4), s2, b2, o + 4 * c);
      }
      const uint32_t slot = it & (NSLOT - 1);
      mbar_wait(empty_a(slot), ((it >> 2) & 1u) ^ 1u);
      tc_fence_after();
      if (!(kDebug & 8)) {
        tmem_st32(a_dst + slot * 64, o);
        tmem_st_wait();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(full_a(slot));
      it += 2;
    };

    for (int u = u_begin; u < u_end;) {
      const int tile = u / p.num_st, s0 = u % p.num_st;
      const int ns = min(p.num_st - s0, u_end - u);
      if (p.sb_tma) {
        for (int st = s0; st < s0 + ns; ++st) {
          if (st == s0 || (st % SPG) == 0) {             // next group of 8 k-blocks: scales / biases of this row
            mbar_wait(sb_full(sbi), (uint32_t)sbph);
            const uint32_t src = sb_ring + sbi * SB_SLOT_BYTES + sb_off;
            s8 = lds128(src);
            b8 = lds128(src + SB_TILE_BYTES);
            __syncwarp();
            if (lane == 0) mbar_arrive(sb_empty(sbi));
            if (++sbi == NSB) { sbi = 0; sbph ^= 1; }
          }
          fetch_stage();
          if constexpr (KPS == 4) {
            // stage st holds k-blocks 4*(st&1) + {0..3} of the group: words 2*(st&1) + {0,1}
            const bool hi = (st & 1) != 0;
            const uint32_t sw0 = hi ? s8.z : s8.x, sw1 = hi ? s8.w : s8.y;
            const uint32_t bw0 = hi ? b8.z : b8.x, bw1 = hi ? b8.w : b8.y;
            kblock(0, prmt(sw0, 0u, half_sel), prmt(bw0, 0u, half_sel));
            kblock(1, prmt(sw1, 0u, half_sel), prmt(bw1, 0u, half_sel));
          } else {
            // stage st holds k-blocks 2*(st&3) + {0,1}: word st&3
            const int wsel = st & 3;
            const uint32_t sw = wsel == 0 ? s8.x : (wsel == 1 ? s8.y : (wsel == 2 ? s8.z : s8.w));
            const uint32_t bw = wsel == 0 ? b8.x : (wsel == 1 ? b8.y : (wsel == 2 ? b8.z : b8.w));
            kblock(0, prmt(sw, 0u, half_sel), prmt(bw, 0u, half_sel));
          }
        }
      } else {
        // scale rows are not 16-byte multiples (K % 512 != 0, e.g. a TP shard): read them from global memory
        const int pair = tile / p.tiles_m;
        const int grow = (blk ? (silu ? p.N / 2 + pair * 128 : pair * 256 + 128) : (silu ? pair * 128 : pair * 256)) + r;
        const bool row_ok = grow < p.N;
        const T* sp = reinterpret_cast<const T*>(p.scales) + (int64_t)(row_ok ? grow : 0) * p.num_kb;
        const T* bp = reinterpret_cast<const T*>(p.biases) + (int64_t)(row_ok ? grow : 0) * p.num_kb;
        const T zero = Tr<T>::from_f(0.f);
        for (int st = s0; st < s0 + ns; ++st) {
          T sv[KPS / 2], bv[KPS / 2];
#pragma unroll
          for (int j2 = 0; j2 < KPS / 2; ++j2) {
            const int kb = st * KPS + kpar + 2 * j2;
            const bool ok = row_ok && kb < p.num_kb;          // k-blocks past K are zero-filled by TMA: A = 0
            sv[j2] = ok ? __ldg(sp + kb) : zero;
            bv[j2] = ok ? __ldg(bp + kb) : zero;
          }
          fetch_stage();
#pragma unroll
          for (int j2 = 0; j2 < KPS / 2; ++j2) kblock(j2, dup16(sv[j2]), dup16(bv[j2]));
        }
      }
      u += ns;
    }
  } else {
    // ===================================== epilogue warps ===================================
    // Thread = one output column n (weight row r of block A and of block B), EPI_TOK tokens per step: every warp-level
    // load / store below touches one contiguous line per token (32 lanes x 4 B, or x 2 B for 16-bit outputs -- whole
    // sectors either way), so nothing is staged through shared memory and no store is atomic.
    const int e = warp - DQ_WARPS;                     // == warp % 4: the TMEM lane quarter this warp may read
    const int r = e * 32 + lane;                     // output row (weight row) within the 128-row block
    const uint32_t lane_base = tmem_base + ((uint32_t)(e * 32) << 16);
    const int et = threadIdx.x - DQ_THREADS;         // 0..127
    constexpr float DSCALE = FAST ? 16.f : 1.f;      // FAST: A (hence D) was produced at 1/16
    const int64_t slot_elems = (int64_t)2 * TN * 128;   // one partial tile: [2 blocks][TN tokens][128 rows] fp32

    // fused epilogue of EPI_TOK tokens of this thread's two output columns (a: block A, b: block B)
    auto emit = [&](const float (&a)[EPI_TOK], const float (&b)[EPI_TOK], int rowA0, int rowB0, int tokb) {
      const int n0 = rowA0 + r, n1 = rowB0 + r;
      if (p.epi == B200P_EPI_ADD_RESIDUAL) {
        // every element of y is owned by exactly one thread of the launch: plain read-modify-write
        float* y = reinterpret_cast<float*>(p.y);
        const bool ok0 = n0 < p.N, ok1 = n1 < p.N;
        float ha[EPI_TOK], hb[EPI_TOK];
#pragma unroll
        for (int j = 0; j < EPI_TOK; ++j) {
          const bool okt = tokb + j < p.M_tok;
          const float* yr = y + (int64_t)(tokb + j) * p.ldy;
          ha[j] = (ok0 && okt) ? __ldcg(yr + n0) : 0.f;
          hb[j] = (ok1 && okt) ? __ldcg(yr + n1) : 0.f;
        }
#pragma unroll
        for (int j = 0; j < EPI_TOK; ++j) {
          const bool okt = tokb + j < p.M_tok;
          float* yr = y + (int64_t)(tokb + j) * p.ldy;
          if (ok0 && okt) __stcg(yr + n0, ha[j] + a[j]);
          if (ok1 && okt) __stcg(yr + n1, hb[j] + b[j]);
        }
      } else if (p.epi == B200P_EPI_STORE_F32) {
        float* y = reinterpret_cast<float*>(p.y);
#pragma unroll
        for (int j = 0; j < EPI_TOK; ++j) {
          if (tokb + j < p.M_tok) {
            float* yr = y + (int64_t)(tokb + j) * p.ldy;
            if (n0 < p.N) __stcg(yr + n0, a[j]);
            if (n1 < p.N) __stcg(yr + n1, b[j]);
          }
        }
      } else if (p.epi == B200P_EPI_STORE) {
        T* y = reinterpret_cast<T*>(p.y);
        const int c0 = p.y_perm ? perm8_pos(n0) : n0, c1 = p.y_perm ? perm8_pos(n1) : n1;
#pragma unroll
        for (int j = 0; j < EPI_TOK; ++j) {
          if (tokb + j < p.M_tok) {
            T* yr = y + (int64_t)(tokb + j) * p.ldy;
            if (n0 < p.N) yr[c0] = Tr<T>::from_f(a[j]);
            if (n1 < p.N) yr[c1] = Tr<T>::from_f(b[j]);
          }
        }
      } else {                                         // SILU_MUL: block A = gate rows, block B = the matching up rows
        T* y = reinterpret_cast<T*>(p.y);
        const int c0 = p.y_perm ? perm8_pos(n0) : n0;
        if (n0 < p.N / 2) {
#pragma unroll
          for (int j = 0; j < EPI_TOK; ++j)
            if (tokb + j < p.M_tok) y[(int64_t)(tokb + j) * p.ldy + c0] = Tr<T>::from_f(silu_tc(a[j]) * b[j]);
        }
      }
    };

    pdl_wait();                                      // y and the workspace belong to the previous kernels until here
    int seg = 0;
    int def0 = -1, def1 = -1;                        // split tiles this CTA contributed to (its first / its last segment)
    for (int u = u_begin; u < u_end; ++seg) {
      const int tile = u / p.num_st, s0 = u % p.num_st;
      const int ns = min(p.num_st - s0, u_end - u);
      const int pair = tile / p.tiles_m, tm = tile % p.tiles_m;
      const int rowA0 = silu ? pair * 128 : pair * 256;
      const int rowB0 = silu ? p.N / 2 + pair * 128 : pair * 256 + 128;
      const int tok0 = tm * TN;
      const int buf = NBUF == 2 ? (seg & 1) : 0;
      const bool full = (s0 == 0 && ns == p.num_st);

      mbar_wait(accum_full(buf), (uint32_t)((seg / NBUF) & 1));
      tc_fence_after();
      const uint32_t dA = lane_base + d_col(buf, 0), dB = lane_base + d_col(buf, 1);
      if (!(kDebug & 4)) {
        if (full) {
          for (int t0 = 0; t0 < TN; t0 += EPI_TOK) {
            float a[EPI_TOK], b[EPI_TOK];
            tmem_ld8(dA + (uint32_t)t0, a);
            tmem_ld8(dB + (uint32_t)t0, b);
            tmem_ld_wait();
            if constexpr (FAST) {
#pragma unroll
              for (int j = 0; j < EPI_TOK; ++j) { a[j] *= DSCALE; b[j] *= DSCALE; }
            }
            emit(a, b, rowA0, rowB0, tok0 + t0);
          }
        } else {
          // partial tile: slot 1 if this CTA owns the tile's first stage (then it is this CTA's LAST segment), else slot 0
          // (its FIRST segment) -- a CTA never needs the same slot twice
          const int slot = (s0 == 0) ? 1 : 0;
          float* part = p.ws_part + ((int64_t)cta * 2 + slot) * slot_elems;
          for (int t0 = 0; t0 < TN; t0 += EPI_TOK) {
            float a[EPI_TOK], b[EPI_TOK];
            tmem_ld8(dA + (uint32_t)t0, a);
            tmem_ld8(dB + (uint32_t)t0, b);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < EPI_TOK; ++j) {
              __stcg(part + (int64_t)(t0 + j) * 128 + r, a[j] * DSCALE);
              __stcg(part + (int64_t)(TN + t0 + j) * 128 + r, b[j] * DSCALE);
            }
          }
          if (slot) def1 = tile; else def0 = tile;
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(accum_empty(buf));    // this D buffer may be overwritten by a later segment's MMAs
      u += ns;
    }

    // ---- split tiles: meet the other contributors, then reduce and finish a share of each tile
    if (def0 >= 0 || def1 >= 0) {
      __threadfence();                                 // this thread's partial stores are performed device-wide ...
      named_bar_sync(1, EPI_THREADS);                  // ... for all 128 epilogue threads before the arrival is published
      if (et == 0) {
        if (def0 >= 0) atomicAdd(p.ws_counter + 2 * def0, 1);
        if (def1 >= 0) atomicAdd(p.ws_counter + 2 * def1, 1);
      }
      for (int i = 0; i < 2; ++i) {
        const int tile = i ? def1 : def0;
        if (tile < 0) continue;
        const int c_first = owner_of(tile * p.num_st, p.U, G);
        const int c_last = owner_of(tile * p.num_st + p.num_st - 1, p.U, G);
        const int S = c_last - c_first + 1, me = cta - c_first;
        if (et == 0) {
          const volatile int* cnt = p.ws_counter + 2 * tile;
          unsigned long long t_start;
          asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_start));
          while (*cnt < S) {
            __nanosleep(40);
            unsigned long long t_now;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_now));
            if (t_now - t_start > 4000000000ull) __trap();      // a contributor never arrived: fail loudly, do not hang
          }
          __threadfence();
        }
        named_bar_sync(2, EPI_THREADS);
        const int pair = tile / p.tiles_m, tm = tile % p.tiles_m;
        const int rowA0 = silu ? pair * 128 : pair * 256;
        const int rowB0 = silu ? p.N / 2 + pair * 128 : pair * 256 + 128;
        const int tok0 = tm * TN;
        for (int t0 = me * EPI_TOK; t0 < TN; t0 += S * EPI_TOK) {
          if (tok0 + t0 >= p.M_tok) break;             // padding tokens of the last token tile
          float a[EPI_TOK], b[EPI_TOK];
#pragma unroll
          for (int j = 0; j < EPI_TOK; ++j) a[j] = b[j] = 0.f;
#pragma unroll 2
          for (int c = c_first; c <= c_last; ++c) {     // fixed order: bit-reproducible sums
            const float* part = p.ws_part + ((int64_t)c * 2 + (c == c_first ? 1 : 0)) * slot_elems;
            float va[EPI_TOK], vb[EPI_TOK];
#pragma unroll
            for (int j = 0; j < EPI_TOK; ++j) {         // 16 independent L2 loads in flight
              va[j] = __ldcg(part + (int64_t)(t0 + j) * 128 + r);
              vb[j] = __ldcg(part + (int64_t)(TN + t0 + j) * 128 + r);
            }
#pragma unroll
            for (int j = 0; j < EPI_TOK; ++j) { a[j] += va[j]; b[j] += vb[j]; }
          }
          emit(a, b, rowA0, rowB0, tok0 + t0);
        }
        // the last contributor to finish re-arms the tile's counters for the next launch (all S have passed the wait)
        if (et == 0) {
          const int old = atomicAdd(p.ws_counter + 2 * tile + 1, 1);
          if (old == S - 1) {
            p.ws_counter[2 * tile] = 0;
            p.ws_counter[2 * tile + 1] = 0;
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == WARP_MMA) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------------ host
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

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
constexpr int SMEM_FIXED = 1024 /*align*/ + NSB * SB_SLOT_BYTES + 512 /*barriers*/;
static_assert(8 * (4 * MAX_STAGES + 2 * NSLOT + 2 * NSB + 4) + 8 <= 512, "barrier block overflows its slot");
// smallest M that may take the tensor path.  1: the tile kernel beats the CUDA-core GEMV at EVERY M (8.4 vs 10.2 us at
// M = 1 on q|k|v, 22.5 vs 33.0 us on gate|up: profiles/r1_run11_qmm_sweep_gemv_vs_tensor.txt), so the CUDA-core GEMV only
// serves float32 models, group sizes != 64 and odd shapes.
// B200P_MIN_TC_M / b200p_qmm_set_min_tensor_m override it.
static int g_min_tc_m = [] { const char* e = getenv("B200P_MIN_TC_M"); return e ? atoi(e) : 1; }();
void qmm_tc_set_min_m(int m) { g_min_tc_m = m < 1 ? 1 : m; }
constexpr int MIN_UNITS_PER_CTA = 2;

struct TcPlan {
  int TN, tiles_m, pairs, grid, w_stages, x_slots, smem_bytes;
  long long U;
};

static TcPlan make_plan(int64_t M, int64_t N, int64_t K, int bits, int epi) {
  TcPlan pl;
  const int tn = (int)(M >= MAX_TN ? MAX_TN : ((M + 15) / 16) * 16);
  pl.TN = tn;
  pl.tiles_m = (int)((M + tn - 1) / tn);
  pl.pairs = (int)(epi == B200P_EPI_SILU_MUL ? (N / 2 + 127) / 128 : (N + 255) / 256);
  const int kps = bits == 4 ? 4 : 2;                       // k-blocks per 128-byte stage row
  const int num_st = (int)((K / 64 + kps - 1) / kps);
  pl.U = (long long)pl.pairs * pl.tiles_m * num_st;     // qmm_tc_supported bounds it to 32-bit unit arithmetic
  long long g = (pl.U + MIN_UNITS_PER_CTA - 1) / MIN_UNITS_PER_CTA;
  const int sms = sm_count();
  pl.grid = (int)(g > sms ? sms : (g < 1 ? 1 : g));
  // X ring: one [TN x 64] tile per k-block, 8 slots (4 at TN = 128) -- about 1 us of MMA work in flight, enough to
  // cover the L2 latency of the tile loads; the weight ring gets the rest of shared memory
  pl.x_slots = tn * 128 * 8 <= X_RING_BYTES ? 8 : 4;
  int st = (SMEM_BUDGET - SMEM_FIXED - pl.x_slots * tn * 128) / W_STAGE_BYTES;
  if (st > MAX_STAGES) st = MAX_STAGES;
  pl.w_stages = st;
  pl.smem_bytes = SMEM_FIXED + pl.x_slots * tn * 128 + st * W_STAGE_BYTES;
  return pl;
}

bool qmm_tc_supported(int64_t M, int64_t N, int64_t K, int gs, int bits, int dtype, int epi) {
  if (gs != 64) return false;
  if (dtype != B200P_F16 && dtype != B200P_BF16) return false;
  if (bits == 8 && dtype == B200P_BF16) return false;        // 128 + q is not exact in bf16 for q >= 128
  if (M < g_min_tc_m) return false;
  if (K % 64 != 0 || (K * bits / 8) % 16 != 0) return false;
  if (N % 8 != 0) return false;
  if (epi == B200P_EPI_SILU_MUL && (N / 2) % 8 != 0) return false;
  if (N >= (1ll << 31) || M >= (1ll << 31)) return false;
  return true;
}

static size_t ws_counter_bytes(int64_t M, int64_t N) {
  const int64_t tn = M >= MAX_TN ? MAX_TN : ((M + 15) / 16) * 16;
  const int64_t tiles_m = (M + tn - 1) / tn;
  return ((size_t)(tiles_m * ((N + 255) / 256 + 1)) * 2 * sizeof(int) + 255) & ~(size_t)255;   // {arrived, done} per tile
}

size_t qmm_tc_workspace_bytes(int64_t M, int64_t N, int64_t K) {
  (void)K;
  if (M < 1) return 0;
  const int64_t tn = M >= MAX_TN ? MAX_TN : ((M + 15) / 16) * 16;
  // per-tile counters (ZERO on first use; the kernel leaves them zero) + two fp32 [2*TN x 128] partial-tile slots per CTA
  // (a CTA's first and last segment; no initial state)
  return ws_counter_bytes(M, N) + (size_t)sm_count() * 2 * 2 * tn * 128 * sizeof(float);
}

template <typename T, int BITS, bool XPERM, bool FAST>
static int launch_tc(const CUtensorMap* tms, const TcParams& p, const TcPlan& pl, cudaStream_t st) {
  auto kern = qmm_tc_kernel<T, BITS, XPERM, FAST>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BUDGET);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(qmm_tc_kernel)");
    attr_set = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)pl.grid);
  cfg.blockDim = dim3(TC_THREADS);
  cfg.dynamicSmemBytes = (size_t)pl.smem_bytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled_gemm() ? 1 : 0;
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, tms[0], tms[1], tms[2], tms[3], p);
  if (e != cudaSuccess) return cuda_fail(e, "qmm_tc_kernel launch");
  return B200P_OK;
}

static int encode_2d(EncodeTiledFn enc, CUtensorMap* tm, CUtensorMapDataType dt, const void* base, uint64_t d0, uint64_t d1,
                     uint64_t stride_bytes, uint32_t b0, uint32_t b1, CUtensorMapSwizzle sw, const char* what) {
  cuuint64_t dims[2] = {(cuuint64_t)d0, (cuuint64_t)d1};
  cuuint64_t strides[1] = {(cuuint64_t)stride_bytes};
  cuuint32_t box[2] = {b0, b1};
  cuuint32_t estr[2] = {1u, 1u};
  CUresult r = enc(tm, dt, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B200P_REQUIRE(r == CUDA_SUCCESS, B200P_ERR_CUDA, "cuTensorMapEncodeTiled(%s) failed: %d", what, (int)r);
  return B200P_OK;
}

int qmm_tc_launch(const void* x, const uint32_t* w, const void* s, const void* b, void* y, int64_t M, int64_t N,
                  int64_t K, int64_t ldx, int64_t ldy, int bits, int dtype, int epi, int flags, void* ws, size_t ws_bytes,
                  cudaStream_t st) {
  EncodeTiledFn enc = get_encode();
  B200P_REQUIRE(enc != nullptr, B200P_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
  TcPlan pl = make_plan(M, N, K, bits, epi);
  B200P_REQUIRE(pl.w_stages >= 2, B200P_ERR_ARG, "shared memory too small for the tensor path");
  const bool x_perm = (flags & B200P_X_PERM8) != 0, y_perm = (flags & B200P_Y_PERM8) != 0;
  B200P_REQUIRE(!x_perm || bits == 4, B200P_ERR_ARG, "B200P_X_PERM8 is the 4-bit activation layout");
  B200P_REQUIRE(!y_perm || epi == B200P_EPI_STORE || epi == B200P_EPI_SILU_MUL, B200P_ERR_ARG,
                "B200P_Y_PERM8 applies to 16-bit outputs only");
  const int num_kb = (int)(K / 64);
  const bool sb_tma = (num_kb % 8) == 0 && aligned(s, 16) && aligned(b, 16);
  const CUtensorMapDataType dt16 = dtype == B200P_F16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  const bool fast = (flags & B200P_W_SMALLSCALE) != 0 && dtype == B200P_F16;
  CUtensorMap tms[4];
  int rc = encode_2d(enc, &tms[0], CU_TENSOR_MAP_DATA_TYPE_UINT8, w, (uint64_t)(K * bits / 8), (uint64_t)N,
                     (uint64_t)(K * bits / 8), 128u, 128u, CU_TENSOR_MAP_SWIZZLE_128B, "weights");
  if (rc) return rc;
  rc = encode_2d(enc, &tms[1], dt16, x, (uint64_t)K, (uint64_t)M, (uint64_t)(ldx * 2), 64u, (uint32_t)pl.TN,
                 CU_TENSOR_MAP_SWIZZLE_128B, "x");
  if (rc) return rc;
  if (sb_tma) {
    rc = encode_2d(enc, &tms[2], dt16, s, (uint64_t)num_kb, (uint64_t)N, (uint64_t)num_kb * 2, 8u, 128u,
                   CU_TENSOR_MAP_SWIZZLE_NONE, "scales");
    if (rc) return rc;
    rc = encode_2d(enc, &tms[3], dt16, b, (uint64_t)num_kb, (uint64_t)N, (uint64_t)num_kb * 2, 8u, 128u,
                   CU_TENSOR_MAP_SWIZZLE_NONE, "biases");
    if (rc) return rc;
  } else {
    tms[2] = tms[0];
    tms[3] = tms[0];
  }
  TcParams p;
  p.scales = s; p.biases = b; p.y = y;
  const size_t counters = ws_counter_bytes(M, N);
  const size_t need = counters + (size_t)pl.grid * 2 * 2 * pl.TN * 128 * sizeof(float);
  B200P_REQUIRE(ws_bytes >= need && aligned(ws, 16), B200P_ERR_ARG, "qmm workspace too small (%zu < %zu)", ws_bytes, need);
  {
    const bool f32out = epi == B200P_EPI_STORE_F32 || epi == B200P_EPI_ADD_RESIDUAL;
    B200P_REQUIRE(aligned(y, 4) && (f32out || ldy % 2 == 0), B200P_ERR_ALIGN, "y rows must be 4-byte aligned for the tensor path (ldy=%lld)",
                  (long long)ldy);
  }
  p.ws_counter = reinterpret_cast<int*>(ws);
  p.ws_part = reinterpret_cast<float*>(reinterpret_cast<char*>(ws) + counters);
  p.M_tok = (int)M; p.N = (int)N; p.K = (int)K; p.TN = pl.TN; p.ldy = ldy; p.epi = epi; p.y_perm = y_perm ? 1 : 0;
  B200P_REQUIRE(pl.U * (pl.grid + 1) < (1ll << 31), B200P_ERR_ARG, "problem too large for the tensor path (%lld units)", pl.U);
  p.num_kb = num_kb; p.pairs = pl.pairs; p.tiles_m = pl.tiles_m; p.U = (int)pl.U;
  p.num_st = (p.num_kb + (bits == 4 ? 4 : 2) - 1) / (bits == 4 ? 4 : 2);
  p.sb_tma = sb_tma ? 1 : 0;
  p.w_stages = pl.w_stages; p.x_slots = pl.x_slots;
  if (dtype == B200P_F16) {
    if (bits == 8) return fast ? launch_tc<__half, 8, false, true>(tms, p, pl, st) : launch_tc<__half, 8, false, false>(tms, p, pl, st);
    if (fast) return x_perm ? launch_tc<__half, 4, true, true>(tms, p, pl, st) : launch_tc<__half, 4, false, true>(tms, p, pl, st);
    return x_perm ? launch_tc<__half, 4, true, false>(tms, p, pl, st) : launch_tc<__half, 4, false, false>(tms, p, pl, st);
  } else {
    B200P_REQUIRE(bits == 4, B200P_ERR_ARG, "bf16 8-bit is not on the tensor path");
    return x_perm ? launch_tc<__nv_bfloat16, 4, true, false>(tms, p, pl, st) : launch_tc<__nv_bfloat16, 4, false, false>(tms, p, pl, st);
  }
}

}  // namespace b200p
