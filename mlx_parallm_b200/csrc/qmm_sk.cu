// Stream-K variant of the tensor-core GEMM for the fp32 ACCUMULATE epilogue (B200P_EPI_ADD_RESIDUAL):
//
//   Y[tok, n] += sum_k X[tok, k] * (q[n,k] * s[n, k/64] + b[n, k/64])          Y float32
//
// Same main loop as qmm_tc.cu (swap-AB, weights dequantised straight into TMEM as the A operand, X tiles by TMA, 24 warps:
// 16 dequantisers, 4 epilogue, 2 TMA producers, 2 MMA issuers -- one per 128-row block of the pair, see
// umma_half_kblock_f16_ts in tc_ptx.cuh for why two).  What differs is the schedule and the way out:
//   * persistent stream-K: the work of a launch is the linear sequence of units (output tile of 2 x 128 weight rows,
//     K stage of 128 packed bytes per row); CTA c of G takes the contiguous range [c*U/G, (c+1)*U/G), so every SM streams
//     the same number of weight bytes whatever N and K are -- o_proj / down_proj have 16 tiles for 148 SMs;
//   * every segment (whole tile or not) leaves through shared memory and ONE cp.reduce.async.bulk.tensor (.add.f32) per
//     [16 tokens x 128 rows] staging tile: partial tiles of different CTAs are summed by the L2 atomic units, nothing is
//     waited for, there is no workspace and no tail.  Sums are NOT taken in a fixed order (fp32 adds of <= 148 partials).
// Measured against the whole-tile kernel's K split through the workspace (profiles/r2_run11_*): o_proj M=64 8.5 vs 13.4 us,
// down_proj 16.7 vs 20.3 us.  The model also sends q|k|v through this epilogue at decode sizes (fp32 accumulator that
// rope_kv_append rounds and re-zeroes).
//
// Algorithmic HBM bytes per launch: N*K*bits/8 + 2*N*(K/64)*2.
#include <cuda.h>

#include <stdlib.h>
#include <type_traits>

#include "common.cuh"
#include "dequant.cuh"
#include "tc_ptx.cuh"

namespace b200p {

int sm_count();
bool pdl_enabled_gemm();

namespace sk {

// ------------------------------------------------------------------------------------------------ kernel
constexpr int DQ_WARPS = 16;
constexpr int DQ_THREADS = DQ_WARPS * 32;
constexpr int EPI_WARPS = 4;
constexpr int EPI_THREADS = EPI_WARPS * 32;
constexpr int WARP_W = DQ_WARPS + EPI_WARPS;     // TMA producer: weights + scale/bias tiles (never waits for the previous grid)
constexpr int WARP_X = WARP_W + 1;               // TMA producer: activation tiles (after griddepcontrol.wait)
constexpr int WARP_MMA = WARP_X + 1;
constexpr int MMA_WARPS = 2;                      // one issuing warp per 128-row block of the pair (see umma_half_kblock_f16_ts)
constexpr int TC_THREADS = (WARP_MMA + MMA_WARPS) * 32;
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
constexpr int EPI_TOK = 16;                      // tokens per epilogue staging tile
constexpr int EPI_NBUF = 2;
constexpr int EPI_BUF_BYTES = EPI_TOK * 128 * 4; // [16 tokens x 128 rows] fp32 (16-bit tiles use the first half)

struct TcParams {
  const void* scales;
  const void* biases;
  void* y;
  int M_tok, N, K, TN;
  int64_t ldy;
  int epi;
  int y_perm;          // 1: 16-bit outputs (STORE / SILU_MUL) are written in the PERM8 column order
  int num_kb;          // K / 64
  int num_st;          // stages along K = ceil(num_kb / KPS)
  int pairs, tiles_m;  // output tile t = pair * tiles_m + token tile
  long long U;         // total units = pairs * tiles_m * num_st
  int sb_tma;          // 1: scale/bias tiles come through TMA (num_kb % 8 == 0)
  int w_stages;        // weight ring depth
  int x_slots;         // X ring depth (4 or 8; one slot = the [TN x 64] tile of one k-block)
};


// FAST: the small-scale dequant of dequant.cuh (f16, caller promised B200P_W_SMALLSCALE): A and D are produced at 1/16.
template <typename T, int BITS, bool XPERM, bool FAST>
__global__ void __launch_bounds__(TC_THREADS, 1)
qmm_sk_kernel(const __grid_constant__ CUtensorMap tm_w, const __grid_constant__ CUtensorMap tm_x,
              const __grid_constant__ CUtensorMap tm_s, const __grid_constant__ CUtensorMap tm_b,
              const __grid_constant__ CUtensorMap tm_y, const TcParams p) {
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
  const uint32_t epi_ring = sb_ring + NSB * SB_SLOT_BYTES;
  const uint32_t bar_base = epi_ring + EPI_NBUF * EPI_BUF_BYTES;
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
  const long long u_begin = (p.U * cta) / G, u_end = (p.U * (cta + 1)) / G;

  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) { mbar_init(full_w(s), 1); mbar_init(empty_w(s), DQ_WARPS); }
    for (int s = 0; s < p.x_slots; ++s) { mbar_init(full_x(s), 1); mbar_init(empty_x(s), MMA_WARPS); }
    for (int s = 0; s < NSLOT; ++s) { mbar_init(full_a(s), 8); mbar_init(empty_a(s), MMA_WARPS); }
    for (int s = 0; s < NSB; ++s) { mbar_init(sb_full(s), 1); mbar_init(sb_empty(s), DQ_WARPS); }
    for (int d = 0; d < 2; ++d) { mbar_init(accum_full(d), MMA_WARPS); mbar_init(accum_empty(d), EPI_WARPS); }
    fence_barrier_init();
    tma_prefetch_desc(&tm_w);
    tma_prefetch_desc(&tm_x);
    if (p.sb_tma) { tma_prefetch_desc(&tm_s); tma_prefetch_desc(&tm_b); }
    tma_prefetch_desc(&tm_y);
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
    for (long long u = u_begin; u < u_end; ++u) {
      const int tile = (int)(u / p.num_st), st = (int)(u % p.num_st);
      const int pair = tile / p.tiles_m;
      const int rowA0 = pair * 256;
      const int rowB0 = pair * 256 + 128;
      if (p.sb_tma && (u == u_begin || (st % SPG) == 0)) {     // a new group of 8 k-blocks starts with this stage
        mbar_wait(sb_empty(sbi), sbph ^ 1);
        if (elect_one()) {
          const uint32_t sdst = sb_ring + sbi * SB_SLOT_BYTES;
          const int kb0 = (st / SPG) * 8;
          mbar_arrive_expect_tx(sb_full(sbi), (uint32_t)SB_SLOT_BYTES);
          tma_load_2d(sdst, &tm_s, kb0, rowA0, sb_full(sbi));
          tma_load_2d(sdst + SB_TILE_BYTES, &tm_b, kb0, rowA0, sb_full(sbi));
          tma_load_2d(sdst + 2 * SB_TILE_BYTES, &tm_s, kb0, rowB0, sb_full(sbi));
          tma_load_2d(sdst + 3 * SB_TILE_BYTES, &tm_b, kb0, rowB0, sb_full(sbi));
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
    for (long long u = u_begin; u < u_end; ++u) {
      const int tile = (int)(u / p.num_st), st = (int)(u % p.num_st);
      const int tm = tile % p.tiles_m;
#pragma unroll
      for (int jj = 0; jj < KPS; ++jj, ++it) {
        const uint32_t xs = it & x_mask;
        mbar_wait(empty_x(xs), ((it >> x_shift) & 1u) ^ 1u);
        if (elect_one()) {
          mbar_arrive_expect_tx(full_x(xs), x_slot_bytes);
          tma_load_2d(x_ring + xs * x_slot_bytes, &tm_x, (st * KPS + jj) * 64, tm * TN, full_x(xs));
        }
        __syncwarp();
      }
    }
  } else if (warp >= WARP_MMA) {
    // ===================================== MMA issuers (warp WARP_MMA + b: block b of the pair) =====================
    // instruction descriptor (cute::UMMA::InstrDescriptor): c_format F32 (1<<4), a/b format (0 f16, 1 bf16) at
    // bits 7/10, K-major A and B, N>>3 at bit 17, M>>4 at bit 24
    const uint32_t fmt = std::is_same<T, __nv_bfloat16>::value ? 1u : 0u;
    const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(TN >> 3) << 17) | ((128u >> 4) << 24);
    int seg = 0;
    uint32_t it = 0;
    const uint64_t xdesc0 = make_desc_sw128(x_ring);            // X slot 0, first 16 k; slot s adds s * (slot bytes >> 4)
    const uint32_t xdesc_step = x_slot_bytes >> 4;
    const int mw = warp - WARP_MMA;                    // block of the pair this warp issues for
    const uint32_t a_base = tmem_base + A_COL0 + (uint32_t)(mw * 32);
    for (long long u = u_begin; u < u_end; ++seg) {
      const int s0 = (int)(u % p.num_st);
      const int ns = (int)min((long long)(p.num_st - s0), u_end - u);
      const int buf = NBUF == 2 ? (seg & 1) : 0;
      if (seg >= NBUF) {                               // the epilogue of the segment that last used this D buffer has drained it
        mbar_wait(accum_empty(buf), (uint32_t)(((seg / NBUF) - 1) & 1));
        tc_fence_after();
      }
      const uint32_t dmine = tmem_base + d_col(buf, mw);
      const int nkb = ns * KPS;
      for (int j = 0; j < nkb; ++j, ++it) {
        const uint32_t slot = it & (NSLOT - 1), xs = it & x_mask;
        mbar_wait(full_x(xs), (it >> x_shift) & 1u);
        mbar_wait(full_a(slot), (it >> 2) & 1u);
        tc_fence_after();
        if (elect_one()) {
          umma_half_kblock_f16_ts(dmine, a_base + slot * 64, xdesc0 + (uint64_t)(xs * xdesc_step), idesc, j > 0 ? 1u : 0u);
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
      {
        tmem_st32(a_dst + slot * 64, o);
        tmem_st_wait();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(full_a(slot));
      it += 2;
    };

    for (long long u = u_begin; u < u_end;) {
      const int tile = (int)(u / p.num_st), s0 = (int)(u % p.num_st);
      const int ns = (int)min((long long)(p.num_st - s0), u_end - u);
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
        const int grow = pair * 256 + blk * 128 + r;
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
    // TMEM -> registers -> a [16 tokens x 128 rows] staging tile in shared memory (the transpose is free: thread = row
    // writes column r of every token line, bank = lane) -> ONE TMA tensor store / reduce-add per tile.  Stream-K
    // partial tiles are summed by the TMA unit (cp.reduce.async.bulk .add.f32) instead of 16K scalar atomics per
    // segment, which kept the LSU busy for ~10 us.
    const int e = warp - DQ_WARPS;                     // == warp % 4: the TMEM lane quarter this warp may read
    const int r = e * 32 + lane;                     // output row (weight row) within the 128-row block
    const uint32_t lane_base = tmem_base + ((uint32_t)(e * 32) << 16);
    const int et = threadIdx.x - DQ_THREADS;         // 0..127
    constexpr float DSCALE = FAST ? 16.f : 1.f;      // FAST: A (hence D) was produced at 1/16
    pdl_wait();                                      // y belongs to the previous kernels until here
    int seg = 0;
    uint32_t nchunk = 0;
    for (long long u = u_begin; u < u_end; ++seg) {
      const int tile = (int)(u / p.num_st), s0 = (int)(u % p.num_st);
      const int ns = (int)min((long long)(p.num_st - s0), u_end - u);
      const int pair = tile / p.tiles_m, tm = tile % p.tiles_m;
      const int rowA0 = pair * 256;
      const int rowB0 = pair * 256 + 128;
      const int tok0 = tm * TN;
      const int buf = NBUF == 2 ? (seg & 1) : 0;

      mbar_wait(accum_full(buf), (uint32_t)((seg / NBUF) & 1));
      tc_fence_after();
      const uint32_t dA = lane_base + d_col(buf, 0), dB = lane_base + d_col(buf, 1);
      for (int t0 = 0; t0 < TN; t0 += EPI_TOK) {
#pragma unroll
        for (int blk = 0; blk < 2; ++blk) {
          float v[EPI_TOK];
          tmem_ld16((blk ? dB : dA) + (uint32_t)t0, v);
          tmem_ld_wait();
          // claim staging buffer (nchunk & 1): the bulk group that last read it must be done with shared memory
          if (et == 0) bulk_wait_read<EPI_NBUF - 1>();
          named_bar_sync(1, EPI_THREADS);
          const uint32_t sb = epi_ring + (nchunk & (EPI_NBUF - 1)) * EPI_BUF_BYTES;
#pragma unroll
          for (int j = 0; j < EPI_TOK; ++j) sts32(sb + (uint32_t)(j * 128 + r) * 4u, __float_as_uint(v[j] * DSCALE));
          fence_proxy_async_smem();
          named_bar_sync(2, EPI_THREADS);
          if (et == 0) {                               // rows / tokens outside the tensor are clipped by the TMA unit
            tma_reduce_add_2d(&tm_y, blk ? rowB0 : rowA0, tok0 + t0, sb);
            bulk_commit();
          }
          ++nchunk;
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(accum_empty(buf));    // this D buffer may be overwritten by a later segment's MMAs
      u += ns;
    }
    if (et == 0) bulk_wait_all();                      // staging memory and outputs are settled before the CTA retires
  }
  tc_fence_before();
  __syncthreads();
  if (warp == WARP_MMA) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------------ host
constexpr int SMEM_BUDGET = 227 * 1024;
constexpr int SMEM_FIXED = 1024 /*align*/ + NSB * SB_SLOT_BYTES + EPI_NBUF * EPI_BUF_BYTES + 512 /*barriers*/;
static_assert(8 * (4 * MAX_STAGES + 2 * NSLOT + 2 * NSB + 4) + 8 <= 512, "barrier block overflows its slot");
constexpr int MIN_UNITS_PER_CTA = 2;

struct TcPlan {
  int TN, tiles_m, pairs, grid, w_stages, x_slots, smem_bytes;
  long long U;
};

static TcPlan make_plan(int64_t M, int64_t N, int64_t K, int bits) {
  TcPlan pl;
  const int tn = (int)(M >= MAX_TN ? MAX_TN : ((M + 15) / 16) * 16);
  pl.TN = tn;
  pl.tiles_m = (int)((M + tn - 1) / tn);
  pl.pairs = (int)((N + 255) / 256);
  const int kps = bits == 4 ? 4 : 2;                       // k-blocks per 128-byte stage row
  const int num_st = (int)((K / 64 + kps - 1) / kps);
  pl.U = (long long)pl.pairs * pl.tiles_m * num_st;
  long long g = (pl.U + MIN_UNITS_PER_CTA - 1) / MIN_UNITS_PER_CTA;
  const int sms = sm_count();
  pl.grid = (int)(g > sms ? sms : (g < 1 ? 1 : g));
  // X ring: one [TN x 64] tile per k-block, 8 slots (4 at TN = 128); the weight ring gets the rest of shared memory
  pl.x_slots = tn * 128 * 8 <= X_RING_BYTES ? 8 : 4;
  int st = (SMEM_BUDGET - SMEM_FIXED - pl.x_slots * tn * 128) / W_STAGE_BYTES;
  if (st > MAX_STAGES) st = MAX_STAGES;
  pl.w_stages = st;
  pl.smem_bytes = SMEM_FIXED + pl.x_slots * tn * 128 + st * W_STAGE_BYTES;
  return pl;
}

template <typename T, int BITS, bool XPERM, bool FAST>
static int launch_sk(const CUtensorMap* tms, const TcParams& p, const TcPlan& pl, cudaStream_t st) {
  auto kern = qmm_sk_kernel<T, BITS, XPERM, FAST>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BUDGET);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(qmm_sk_kernel)");
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
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, tms[0], tms[1], tms[2], tms[3], tms[4], p);
  if (e != cudaSuccess) return cuda_fail(e, "qmm_sk_kernel launch");
  return B200P_OK;
}

}  // namespace sk

// the caller (c_api.cu) has already checked qmm_tc_supported(): same operand requirements, plus 16-byte aligned y rows
bool qmm_sk_supported(int64_t M, int64_t N, int64_t K, int bits, int epi, const void* y, int64_t ldy) {
  (void)M; (void)K; (void)bits;
  return epi == B200P_EPI_ADD_RESIDUAL && N % 4 == 0 && aligned(y, 16) && (ldy * 4) % 16 == 0;
}

int qmm_sk_launch(const void* x, const uint32_t* w, const void* s, const void* b, void* y, int64_t M, int64_t N,
                  int64_t K, int64_t ldx, int64_t ldy, int bits, int dtype, int flags, cudaStream_t st) {
  using namespace sk;
  TcEncodeFn enc = tc_get_encode();
  B200P_REQUIRE(enc != nullptr, B200P_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
  TcPlan pl = make_plan(M, N, K, bits);
  B200P_REQUIRE(pl.w_stages >= 2, B200P_ERR_ARG, "shared memory too small for the tensor path");
  const bool x_perm = (flags & B200P_X_PERM8) != 0;
  B200P_REQUIRE(!x_perm || bits == 4, B200P_ERR_ARG, "B200P_X_PERM8 is the 4-bit activation layout");
  B200P_REQUIRE((flags & B200P_Y_PERM8) == 0, B200P_ERR_ARG, "B200P_Y_PERM8 applies to 16-bit outputs only");
  const int num_kb = (int)(K / 64);
  const bool sb_tma = (num_kb % 8) == 0 && aligned(s, 16) && aligned(b, 16);
  const bool fast = (flags & B200P_W_SMALLSCALE) != 0 && dtype == B200P_F16;
  const CUtensorMapDataType dt16 = dtype == B200P_F16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  CUtensorMap tms[5];
  int rc = tc_encode_2d(enc, &tms[0], CU_TENSOR_MAP_DATA_TYPE_UINT8, w, (uint64_t)(K * bits / 8), (uint64_t)N,
                        (uint64_t)(K * bits / 8), 128u, 128u, CU_TENSOR_MAP_SWIZZLE_128B, "weights");
  if (rc) return rc;
  rc = tc_encode_2d(enc, &tms[1], dt16, x, (uint64_t)K, (uint64_t)M, (uint64_t)(ldx * 2), 64u, (uint32_t)pl.TN,
                    CU_TENSOR_MAP_SWIZZLE_128B, "x");
  if (rc) return rc;
  if (sb_tma) {
    rc = tc_encode_2d(enc, &tms[2], dt16, s, (uint64_t)num_kb, (uint64_t)N, (uint64_t)num_kb * 2, 8u, 128u,
                      CU_TENSOR_MAP_SWIZZLE_NONE, "scales");
    if (rc) return rc;
    rc = tc_encode_2d(enc, &tms[3], dt16, b, (uint64_t)num_kb, (uint64_t)N, (uint64_t)num_kb * 2, 8u, 128u,
                      CU_TENSOR_MAP_SWIZZLE_NONE, "biases");
    if (rc) return rc;
  } else {
    tms[2] = tms[0];
    tms[3] = tms[0];
  }
  // output tensor: [M tokens x N columns] fp32; staging tiles are [16 tokens x 128 columns]
  rc = tc_encode_2d(enc, &tms[4], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, y, (uint64_t)N, (uint64_t)M, (uint64_t)ldy * 4, 128u,
                    (uint32_t)EPI_TOK, CU_TENSOR_MAP_SWIZZLE_NONE, "y");
  if (rc) return rc;
  TcParams p;
  p.scales = s; p.biases = b; p.y = y;
  p.M_tok = (int)M; p.N = (int)N; p.K = (int)K; p.TN = pl.TN; p.ldy = ldy; p.epi = B200P_EPI_ADD_RESIDUAL; p.y_perm = 0;
  p.num_kb = num_kb; p.pairs = pl.pairs; p.tiles_m = pl.tiles_m; p.U = pl.U;
  p.num_st = (p.num_kb + (bits == 4 ? 4 : 2) - 1) / (bits == 4 ? 4 : 2);
  p.sb_tma = sb_tma ? 1 : 0;
  p.w_stages = pl.w_stages; p.x_slots = pl.x_slots;
  if (dtype == B200P_F16) {
    if (bits == 8) return fast ? launch_sk<__half, 8, false, true>(tms, p, pl, st) : launch_sk<__half, 8, false, false>(tms, p, pl, st);
    if (fast) return x_perm ? launch_sk<__half, 4, true, true>(tms, p, pl, st) : launch_sk<__half, 4, false, true>(tms, p, pl, st);
    return x_perm ? launch_sk<__half, 4, true, false>(tms, p, pl, st) : launch_sk<__half, 4, false, false>(tms, p, pl, st);
  }
  if (bits == 8) return launch_sk<__nv_bfloat16, 8, false, false>(tms, p, pl, st);
  return x_perm ? launch_sk<__nv_bfloat16, 4, true, false>(tms, p, pl, st) : launch_sk<__nv_bfloat16, 4, false, false>(tms, p, pl, st);
}

}  // namespace b200p
