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
// warp roles (768 threads):
//   warps 0..15  dequantisers: 4 groups of 4 warps, group g handles weight block g&1 on k-blocks of parity g>>1;
//                a thread owns one weight row: LDS packed codes -> 16-bit q*s+b -> tcgen05.st into the A ring;
//   warps 16..19 epilogue: TMEM -> registers -> global (thread = output column n: every warp store is one coalesced
//                line per token), stream-K partial tiles and their final reduction;
//   warp 20      TMA producer of weights and scale/bias tiles (independent of the previous kernel: starts at once);
//   warp 21      TMA producer of X tiles (after griddepcontrol.wait);
//   warps 22, 23 MMA issuers, one per 128-row block of the pair (elect.sync, operands in uniform registers); warp 22 owns TMEM.
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
#include <initializer_list>
#include <mutex>
#include <type_traits>
#include <unordered_map>

#include "common.cuh"
#include "dequant.cuh"
#include "tc_ptx.cuh"

namespace b200p {

int sm_count();
bool pdl_enabled_gemm();

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
constexpr int NSLOT = 8;                         // upper bound of A ring slots (each: 2 blocks x 32 TMEM columns); a launch
                                                 // uses p.a_slots (even) = what the D buffers leave of the 512 TMEM columns
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
// Profiling build only (-DB200P_TC_TIMING=1): CTA 0 accumulates the cycles each role spends in each barrier wait into
// g_tc_timing (read back with b200p_debug_tc_timing); slots are listed in scripts/tc_timing.py.
#ifndef B200P_TC_TIMING
#define B200P_TC_TIMING 0
#endif
#if B200P_TC_TIMING
__device__ unsigned long long g_tc_timing[32];
#define TWAIT(slot, bar, par)                                   \
  do {                                                          \
    const long long t0_ = clock64();                            \
    mbar_wait(bar, par);                                        \
    tacc[slot] += (unsigned long long)(clock64() - t0_);        \
  } while (0)
#define TDECL unsigned long long tacc[4] = {0ull, 0ull, 0ull, 0ull}; const long long tstart_ = clock64()
// section timers of the dequant warps (slots 16..19: LDS fetch, math, TMEM store + wait::st, fence + publish)
#define TSEC_DECL unsigned long long tsec[4] = {0ull, 0ull, 0ull, 0ull}; long long tmark_ = 0
#define TSEC_BEGIN() tmark_ = clock64()
#define TSEC_END(i) do { const long long n_ = clock64(); tsec[i] += (unsigned long long)(n_ - tmark_); tmark_ = n_; } while (0)
#define TSEC_FLUSH(cond) do { if (blockIdx.x == 0 && (cond)) for (int i_ = 0; i_ < 4; ++i_) g_tc_timing[16 + i_] = tsec[i_]; } while (0)
#define TFLUSH(base, n, cond)                                                                                   \
  do {                                                                                                          \
    if (blockIdx.x == 0 && (cond)) {                                                                            \
      for (int i_ = 0; i_ < (n); ++i_) g_tc_timing[(base) + i_] = tacc[i_];                                     \
      g_tc_timing[(base) + (n)] = (unsigned long long)(clock64() - tstart_);                                    \
    }                                                                                                           \
  } while (0)
#else
#define TWAIT(slot, bar, par) mbar_wait(bar, par)
#define TDECL
#define TFLUSH(base, n, cond)
#define TSEC_DECL
#define TSEC_BEGIN()
#define TSEC_END(i)
#define TSEC_FLUSH(cond)
#endif

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
  int R;               // weight rows per MMA block of a tile (multiple of 8, <= 128): a tile is R + R rows
  int S;               // K splits per tile (1: whole tiles; > 1 only when tiles * S <= grid, one item per CTA)
  int items;           // work items = tiles * S; CTA c takes items c, c + G, c + 2G, ...
  int rotate;          // 1: an item's stages are visited in rotated order (seg_of)
  int nbuf;            // D buffers (2: the epilogue of one item overlaps the main loop of the next; 1 frees TMEM for the A ring)
  int a_slots;         // A ring depth: (512 - D columns) / 64, even
  int sb_tma;          // 1: scale/bias tiles come through TMA (num_kb % 8 == 0)
  int w_stages;        // weight ring depth
  int x_slots;         // X ring depth (4 or 8; one slot = the [TN x 64] tile of one k-block)
};

// work item -> (output tile, stage range): item = tile * S + split, split covers stages [split*num_st/S, (split+1)*num_st/S)
// The stages of an item are visited in ROTATED order, starting at stage s0 + rot: CTAs that walk K in lockstep (all at the
// same 128-byte column of their 2 KB rows) camp on a few HBM channels -- measured 1.0 instead of 2.6 TB/s on gate|up.
struct TcSeg { int tile, s0, ns, rot; };
__device__ __forceinline__ TcSeg seg_of(const TcParams& p, int item) {
  TcSeg g;
  g.tile = item / p.S;
  const int split = item - g.tile * p.S;
  g.s0 = (split * p.num_st) / p.S;
  g.ns = ((split + 1) * p.num_st) / p.S - g.s0;
  g.rot = p.rotate ? item % g.ns : 0;
  return g;
}
__device__ __forceinline__ int seg_stage(const TcSeg& g, int i) {
  const int j = i + g.rot;
  return g.s0 + (j >= g.ns ? j - g.ns : j);
}

// Fused epilogue of EPI_TOK tokens of one thread's two output columns (a: block A row r, b: block B row r).  Thread = one
// output column, so every warp-level store below is one contiguous line per token (32 lanes x 4 B, or x 2 B for 16-bit
// outputs -- whole sectors either way): nothing is staged through shared memory and no store is atomic.
template <typename T>
__device__ __forceinline__ void tc_emit(const TcParams& p, bool silu, int r, const float (&a)[EPI_TOK], const float (&b)[EPI_TOK],
                                        int rowA0, int rowB0, int tokb) {
    // rows r >= R are not part of this tile (their TMEM lanes hold garbage): pushing n past N disables every store below
    const int n0 = r < p.R ? rowA0 + r : p.N, n1 = r < p.R ? rowB0 + r : p.N;
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
  }

// Drain one WHOLE accumulator tile (both blocks) of this thread's TMEM lane through the fused epilogue: token groups
// worker, worker + nworkers, ... of EPI_TOK tokens each.  The epilogue warps call it with (0, 1); on a CTA's LAST item the
// dequant warps of the same lane quarter have nothing left to do and join as workers 1..4 (nworkers = 5), which cuts the
// tail of a one-tile-per-CTA launch (gate|up at 64 rows: 8 groups, 2 per worker instead of 8 in a row).
template <typename T, bool FAST>
__device__ __forceinline__ void tc_drain_tile(const TcParams& p, bool silu, int r, uint32_t dA, uint32_t dB, int TN, int rowA0,
                                              int rowB0, int tok0, int worker, int nworkers) {
  constexpr float DSCALE = FAST ? 16.f : 1.f;        // FAST: A (hence D) was produced at 1/16
  for (int t0 = worker * EPI_TOK; t0 < TN; t0 += nworkers * EPI_TOK) {
    if (tok0 + t0 >= p.M_tok) break;                 // padding tokens of the last token tile
    float a[EPI_TOK], b[EPI_TOK];
    tmem_ld8(dA + (uint32_t)t0, a);
    tmem_ld8(dB + (uint32_t)t0, b);
    tmem_ld_wait();
    if constexpr (FAST) {
#pragma unroll
      for (int j = 0; j < EPI_TOK; ++j) { a[j] *= DSCALE; b[j] *= DSCALE; }
    }
    tc_emit<T>(p, silu, r, a, b, rowA0, rowB0, tok0 + t0);
  }
}

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
  const uint32_t w_tile_bytes = (uint32_t)p.R * 128u;          // one block of one stage: R rows x 128 packed bytes
  const uint32_t w_stage_bytes = 2u * w_tile_bytes;
  const uint32_t x_ring = w_ring + S * w_stage_bytes;
  const uint32_t x_slot_bytes = (uint32_t)(TN * 128);
  const uint32_t x_mask = (uint32_t)(p.x_slots - 1);
  const int x_shift = p.x_slots == 8 ? 3 : 2;
  const uint32_t sb_ring = x_ring + p.x_slots * x_slot_bytes;
  const uint32_t bar_base = sb_ring + NSB * SB_SLOT_BYTES;
  const int NBUF = p.nbuf;                       // D buffers: buffer d, block b at columns (2d + b) * d_stride
  const uint32_t d_stride = TN <= 32 ? 32u : (TN <= 64 ? 64u : 128u);
  const uint32_t A_COL0 = 512u - (uint32_t)p.a_slots * 64u;   // the A ring sits at the top of TMEM
  const uint32_t a_slots = (uint32_t)p.a_slots;
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
  auto d_col = [&](int d, int b) { return (uint32_t)(2 * d + b) * d_stride; };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int G = (int)gridDim.x, cta = (int)blockIdx.x;
  const bool silu = p.epi == B200P_EPI_SILU_MUL;

  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) { mbar_init(full_w(s), 1); mbar_init(empty_w(s), DQ_WARPS); }
    for (int s = 0; s < p.x_slots; ++s) { mbar_init(full_x(s), 1); mbar_init(empty_x(s), MMA_WARPS); }
    for (int s = 0; s < p.a_slots; ++s) { mbar_init(full_a(s), 8); mbar_init(empty_a(s), MMA_WARPS); }
    for (int s = 0; s < NSB; ++s) { mbar_init(sb_full(s), 1); mbar_init(sb_empty(s), DQ_WARPS); }
    for (int d = 0; d < 2; ++d) { mbar_init(accum_full(d), MMA_WARPS); mbar_init(accum_empty(d), EPI_WARPS); }
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
    TDECL;
    for (int item = cta; item < p.items; item += G) {
     const TcSeg sg = seg_of(p, item);
     const int pair = sg.tile / p.tiles_m;
     const int rowA0 = silu ? pair * p.R : pair * 2 * p.R;
     const int rowB0 = silu ? p.N / 2 + pair * p.R : pair * 2 * p.R + p.R;
     for (int i = 0; i < sg.ns; ++i) {
      const int st = seg_stage(sg, i);
      if (p.sb_tma && (i == 0 || st == sg.s0 || (st % SPG) == 0)) {     // a new group of 8 k-blocks starts with this stage
        TWAIT(1, sb_empty(sbi), sbph ^ 1);
        if (elect_one()) {
          const uint32_t sdst = sb_ring + sbi * SB_SLOT_BYTES;
          const int kb0 = (st / SPG) * 8;
          if (kDebug & 32) {
            mbar_arrive(sb_full(sbi));
          } else {
            mbar_arrive_expect_tx(sb_full(sbi), (uint32_t)(4 * p.R * 16));
            tma_load_2d(sdst, &tm_s, kb0, rowA0, sb_full(sbi));
            tma_load_2d(sdst + SB_TILE_BYTES, &tm_b, kb0, rowA0, sb_full(sbi));
            tma_load_2d(sdst + 2 * SB_TILE_BYTES, &tm_s, kb0, rowB0, sb_full(sbi));
            tma_load_2d(sdst + 3 * SB_TILE_BYTES, &tm_b, kb0, rowB0, sb_full(sbi));
          }
        }
        __syncwarp();
        if (++sbi == NSB) { sbi = 0; sbph ^= 1; }
      }
      TWAIT(0, empty_w(si), ph ^ 1);
      if (elect_one()) {
        const uint32_t dst = w_ring + si * w_stage_bytes;
        mbar_arrive_expect_tx(full_w(si), w_stage_bytes);
        tma_load_2d(dst, &tm_w, st * 128, rowA0, full_w(si));
        tma_load_2d(dst + w_tile_bytes, &tm_w, st * 128, rowB0, full_w(si));
      }
      __syncwarp();
      if (++si == S) { si = 0; ph ^= 1; }
     }
    }
    TFLUSH(0, 2, lane == 0);                         // 0 empty_w, 1 sb_empty, 2 total
  } else if (warp == WARP_X) {
    // ===================================== activation producer ==============================
    pdl_wait();                                      // X is the previous kernel's output
    uint32_t it = 0;
    TDECL;
    for (int item = cta; item < p.items; item += G) {
     const TcSeg sg = seg_of(p, item);
     const int tm = sg.tile % p.tiles_m;
     for (int i = 0; i < sg.ns; ++i) {
      const int st = seg_stage(sg, i);
#pragma unroll
      for (int jj = 0; jj < KPS; ++jj, ++it) {
        const uint32_t xs = it & x_mask;
        TWAIT(0, empty_x(xs), ((it >> x_shift) & 1u) ^ 1u);
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
    }
    TFLUSH(3, 1, lane == 0);                         // 3 empty_x, 4 total
  } else if (warp >= WARP_MMA) {
    // ===================================== MMA issuers (warp WARP_MMA + b: block b of the pair) =====================
    // instruction descriptor (cute::UMMA::InstrDescriptor): c_format F32 (1<<4), a/b format (0 f16, 1 bf16) at
    // bits 7/10, K-major A and B, N>>3 at bit 17, M>>4 at bit 24
    const uint32_t fmt = std::is_same<T, __nv_bfloat16>::value ? 1u : 0u;
    const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(TN >> 3) << 17) | ((128u >> 4) << 24);
    int seg = 0;
    uint32_t slot = 0, aph = 0, xs = 0, xph = 0;
    TDECL;
    const uint64_t xdesc0 = make_desc_sw128(x_ring);            // X slot 0, first 16 k; slot s adds s * (slot bytes >> 4)
    const uint32_t xdesc_step = x_slot_bytes >> 4;
    const int mw = warp - WARP_MMA;                    // block of the pair this warp issues for
    const uint32_t a_base = tmem_base + A_COL0 + (uint32_t)(mw * 32);
    const uint32_t x_slots = (uint32_t)p.x_slots;
    for (int item = cta; item < p.items; item += G, ++seg) {
      const int ns = seg_of(p, item).ns;
      const int buf = NBUF == 2 ? (seg & 1) : 0;
      if (seg >= NBUF) {                               // the epilogue of the segment that last used this D buffer has drained it
        TWAIT(2, accum_empty(buf), (uint32_t)(((seg / NBUF) - 1) & 1));
        tc_fence_after();
      }
      const uint32_t dmine = tmem_base + d_col(buf, mw);
      const int nkb = ns * KPS;
      for (int j = 0; j < nkb; ++j) {
        TWAIT(0, full_x(xs), xph);
        TWAIT(1, full_a(slot), aph);
        tc_fence_after();
        if (elect_one()) {
          if (!(kDebug & 2))
            umma_half_kblock_f16_ts(dmine, a_base + slot * 64, xdesc0 + (uint64_t)(xs * xdesc_step), idesc, j > 0 ? 1u : 0u);
          umma_commit(empty_a(slot));
          umma_commit(empty_x(xs));
          if (j == nkb - 1) umma_commit(accum_full(buf));
        }
        __syncwarp();
        if (++slot == a_slots) { slot = 0; aph ^= 1u; }
        if (++xs == x_slots) { xs = 0; xph ^= 1u; }
      }
    }
    TFLUSH(5, 3, lane == 0 && mw == 0);              // 5 full_x, 6 full_a, 7 accum_empty, 8 total
  } else if (warp < DQ_WARPS) {
    // ===================================== dequantisers =====================================
    const int grp = warp >> 2;                         // 0..3
    const int blk = grp & 1, kpar = grp >> 1;
    const int q = warp & 3;                          // TMEM lane quarter this warp may touch
    const int r = q * 32 + lane;                     // weight row within the 128-row block
    const uint32_t a_dst = tmem_base + ((uint32_t)(q * 32) << 16) + A_COL0 + (uint32_t)(blk * 32);
    // SWIZZLE_128B tile [128 rows x 128 B]: row r at (r/8)*1024 + (r%8)*128, 16-byte chunk c at (c ^ (r%8))
    const uint32_t w_row_off = (uint32_t)blk * w_tile_bytes + (uint32_t)((r >> 3) * 1024 + (r & 7) * 128);
    // rows r >= R of a block carry no weights: a warp with none of them skips the math and the TMEM store (its D rows are
    // never read) but keeps every barrier protocol; single idle lanes of an active warp compute on stale shared memory
    const bool warp_active = q * 32 < p.R;
    const uint32_t w_xor = (uint32_t)(r & 7);
    const uint32_t sb_off = (uint32_t)(blk * 2 * SB_TILE_BYTES + r * 16);
    const uint32_t half_sel = kpar ? 0x3232u : 0x1010u;   // which 16-bit half of a scale word is this thread's k-block

    int si = 0, ph = 0, sbi = 0, sbph = 0;
    uint32_t slot = (uint32_t)kpar, aph = 0;         // A ring slot of this thread's next k-block (it advances by 2) + phase
    uint4 s8 = make_uint4(0u, 0u, 0u, 0u), b8 = s8;
    TDECL;
    TSEC_DECL;

    constexpr int CPK = PKB / 16;                    // 16-byte chunks of packed codes per k-block
    constexpr int KPT = KPS / 2;                     // k-blocks per thread per stage
    uint4 pk[KPT * CPK];
    // wait for the stage's weights, pull this thread's codes into registers and hand the slot straight back to the
    // producer: the HBM stream then only waits for LDS, not for the dequant math or the MMAs
    auto fetch_stage = [&]() {
      TWAIT(0, full_w(si), (uint32_t)ph);
      TSEC_BEGIN();
      const uint32_t wsrc = w_ring + si * w_stage_bytes + w_row_off;
#pragma unroll
      for (int j2 = 0; j2 < KPT; ++j2)
#pragma unroll
        for (int c = 0; c < CPK; ++c)
          pk[j2 * CPK + c] = lds128(wsrc + ((((uint32_t)((kpar + 2 * j2) * CPK + c)) ^ w_xor) << 4));
      __syncwarp();
      if (lane == 0) mbar_arrive(empty_w(si));
      TSEC_END(0);
      if (++si == S) { si = 0; ph ^= 1; }
    };
    // one k-block: 64 codes of row r -> 32 packed 16-bit pairs -> A ring slot
    auto kblock = [&](int j2, uint32_t s2, uint32_t b2) {
      uint32_t o[32];
      TSEC_BEGIN();
      const uint32_t* pw = reinterpret_cast<const uint32_t*>(pk + j2 * CPK);
      if constexpr (FAST) {
        const SmallScale sc = small_scale_prep(s2, b2);
#pragma unroll
        for (int c = 0; c < 8; ++c) dequant8_small<BITS, XPERM>(pw + c * (BITS / 4), sc, o + 4 * c);
      } else {
#pragma unroll
        for (int c = 0; c < 8; ++c) dequant8<T, BITS, XPERM>(pw + c * (BITS / 4), s2, b2, o + 4 * c);
      }
      TSEC_END(1);
      TWAIT(2, empty_a(slot), aph ^ 1u);
      TSEC_BEGIN();
      tc_fence_after();
      if (!(kDebug & 8)) {
        tmem_st32(a_dst + slot * 64, o);
        tmem_st_wait();
      }
      TSEC_END(2);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(full_a(slot));
      TSEC_END(3);
      slot += 2;
      if (slot >= a_slots) { slot -= a_slots; aph ^= 1u; }
    };

    if (!warp_active) {
      // barrier protocol only: this warp's 32 rows are beyond the R rows of a block
      for (int item = cta; item < p.items; item += G) {
        const TcSeg sg = seg_of(p, item);
        for (int i = 0; i < sg.ns; ++i) {
          const int st = seg_stage(sg, i);
          if (p.sb_tma && (i == 0 || st == sg.s0 || (st % SPG) == 0)) {
            mbar_wait(sb_full(sbi), (uint32_t)sbph);
            __syncwarp();
            if (lane == 0) mbar_arrive(sb_empty(sbi));
            if (++sbi == NSB) { sbi = 0; sbph ^= 1; }
          }
          mbar_wait(full_w(si), (uint32_t)ph);
          __syncwarp();
          if (lane == 0) mbar_arrive(empty_w(si));
          if (++si == S) { si = 0; ph ^= 1; }
          for (int j2 = 0; j2 < KPS / 2; ++j2) {
            mbar_wait(empty_a(slot), aph ^ 1u);
            __syncwarp();
            if (lane == 0) mbar_arrive(full_a(slot));
            slot += 2;
            if (slot >= a_slots) { slot -= a_slots; aph ^= 1u; }
          }
        }
      }
    } else
    for (int item = cta; item < p.items; item += G) {
      const TcSeg sg = seg_of(p, item);
      const int tile = sg.tile, s0 = sg.s0, ns = sg.ns;
      if (p.sb_tma) {
        for (int i = 0; i < ns; ++i) {
          const int st = seg_stage(sg, i);
          if (i == 0 || st == s0 || (st % SPG) == 0) {   // next group of 8 k-blocks: scales / biases of this row
            TWAIT(1, sb_full(sbi), (uint32_t)sbph);
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
        const int grow = (blk ? (silu ? p.N / 2 + pair * p.R : pair * 2 * p.R + p.R) : (silu ? pair * p.R : pair * 2 * p.R)) + r;
        const bool row_ok = grow < p.N && r < p.R;
        const T* sp = reinterpret_cast<const T*>(p.scales) + (int64_t)(row_ok ? grow : 0) * p.num_kb;
        const T* bp = reinterpret_cast<const T*>(p.biases) + (int64_t)(row_ok ? grow : 0) * p.num_kb;
        const T zero = Tr<T>::from_f(0.f);
        for (int i = 0; i < ns; ++i) {
          const int st = seg_stage(sg, i);
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
    }
    TFLUSH(9, 3, threadIdx.x == 0);                  // 9 full_w, 10 sb_full, 11 empty_a, 12 total (dequant warp 0)
    TSEC_FLUSH(threadIdx.x == 0);
    // ---- help the epilogue warp of this lane quarter drain the CTA's LAST whole tile (see tc_drain_tile)
    if (p.S == 1 && warp_active && cta < p.items && !(kDebug & 4)) {
      const int n_seg = (p.items - cta + G - 1) / G, seg = n_seg - 1, item = cta + seg * G;
      const TcSeg sg = seg_of(p, item);
      const int pair = sg.tile / p.tiles_m, tm = sg.tile % p.tiles_m;
      const int buf = NBUF == 2 ? (seg & 1) : 0;
      pdl_wait();                                    // y belongs to the previous kernels until here
      mbar_wait(accum_full(buf), (uint32_t)((seg / NBUF) & 1));
      tc_fence_after();
      const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16);
      tc_drain_tile<T, FAST>(p, silu, r, lane_base + d_col(buf, 0), lane_base + d_col(buf, 1), TN, silu ? pair * p.R : pair * 2 * p.R,
                             silu ? p.N / 2 + pair * p.R : pair * 2 * p.R + p.R, tm * TN, 1 + grp, 1 + DQ_WARPS / 4);
    }
  } else {
    // ===================================== epilogue warps ===================================
    TDECL;
    // Thread = one output column n (weight row r of block A and of block B), EPI_TOK tokens per step: every warp-level
    // load / store below touches one contiguous line per token (32 lanes x 4 B, or x 2 B for 16-bit outputs -- whole
    // sectors either way), so nothing is staged through shared memory and no store is atomic.
    const int e = warp - DQ_WARPS;                     // == warp % 4: the TMEM lane quarter this warp may read
    const int r = e * 32 + lane;                     // output row (weight row) within the 128-row block
    const uint32_t lane_base = tmem_base + ((uint32_t)(e * 32) << 16);
    const int et = threadIdx.x - DQ_THREADS;         // 0..127
    constexpr float DSCALE = FAST ? 16.f : 1.f;      // FAST: A (hence D) was produced at 1/16
    const int64_t slot_elems = (int64_t)2 * TN * 128;   // one partial tile: [2 blocks][TN tokens][128 rows] fp32

    pdl_wait();                                      // y and the workspace belong to the previous kernels until here
    int seg = 0;
    int def0 = -1;                                   // the split tile this CTA contributed to (S > 1: one item per CTA)
    for (int item = cta; item < p.items; item += G, ++seg) {
      const TcSeg sg = seg_of(p, item);
      const int tile = sg.tile;
      const int pair = tile / p.tiles_m, tm = tile % p.tiles_m;
      const int rowA0 = silu ? pair * p.R : pair * 2 * p.R;
      const int rowB0 = silu ? p.N / 2 + pair * p.R : pair * 2 * p.R + p.R;
      const int tok0 = tm * TN;
      const int buf = NBUF == 2 ? (seg & 1) : 0;
      const bool full = p.S == 1;

      TWAIT(0, accum_full(buf), (uint32_t)((seg / NBUF) & 1));
      tc_fence_after();
      const uint32_t dA = lane_base + d_col(buf, 0), dB = lane_base + d_col(buf, 1);
      if (!(kDebug & 4)) {
        if (full) {
          const bool last = item + G >= p.items;       // the dequant warps help with this CTA's last tile
          tc_drain_tile<T, FAST>(p, silu, r, dA, dB, TN, rowA0, rowB0, tok0, 0, last ? 1 + DQ_WARPS / 4 : 1);
        } else {
          // partial tile of a K split: one slot per CTA (the host plans S > 1 only with one item per CTA)
          float* part = p.ws_part + (int64_t)cta * slot_elems;
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
          def0 = tile;
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(accum_empty(buf));    // this D buffer may be overwritten by a later segment's MMAs
    }

    TFLUSH(13, 1, et == 0);                          // 13 accum_full, 14 total up to the end of the tile epilogues
    // ---- split tile: meet the S - 1 other contributors (CTAs tile*S .. tile*S + S - 1: equal work, they arrive together),
    // then reduce and finish a share of the tile
    if (def0 >= 0) {
      __threadfence();                                 // this thread's partial stores are performed device-wide ...
      named_bar_sync(1, EPI_THREADS);                  // ... for all 128 epilogue threads before the arrival is published
      const int tile = def0;
      const int S = p.S, c_first = tile * S, me = cta - c_first;
      if (et == 0) {
        atomicAdd(p.ws_counter + 2 * tile, 1);
        const volatile int* cnt = p.ws_counter + 2 * tile;
        unsigned long long t_start;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_start));
        while (*cnt < S) {
          __nanosleep(20);
          unsigned long long t_now;
          asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_now));
          if (t_now - t_start > 4000000000ull) __trap();      // a contributor never arrived: fail loudly, do not hang
        }
        __threadfence();
      }
      named_bar_sync(2, EPI_THREADS);
      const int pair = tile / p.tiles_m, tm = tile % p.tiles_m;
      const int rowA0 = silu ? pair * p.R : pair * 2 * p.R;
      const int rowB0 = silu ? p.N / 2 + pair * p.R : pair * 2 * p.R + p.R;
      const int tok0 = tm * TN;
      for (int t0 = me * EPI_TOK; t0 < TN; t0 += S * EPI_TOK) {
        if (tok0 + t0 >= p.M_tok) break;               // padding tokens of the last token tile
        float a[EPI_TOK], b[EPI_TOK];
#pragma unroll
        for (int j = 0; j < EPI_TOK; ++j) a[j] = b[j] = 0.f;
#pragma unroll 2
        for (int c = c_first; c < c_first + S; ++c) {   // fixed order: bit-reproducible sums
          const float* part = p.ws_part + (int64_t)c * slot_elems;
          float va[EPI_TOK], vb[EPI_TOK];
#pragma unroll
          for (int j = 0; j < EPI_TOK; ++j) {           // 16 independent L2 loads in flight
            va[j] = __ldcg(part + (int64_t)(t0 + j) * 128 + r);
            vb[j] = __ldcg(part + (int64_t)(TN + t0 + j) * 128 + r);
          }
#pragma unroll
          for (int j = 0; j < EPI_TOK; ++j) { a[j] += va[j]; b[j] += vb[j]; }
        }
        tc_emit<T>(p, silu, r, a, b, rowA0, rowB0, tok0 + t0);
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
  tc_fence_before();
  __syncthreads();
  if (warp == WARP_MMA) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------------ host
TcEncodeFn tc_get_encode() {
  static TcEncodeFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = (TcEncodeFn)p;
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
// A launch is cut into work items (output tile, K split).  An output tile is a PAIR of row blocks of R <= 128 weight rows
// each (for SILU_MUL: R gate rows and the matching R up rows) times one token tile; R is chosen per matrix so that the tiles
// fill the SMs evenly with WHOLE tiles wherever the matrix has enough rows (gate|up of Llama-3-8B: R = 104 -> 138 tiles on
// 148 SMs; lm_head: R = 88 -> 5 tiles per SM) -- no partial tiles, the fused epilogue goes straight from TMEM to global
// memory.  Only matrices with too few rows for that (o_proj, down_proj, q|k|v at decode) split K S ways, S * tiles <= #SMs,
// and sum the S partial tiles through the workspace (kernel tail).
struct TcPlan {
  int TN, tiles_m, pairs, R, S, items, grid, w_stages, x_slots, smem_bytes, nbuf, a_slots;
};

static int env_int(const char* name, int dflt) { const char* e = getenv(name); return e ? atoi(e) : dflt; }
static const int g_force_R = env_int("B200P_TC_R", 0);       // experiments: force the plan
static const int g_force_S = env_int("B200P_TC_S", 0);
static const int g_rotate = env_int("B200P_TC_ROT", 1);
static const int g_force_nbuf = env_int("B200P_TC_NBUF", 0);

static TcPlan make_plan(int64_t M, int64_t N, int64_t K, int bits, int epi) {
  TcPlan pl;
  const int tn = (int)(M >= MAX_TN ? MAX_TN : ((M + 15) / 16) * 16);
  pl.TN = tn;
  pl.tiles_m = (int)((M + tn - 1) / tn);
  const int kps = bits == 4 ? 4 : 2;                       // k-blocks per 128-byte stage row
  const int num_st = (int)((K / 64 + kps - 1) / kps);
  const int sms = sm_count();
  const bool silu = epi == B200P_EPI_SILU_MUL;
  const int64_t rows = silu ? N / 2 : N;                   // rows to deal out (silu: gate rows; each brings its up row)
  // cost model in units of (weight rows x stages) streamed by the busiest CTA.  A stage never costs less than ~96 rows'
  // worth of time (the dequant -> tcgen05.st -> MMA chain of one k-block is latency-bound below that), lanes of a
  // partly filled warp still cost issue slots, and a K split pays for writing + re-reading its fp32 partial tile.
  double best = 1e30;
  pl.R = 128; pl.S = 1;
  for (int R = 8; R <= 128; R += 8) {
    if (g_force_R && R != g_force_R) continue;
    const int64_t pairs = silu ? (rows + R - 1) / R : (rows + 2 * R - 1) / (2 * R);
    const int64_t tiles = pairs * pl.tiles_m;
    const int max_s = tiles <= sms ? (int)(sms / tiles) : 1;
    for (int S = 1; S <= max_s && S <= num_st && S <= 16; ++S) {
      if (g_force_S && S != g_force_S) continue;
      const int64_t items = tiles * S;
      const int64_t per_cta = (items + sms - 1) / sms;
      const int st_max = (num_st + S - 1) / S;
      const int r32 = ((R + 31) / 32) * 32;
      double row_cost = 0.5 * (2 * R) + 0.5 * (2 * r32);
      if (row_cost < 192.0) row_cost = 192.0;
      double cost = (double)per_cta * st_max * row_cost;
      if (S > 1) cost += 16.0 * tn + 256.0;
      cost += 64.0 * per_cta;                              // per-tile epilogue
      if (cost < best) { best = cost; pl.R = R; pl.S = S; }
    }
  }
  const int64_t pairs = silu ? (rows + pl.R - 1) / pl.R : (rows + 2 * pl.R - 1) / (2 * pl.R);
  pl.pairs = (int)pairs;
  const int64_t items = pairs * pl.tiles_m * pl.S;
  pl.items = (int)items;
  pl.grid = (int)(items > sms ? sms : (items < 1 ? 1 : items));
  // TMEM: D buffers first, the A ring takes the rest.  A second D buffer only pays when a CTA runs several items
  // (the epilogue of one overlaps the main loop of the next); with one item per CTA the columns deepen the A ring.
  const int d_stride = tn <= 32 ? 32 : (tn <= 64 ? 64 : 128);
  pl.nbuf = (tn > 64 || items <= sms) ? 1 : 2;
  if (g_force_nbuf && tn <= 64) pl.nbuf = g_force_nbuf;
  pl.a_slots = ((512 - pl.nbuf * 2 * d_stride) / 64) & ~1;
  if (pl.a_slots > NSLOT) pl.a_slots = NSLOT;
  // X ring: one [TN x 64] tile per k-block, 8 slots (4 at TN = 128) -- about 1 us of MMA work in flight, enough to
  // cover the L2 latency of the tile loads; the weight ring gets the rest of shared memory
  pl.x_slots = tn * 128 * 8 <= X_RING_BYTES ? 8 : 4;
  const int stage_bytes = 2 * pl.R * 128;
  int st = (SMEM_BUDGET - SMEM_FIXED - pl.x_slots * tn * 128) / stage_bytes;
  if (st > MAX_STAGES) st = MAX_STAGES;
  pl.w_stages = st;
  pl.smem_bytes = SMEM_FIXED + pl.x_slots * tn * 128 + st * stage_bytes;
  return pl;
}

bool qmm_tc_supported(int64_t M, int64_t N, int64_t K, int gs, int bits, int dtype, int epi) {
  if (gs != 64) return false;
  if (dtype != B200P_F16 && dtype != B200P_BF16) return false;
  if (M < g_min_tc_m) return false;
  if (K % 64 != 0 || (K * bits / 8) % 16 != 0) return false;
  if (N % 8 != 0) return false;
  if (epi == B200P_EPI_SILU_MUL && (N / 2) % 8 != 0) return false;
  if (N >= (1ll << 31) || M >= (1ll << 31)) return false;
  return true;
}

// {arrived, done} per split tile: K splits are planned only when tiles * S <= #SMs, so there are never more than #SMs of them
static size_t ws_counter_bytes(int64_t, int64_t) { return 4096; }

size_t qmm_tc_workspace_bytes(int64_t M, int64_t N, int64_t K) {
  (void)K;
  if (M < 1) return 0;
  const int64_t tn = M >= MAX_TN ? MAX_TN : ((M + 15) / 16) * 16;
  // per-tile counters (ZERO on first use; the kernel leaves them zero) + one fp32 [2*TN x 128] partial-tile slot per CTA
  // (no initial state)
  return ws_counter_bytes(M, N) + (size_t)sm_count() * 2 * tn * 128 * sizeof(float);
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

// Tensor maps are pure functions of (base, dims, stride, box, type, swizzle): encode each distinct one once per process
// (a decode step replays a graph and never gets here; an eager prefill GEMM used to pay four driver calls per launch).
struct TmKey {
  const void* base; uint64_t d0, d1, stride; uint32_t b0, b1; int dt, sw;
  bool operator==(const TmKey& o) const {
    return base == o.base && d0 == o.d0 && d1 == o.d1 && stride == o.stride && b0 == o.b0 && b1 == o.b1 && dt == o.dt && sw == o.sw;
  }
};
struct TmKeyHash {
  size_t operator()(const TmKey& k) const {
    uint64_t h = (uint64_t)(uintptr_t)k.base * 0x9E3779B97F4A7C15ull;
    for (uint64_t v : {k.d0, k.d1, k.stride, (uint64_t)k.b0, (uint64_t)k.b1, (uint64_t)k.dt, (uint64_t)k.sw})
      h = (h ^ v) * 0x100000001B3ull + (h >> 29);
    return (size_t)h;
  }
};
static std::mutex g_tm_mutex;
static std::unordered_map<TmKey, CUtensorMap, TmKeyHash> g_tm_cache;

static int encode_2d_raw(TcEncodeFn enc, CUtensorMap* tm, CUtensorMapDataType dt, const void* base, uint64_t d0, uint64_t d1,
                         uint64_t stride_bytes, uint32_t b0, uint32_t b1, CUtensorMapSwizzle sw, const char* what);

int tc_encode_2d(TcEncodeFn enc, CUtensorMap* tm, CUtensorMapDataType dt, const void* base, uint64_t d0, uint64_t d1,
                     uint64_t stride_bytes, uint32_t b0, uint32_t b1, CUtensorMapSwizzle sw, const char* what) {
  const TmKey key{base, d0, d1, stride_bytes, b0, b1, (int)dt, (int)sw};
  {
    std::lock_guard<std::mutex> lk(g_tm_mutex);
    auto it = g_tm_cache.find(key);
    if (it != g_tm_cache.end()) { *tm = it->second; return B200P_OK; }
  }
  const int rc = encode_2d_raw(enc, tm, dt, base, d0, d1, stride_bytes, b0, b1, sw, what);
  if (rc) return rc;
  std::lock_guard<std::mutex> lk(g_tm_mutex);
  if (g_tm_cache.size() > 8192) g_tm_cache.clear();
  g_tm_cache.emplace(key, *tm);
  return B200P_OK;
}

static int encode_2d_raw(TcEncodeFn enc, CUtensorMap* tm, CUtensorMapDataType dt, const void* base, uint64_t d0, uint64_t d1,
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
  TcEncodeFn enc = tc_get_encode();
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
  int rc = tc_encode_2d(enc, &tms[0], CU_TENSOR_MAP_DATA_TYPE_UINT8, w, (uint64_t)(K * bits / 8), (uint64_t)N,
                     (uint64_t)(K * bits / 8), 128u, (uint32_t)pl.R, CU_TENSOR_MAP_SWIZZLE_128B, "weights");
  if (rc) return rc;
  rc = tc_encode_2d(enc, &tms[1], dt16, x, (uint64_t)K, (uint64_t)M, (uint64_t)(ldx * 2), 64u, (uint32_t)pl.TN,
                 CU_TENSOR_MAP_SWIZZLE_128B, "x");
  if (rc) return rc;
  if (sb_tma) {
    rc = tc_encode_2d(enc, &tms[2], dt16, s, (uint64_t)num_kb, (uint64_t)N, (uint64_t)num_kb * 2, 8u, (uint32_t)pl.R,
                   CU_TENSOR_MAP_SWIZZLE_NONE, "scales");
    if (rc) return rc;
    rc = tc_encode_2d(enc, &tms[3], dt16, b, (uint64_t)num_kb, (uint64_t)N, (uint64_t)num_kb * 2, 8u, (uint32_t)pl.R,
                   CU_TENSOR_MAP_SWIZZLE_NONE, "biases");
    if (rc) return rc;
  } else {
    tms[2] = tms[0];
    tms[3] = tms[0];
  }
  TcParams p;
  p.scales = s; p.biases = b; p.y = y;
  const size_t counters = ws_counter_bytes(M, N);
  const size_t need = counters + (pl.S > 1 ? (size_t)pl.grid * 2 * pl.TN * 128 * sizeof(float) : 0);
  B200P_REQUIRE(ws_bytes >= need && aligned(ws, 16), B200P_ERR_ARG, "qmm workspace too small (%zu < %zu)", ws_bytes, need);
  {
    const bool f32out = epi == B200P_EPI_STORE_F32 || epi == B200P_EPI_ADD_RESIDUAL;
    B200P_REQUIRE(aligned(y, 4) && (f32out || ldy % 2 == 0), B200P_ERR_ALIGN, "y rows must be 4-byte aligned for the tensor path (ldy=%lld)",
                  (long long)ldy);
  }
  p.ws_counter = reinterpret_cast<int*>(ws);
  p.ws_part = reinterpret_cast<float*>(reinterpret_cast<char*>(ws) + counters);
  p.M_tok = (int)M; p.N = (int)N; p.K = (int)K; p.TN = pl.TN; p.ldy = ldy; p.epi = epi; p.y_perm = y_perm ? 1 : 0;
  B200P_REQUIRE((int64_t)pl.pairs * pl.tiles_m * pl.S < (1ll << 30), B200P_ERR_ARG, "problem too large for the tensor path");
  p.num_kb = num_kb; p.pairs = pl.pairs; p.tiles_m = pl.tiles_m; p.R = pl.R; p.S = pl.S; p.items = pl.items; p.rotate = g_rotate; p.nbuf = pl.nbuf; p.a_slots = pl.a_slots;
  p.num_st = (p.num_kb + (bits == 4 ? 4 : 2) - 1) / (bits == 4 ? 4 : 2);
  p.sb_tma = sb_tma ? 1 : 0;
  p.w_stages = pl.w_stages; p.x_slots = pl.x_slots;
  if (dtype == B200P_F16) {
    if (bits == 8) return fast ? launch_tc<__half, 8, false, true>(tms, p, pl, st) : launch_tc<__half, 8, false, false>(tms, p, pl, st);
    if (fast) return x_perm ? launch_tc<__half, 4, true, true>(tms, p, pl, st) : launch_tc<__half, 4, false, true>(tms, p, pl, st);
    return x_perm ? launch_tc<__half, 4, true, false>(tms, p, pl, st) : launch_tc<__half, 4, false, false>(tms, p, pl, st);
  } else {
    if (bits == 8) return launch_tc<__nv_bfloat16, 8, false, false>(tms, p, pl, st);
    return x_perm ? launch_tc<__nv_bfloat16, 4, true, false>(tms, p, pl, st) : launch_tc<__nv_bfloat16, 4, false, false>(tms, p, pl, st);
  }
}

}  // namespace b200p

#if B200P_TC_TIMING
extern "C" int b200p_debug_tc_timing(unsigned long long* out32) {
  return cudaMemcpyFromSymbol(out32, b200p::g_tc_timing, sizeof(unsigned long long) * 32) == cudaSuccess ? 0 : -1;
}
#endif
