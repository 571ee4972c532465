// Small-batch (M <= 16) fused-dequant GEMV of b200p_qmm for 4-bit group-64 weights and 16-bit activations:
//
//   y[m, n] (+)= sum_k x[m, k] * (q[n,k] * s[n, k/64] + b[n, k/64])
//
// Why this shape of kernel.  At M <= 16 the step is a pure weight stream (AI = 3.56*M flop/B).  The ALU budget at HBM speed is
// ~2.8 lane-ops per weight (148 SMs x 128 lanes x 1.9 GHz / 13e12 weights/s); a CUDA-core GEMV needs >= 2 ops to unpack a
// code to fp32 plus M FMAs (the old qmv kernel: 1.4-2.2 TB/s at M = 1, collapsing with M -- profiles/r1_run11_*), while the
// dequant itself costs 1.5 ops per weight when done on packed half2 (LOP3 + exponent magic + one HFMA2 for q*s+b).  So the
// dot products go to the tensor cores -- but with NO shared-memory / TMEM / mbarrier pipeline in front of them: every warp
// streams its own 16 weight rows global -> registers with 128-bit no-allocate loads, dequantises in registers and feeds
// mma.sync.m16n8k16 directly (weights = the 16-row A operand, the <= 8 tokens of an n-tile = the B operand, fp32 accumulators
// in registers).  Legacy mma.sync is the right instruction here: it takes register operands, needs ~0.3 G issues/s per SM, and
// the kernel has no fixed setup cost -- the tcgen05 tile kernel (qmm_tc.cu) pays ~7 us per launch before its first MMA.
//
// Mapping.  Lane (g = lane/4, t = lane%4) of a warp owns rows r0 = 16*rg + g and r0 + 8.  Per step of 256 k it loads the 32
// packed bytes [64t, 64t+64) of both rows (the four lanes of a row read one 128-byte line; the 64 codes are ONE quantisation
// group, so the lane needs one scale and one bias per row per step).  A 4-bit word unpacks with two masks into the pairs
// (k0,k4)(k1,k5)(k2,k6)(k3,k7); pairs 0,1 are the lane's k-slots (2t,2t+1),(2t+8,2t+9) of one MMA, pairs 2,3 of the next --
// the MMA's k index is just a label, as long as the B operand uses the same one: lane (g,t) supplies x[token g][same k], which
// in the PERM8 activation layout is exactly one 16-byte chunk per word (natural layout: 4 PRMT per word to re-pair).
// The D fragment leaves every (row, token) sum in exactly one lane: no shuffle reduction.
// Split-K: K is cut into `splits` ranges so that ~16 warps per SM are in flight even for N = 4096; partial sums are combined
// with fp32 red.global.add into y (ADD_RESIDUAL: y already holds the residual / the zeroed accumulator).
//
// Algorithmic HBM bytes per launch: N*K/2 + 2*N*(K/64)*2.
#include "common.cuh"
#include "dequant.cuh"

namespace b200p {

int sm_count();

template <typename T>
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  if constexpr (std::is_same<T, __half>::value) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  } else {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
}

constexpr int QM_WARPS = 4;
constexpr int QM_STEP_K = 256;

// NT: n-tiles of 8 tokens (M <= 8*NT).  MODE 0: plain fp32 store; 1: combine with red.add (ADD_RESIDUAL / split-K);
// 2: SILU_MUL -- the warp's second row block is not "rows + 8" but the UP rows that match its 8 GATE rows (row n of gate_proj
// is row N/2 + n of the fused matrix), so every lane holds the gate and the up sum of the same (row, token) and stores
// silu(g) * u as one 16-bit value, in the PERM8 column order if y_perm; never K-split.
template <typename T, int NT, bool XPERM, int MODE>
__global__ void __launch_bounds__(QM_WARPS * 32)
qmv_mma_kernel(const T* __restrict__ x, const uint32_t* __restrict__ w, const T* __restrict__ scales,
               const T* __restrict__ biases, float* __restrict__ y, int M, int64_t N, int64_t K, int64_t ldx, int64_t ldy,
               int splits, int steps_per_split, int y_perm) {
  constexpr bool ATOMIC = MODE == 1;
  constexpr bool SILU = MODE == 2;
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int64_t wid = (int64_t)blockIdx.x * QM_WARPS + (threadIdx.x >> 5);
  const int64_t rg = wid / splits;
  const int ks = (int)(wid - rg * splits);
  if (SILU ? rg * 8 >= N / 2 : rg * 16 >= N) return;
  const int64_t r0 = SILU ? rg * 8 + g : rg * 16 + g, r1 = SILU ? N / 2 + r0 : r0 + 8;
  const int64_t wrow = K / 8;                 // uint32 words per weight row
  const int64_t ng = K / 64;
  const int64_t total_steps = K / QM_STEP_K;
  const int64_t s_begin = (int64_t)ks * steps_per_split;
  const int64_t s_end = min(total_steps, s_begin + steps_per_split);

  float acc[NT][4];
#pragma unroll
  for (int n = 0; n < NT; ++n)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[n][j] = 0.f;

  const uint32_t* w0 = w + r0 * wrow + 8 * t;     // 8 words = 64 codes per lane per step
  const uint32_t* w1 = w + r1 * wrow + 8 * t;
  const T* sc0 = scales + r0 * ng + t;
  const T* sc1 = scales + r1 * ng + t;
  const T* bi0 = biases + r0 * ng + t;
  const T* bi1 = biases + r1 * ng + t;
  const T* xr[NT];
  bool xok[NT];
#pragma unroll
  for (int n = 0; n < NT; ++n) {
    const int tok = n * 8 + g;
    xok[n] = tok < M;
    xr[n] = x + (int64_t)(xok[n] ? tok : 0) * ldx + 64 * t;
  }

  // weights, scales and biases never depend on the previous kernel: fetch the first step before griddepcontrol.wait
  uint4 qa[2], qb[2];
  uint32_t s2a = 0, b2a = 0, s2b = 0, b2b = 0;
  auto fetch_w = [&](int64_t s) {
    const int64_t wo = s * (QM_STEP_K / 8);
    ldg_stream256(w0 + wo, qa[0], qa[1]);       // the lane's 64 codes of the step = one 32-byte sector per row
    ldg_stream256(w1 + wo, qb[0], qb[1]);
    s2a = dup16(__ldg(sc0 + 4 * s));
    b2a = dup16(__ldg(bi0 + 4 * s));
    s2b = dup16(__ldg(sc1 + 4 * s));
    b2b = dup16(__ldg(bi1 + 4 * s));
  };
  if (s_begin < s_end) fetch_w(s_begin);
  pdl_wait();                                  // x is the previous kernel's output; y its accumulator

  for (int64_t s = s_begin; s < s_end; ++s) {
    const uint4 ca[2] = {qa[0], qa[1]}, cb[2] = {qb[0], qb[1]};
    const uint32_t cs2a = s2a, cb2a = b2a, cs2b = s2b, cb2b = b2b;
    // activations of this step: 64 k of the lane's token per n-tile, 8 chunks of 16 bytes
    uint4 xv[NT][8];
#pragma unroll
    for (int n = 0; n < NT; ++n) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        if (xok[n]) xv[n][i] = *reinterpret_cast<const uint4*>(xr[n] + s * QM_STEP_K + 8 * i);
        else xv[n][i] = make_uint4(0u, 0u, 0u, 0u);
      }
    }
    if (s + 1 < s_end) fetch_w(s + 1);          // next step's weights are in flight during this step's math
    const uint32_t* pa = reinterpret_cast<const uint32_t*>(ca);
    const uint32_t* pb = reinterpret_cast<const uint32_t*>(cb);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      uint32_t da[4], db[4];
      dequant8<T, 4, true>(pa + i, cs2a, cb2a, da);      // pairs (k0,k4) (k1,k5) (k2,k6) (k3,k7) of word i, row r0
      dequant8<T, 4, true>(pb + i, cs2b, cb2b, db);      // same, row r0 + 8
      const uint32_t a_lo[4] = {da[0], db[0], da[1], db[1]};
      const uint32_t a_hi[4] = {da[2], db[2], da[3], db[3]};
#pragma unroll
      for (int n = 0; n < NT; ++n) {
        uint4 u = xv[n][i];
        if constexpr (!XPERM) {                           // natural order -> (x0,x4) (x1,x5) (x2,x6) (x3,x7)
          const uint4 v = u;
          u.x = prmt(v.x, v.z, 0x5410u);
          u.y = prmt(v.x, v.z, 0x7632u);
          u.z = prmt(v.y, v.w, 0x5410u);
          u.w = prmt(v.y, v.w, 0x7632u);
        }
        mma16816<T>(acc[n], a_lo, u.x, u.y);
        mma16816<T>(acc[n], a_hi, u.z, u.w);
      }
    }
  }
  pdl_launch_dependents();
  // D fragment: acc[n][0], [1] = (row r0, tokens 8n + 2t, 2t+1); acc[n][2], [3] = (row r0 + 8, same tokens)
#pragma unroll
  for (int n = 0; n < NT; ++n) {
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int tok = n * 8 + 2 * t + j;
      if (tok < M) {
        if constexpr (SILU) {
          const float gte = acc[n][j], up = acc[n][2 + j];
          T* yt = reinterpret_cast<T*>(y) + (int64_t)tok * ldy;
          yt[y_perm ? perm8_pos((int)r0) : (int)r0] = Tr<T>::from_f(gte / (1.0f + __expf(-gte)) * up);   // silu_tc of tc_ptx.cuh
          continue;
        }
        float* yr = y + (int64_t)tok * ldy;
        if (ATOMIC) {
          atomicAdd(yr + r0, acc[n][j]);
          atomicAdd(yr + r1, acc[n][2 + j]);
        } else {
          yr[r0] = acc[n][j];
          yr[r1] = acc[n][2 + j];
        }
      }
    }
  }
}

bool qmv_mma_supported(int64_t M, int64_t N, int64_t K, int gs, int bits, int dtype, int epi) {
  if (bits != 4 || gs != 64) return false;
  if (dtype != B200P_F16 && dtype != B200P_BF16) return false;
  if (M < 1 || M > 16) return false;
  if (K % QM_STEP_K != 0 || N % 16 != 0) return false;
  if (epi == B200P_EPI_SILU_MUL) return true;                        // N % 16 == 0 -> N / 2 gate rows in whole groups of 8
  return epi == B200P_EPI_STORE_F32 || epi == B200P_EPI_ADD_RESIDUAL;
}

template <typename T>
static int launch_qmv_mma(const void* x, const uint32_t* w, const void* s, const void* b, void* y, int64_t M, int64_t N,
                          int64_t K, int64_t ldx, int64_t ldy, int epi, bool xperm, bool yperm, cudaStream_t st) {
  const bool silu = epi == B200P_EPI_SILU_MUL;
  const int64_t row_groups = silu ? N / 2 / 8 : N / 16;          // SILU_MUL: a warp takes 8 gate rows and their 8 up rows
  const int64_t total_steps = K / QM_STEP_K;
  // Split-K so that every SM holds ~16 warps, but never a grid that spills a few CTAs into a second wave (4096 x 14336 at ten
  // splits was 640 CTAs on 592 slots: 1.08 waves, 30 % slower than 8 splits in one wave): among the split counts pick the
  // one with the lowest waves x (steps per warp + ~1.5 steps of fixed latency); ties go to fewer splits (fewer red.adds).  Plain stores cannot
  // be split.  (More weights in flight per lane instead of more warps was tried and is slower:
  // profiles/r1_run33_gemv_prefetch_depth_experiment.txt.)
  int64_t splits = 1;
  if (epi == B200P_EPI_ADD_RESIDUAL) {
    const int64_t capacity = (int64_t)sm_count() * 4;            // resident CTAs: 4 per SM at 121 registers x 128 threads
    const int64_t max_splits = total_steps / 2 > 1 ? total_steps / 2 : 1;   // at least two steps per warp
    double best = 1e30;
    for (int64_t sp = 1; sp <= max_splits; ++sp) {
      const int64_t sps_c = (total_steps + sp - 1) / sp;
      const int64_t sp_eff = (total_steps + sps_c - 1) / sps_c;
      const int64_t ctas = (row_groups * sp_eff + QM_WARPS - 1) / QM_WARPS;
      // a partly filled wave after the first costs its fill fraction, but never less than half a wave (measured: 28672 x 4096 runs
      // 24.1 us as 1.51 waves of 8 steps against 25.5 us as 0.76 wave of 16; 4096 x 14336 16.9 us as 1.08 waves of 6 steps
      // against 12.9 us as 0.86 wave of 7 -- profiles/r1_run34_gemv_splits.txt)
      const int64_t full = ctas / capacity;
      const double frac = (double)(ctas - full * capacity) / (double)capacity;
      const double waves = full == 0 ? 1.0 : (double)full + (frac > 0.0 ? (frac > 0.5 ? frac : 0.5) : 0.0);
      const double cost = waves * ((double)sps_c + 1.5) * (full == 0 ? 1.0 : 1.1);   // near-ties go to the single wave
      if (cost < best - 1e-9) { best = cost; splits = sp_eff; }
    }
  }
  const int64_t sps = (total_steps + splits - 1) / splits;
  splits = (total_steps + sps - 1) / sps;
  const int64_t warps = row_groups * splits;
  dim3 grid((unsigned)((warps + QM_WARPS - 1) / QM_WARPS)), block(QM_WARPS * 32);
  const int mode = silu ? 2 : (epi == B200P_EPI_ADD_RESIDUAL ? 1 : 0);
#define GO(NT_, XP_, MD_)                                                                                                \
  launch_pdl(qmv_mma_kernel<T, NT_, XP_, MD_>, grid, block, 0, st, (const T*)x, w, (const T*)s, (const T*)b, (float*)y,   \
             (int)M, N, K, ldx, ldy, (int)splits, (int)sps, yperm ? 1 : 0)
#define GO_MODE(NT_, XP_) do { if (mode == 2) GO(NT_, XP_, 2); else if (mode == 1) GO(NT_, XP_, 1); else GO(NT_, XP_, 0); } while (0)
  if (M <= 8) {
    if (xperm) GO_MODE(1, true); else GO_MODE(1, false);
  } else {
    if (xperm) GO_MODE(2, true); else GO_MODE(2, false);
  }
#undef GO_MODE
#undef GO
  B200P_LAUNCH_CHECK("qmv_mma_kernel");
  return B200P_OK;
}

int qmv_mma_launch(const void* x, const uint32_t* w, const void* s, const void* b, void* y, int64_t M, int64_t N, int64_t K,
                   int64_t ldx, int64_t ldy, int dtype, int epi, int flags, cudaStream_t st) {
  B200P_REQUIRE(aligned(s, 2) && aligned(b, 2) && aligned(y, epi == B200P_EPI_SILU_MUL ? 2 : 4) && aligned(w, 32), B200P_ERR_ALIGN,
                "bad alignment (weights: 32 bytes)");
  const bool xperm = (flags & B200P_X_PERM8) != 0, yperm = (flags & B200P_Y_PERM8) != 0;
  if (dtype == B200P_F16) return launch_qmv_mma<__half>(x, w, s, b, y, M, N, K, ldx, ldy, epi, xperm, yperm, st);
  return launch_qmv_mma<__nv_bfloat16>(x, w, s, b, y, M, N, K, ldx, ldy, epi, xperm, yperm, st);
}

}  // namespace b200p
