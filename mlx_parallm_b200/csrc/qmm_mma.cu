// Pipeline-free tensor-core GEMM of b200p_qmm for SMALL 4-bit group-64 matrices at a FEW rows (fp32 accumulate epilogue):
//
//   y[m, n] += sum_k x[m, k] * (q[n,k] * s[n, k/64] + b[n, k/64])
//
// Why a third GEMM kernel.  The tcgen05 tile kernels (qmm_tc.cu, qmm_sk.cu) pay ~5 us per launch before and after their
// streaming part (TMEM allocation, ~50 mbarriers, tensor-map fetches, filling a TMA -> dequant -> TMEM -> MMA pipeline whose
// k-block time is ~620 cycles of round-trip latency): on a 9-14 MB matrix (q|k|v, o_proj) or on a 1-7 MB tensor-parallel shard
// that IS the launch.  This kernel is the register-level mma.sync GEMV of qmv_mma.cu grown to 64 rows: every warp streams its
// own 16 weight rows global -> registers (256-bit no-allocate loads, next step in flight during this step's math), dequantises
// in registers (dequant.cuh) and feeds mma.sync.m16n8k16 directly; no TMEM, no mbarriers, no tensor maps.  What changes at 64
// rows is the B operand: 8 n-tiles of 8 tokens would be 32 KB of activations per warp and step from L2, so the CTA's 8 warps
// share ONE [8*NT tokens x 256 k] tile per step in shared memory (cp.async, double-buffered, 16-byte chunks XOR-swizzled so
// that the fragment loads of a quarter warp hit 8 different bank groups).
// Legacy mma.sync runs at 1020 MAC/cycle/SM on sm_100a (scripts/probe_hmma.cu, profiles/r2_run34_probe_hmma.txt): 128 weights
// per cycle and SM at 8 rows, 16 at 64 rows.  Measured against the tile kernels (profiles/r2_run35_small_gemm_paths.txt, us per
// launch in a CUDA graph over weight copies larger than L2): q|k|v 6144 x 4096 6.2 vs 7.2 at 8 rows, 6.9 vs 7.3 at 16, 9.0 vs
// 7.9 at 32, 14.6 vs 8.9 at 64 (640 MAC/cycle/SM: bound by the mma.sync rate and the fragment loads); o_proj 4.3 vs 5.7 at 8
// rows; TP=8 shards 2.6-4.1 vs 3.8-5.3.  The auto rule in c_api.cu therefore gives it 5..16 rows on matrices up to 32 M weights
// (B200P_MMA_MAX_M, B200P_MMA_MAX_WEIGHTS); the kernel itself runs up to 64 rows (tests, path 4).
//
// Mapping: CTA = (block of 128 weight rows, K split); warp w owns rows 16w + g and 16w + g + 8 (lane = 4g + t), per step of
// 256 k the lane holds the 64 codes [64t, 64t + 64) of both rows = one quantisation group each.  Fragment labelling as in
// qmv_mma.cu: the MMA's k index is a label shared by A and B, so word i of the lane pairs with the 16-byte activation chunk
// 8t + i of token 8n + g (PERM8 layout: as is; natural layout: 4 PRMT).
// Epilogue: the [tokens x 128 rows] partial tile goes to shared memory and leaves as ONE cp.reduce.async.bulk (.add.f32) of
// 512 bytes per token -- an SM retires scalar fp32 reds at ~1.3 cycles per lane, 8192 of them per CTA would cost more than the
// main loop.  Sums of different K splits meet in L2 in arrival order: results may differ in the last fp32 bit between runs
// (like the stream-K kernel's).
//
// Algorithmic HBM bytes per launch: N*K/2 + 2*N*(K/64)*2.
#include <stdlib.h>

#include "common.cuh"
#include "dequant.cuh"
#include "mma_util.cuh"

namespace b200p {

int sm_count();
bool pdl_enabled_gemm();

constexpr int QG_WARPS = 8;
constexpr int QG_ROWS = QG_WARPS * 16;            // weight rows per CTA
constexpr int QG_STEP_K = 256;
constexpr int QG_ROW_BYTES = QG_STEP_K * 2;       // one token's k range of a step: 32 chunks of 16 bytes
constexpr int QG_D_STRIDE = QG_ROWS + 4;          // floats per token line of the staged partial tile (bank spread, 16-byte multiple)

// physical 16-byte chunk of logical chunk c (0..31) in token line `row`: the chunk-in-segment index is XOR-ed with the segment
// (= lane's t) and with the row parity, so the 8 lanes of a quarter warp (2 tokens x 4 t) read 8 different bank groups
__device__ __forceinline__ uint32_t qg_chunk(uint32_t c, uint32_t row) {
  return (c & ~7u) | ((c ^ (c >> 3) ^ ((row & 1u) << 2)) & 7u);
}

__device__ __forceinline__ uint4 qg_lds128(uint32_t addr) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(addr));
  return r;
}
__device__ __forceinline__ void bulk_reduce_add_f32(float* dst, uint32_t src_smem, uint32_t bytes) {
  asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;" ::"l"(dst), "r"(src_smem), "r"(bytes)
               : "memory");
}

template <typename T, int NT, bool XPERM>
__global__ void __launch_bounds__(QG_WARPS * 32, 2)
qmm_mma_kernel(const T* __restrict__ x, const uint32_t* __restrict__ w, const T* __restrict__ scales,
               const T* __restrict__ biases, float* __restrict__ y, int M, int64_t N, int64_t K, int64_t ldx, int64_t ldy,
               int splits, int steps_per_split) {
  extern __shared__ __align__(128) uint8_t smem[];
  constexpr int TOK = NT * 8;
  constexpr uint32_t XBUF = TOK * QG_ROW_BYTES;
  const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(smem);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int64_t rb = blockIdx.x / splits;
  const int ks = (int)(blockIdx.x - rb * splits);
  const int64_t r0 = rb * QG_ROWS + warp * 16 + g, r1 = r0 + 8;
  const int64_t wrow = K / 8, ng = K / 64;
  const int64_t total_steps = K / QG_STEP_K;
  const int64_t s_begin = (int64_t)ks * steps_per_split;
  const int64_t s_end = min(total_steps, s_begin + steps_per_split);

  float acc[NT][4];
#pragma unroll
  for (int n = 0; n < NT; ++n)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[n][j] = 0.f;

  const uint32_t* w0 = w + r0 * wrow + 8 * t;
  const uint32_t* w1 = w + r1 * wrow + 8 * t;
  const T* sc0 = scales + r0 * ng + t;
  const T* sc1 = scales + r1 * ng + t;
  const T* bi0 = biases + r0 * ng + t;
  const T* bi1 = biases + r1 * ng + t;

  uint4 qa[2], qb[2];
  uint32_t s2a = 0, b2a = 0, s2b = 0, b2b = 0;
  auto fetch_w = [&](int64_t s) {
    const int64_t wo = s * (QG_STEP_K / 8);
    ldg_stream256(w0 + wo, qa[0], qa[1]);
    ldg_stream256(w1 + wo, qb[0], qb[1]);
    s2a = dup16(__ldg(sc0 + 4 * s));
    b2a = dup16(__ldg(bi0 + 4 * s));
    s2b = dup16(__ldg(sc1 + 4 * s));
    b2b = dup16(__ldg(bi1 + 4 * s));
  };
  // the activation tile of step s -> buffer `buf`: TOK token lines x 32 chunks, tokens >= M zero-filled
  auto issue_x = [&](int64_t s, int buf) {
    const uint32_t dst0 = sbase + (uint32_t)buf * XBUF;
#pragma unroll
    for (int i = 0; i < TOK * 32 / (QG_WARPS * 32); ++i) {
      const uint32_t idx = (uint32_t)i * (QG_WARPS * 32) + threadIdx.x;
      const uint32_t row = idx >> 5, c = idx & 31u;
      const bool ok = (int)row < M;
      const T* src = x + (int64_t)(ok ? row : 0) * ldx + s * QG_STEP_K + c * 8;
      cp_async16(dst0 + row * QG_ROW_BYTES + (qg_chunk(c, row) << 4), src, ok);
    }
  };

  // weights, scales and biases never depend on the previous kernel: fetch the first step before griddepcontrol.wait
  if (s_begin < s_end) fetch_w(s_begin);
  pdl_wait();                                  // x is the previous kernel's output; y its accumulator
  if (s_begin < s_end) issue_x(s_begin, 0);
  cp_async_commit();

  int buf = 0;
  for (int64_t s = s_begin; s < s_end; ++s, buf ^= 1) {
    if (s + 1 < s_end) issue_x(s + 1, buf ^ 1);
    cp_async_commit();
    cp_async_wait<1>();                        // this step's tile has landed (the next one may still be in flight)
    __syncthreads();
    const uint4 ca[2] = {qa[0], qa[1]}, cb[2] = {qb[0], qb[1]};
    const uint32_t cs2a = s2a, cb2a = b2a, cs2b = s2b, cb2b = b2b;
    if (s + 1 < s_end) fetch_w(s + 1);         // next step's weights are in flight during this step's math
    const uint32_t* pa = reinterpret_cast<const uint32_t*>(ca);
    const uint32_t* pb = reinterpret_cast<const uint32_t*>(cb);
    const uint32_t xb = sbase + (uint32_t)buf * XBUF + (uint32_t)g * QG_ROW_BYTES;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      uint32_t da[4], db[4];
      dequant8<T, 4, true>(pa + i, cs2a, cb2a, da);      // pairs (k0,k4) (k1,k5) (k2,k6) (k3,k7) of word i, row r0
      dequant8<T, 4, true>(pb + i, cs2b, cb2b, db);      // same, row r0 + 8
      const uint32_t a_lo[4] = {da[0], db[0], da[1], db[1]};
      const uint32_t a_hi[4] = {da[2], db[2], da[3], db[3]};
      // chunk 8t + i of token 8n + g: the row parity of 8n + g is the parity of g
      const uint32_t coff = qg_chunk((uint32_t)(8 * t + i), (uint32_t)g) << 4;
#pragma unroll
      for (int n = 0; n < NT; ++n) {
        uint4 u = qg_lds128(xb + (uint32_t)n * 8u * QG_ROW_BYTES + coff);
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
    __syncthreads();                           // everyone is done with this buffer before the step after next refills it
  }
  cp_async_wait<0>();
  pdl_launch_dependents();
  // ---- epilogue: partial tile -> shared memory [token][128 rows] -> one bulk reduce-add of 512 bytes per token ----------------
  // D fragment: acc[n][0], [1] = (row 16w + g, tokens 8n + 2t, 2t+1); acc[n][2], [3] = (row + 8, same tokens)
  float* dst = reinterpret_cast<float*>(smem);
#pragma unroll
  for (int n = 0; n < NT; ++n) {
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int tok = n * 8 + 2 * t + j;
      dst[tok * QG_D_STRIDE + warp * 16 + g] = acc[n][j];
      dst[tok * QG_D_STRIDE + warp * 16 + g + 8] = acc[n][2 + j];
    }
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // generic-proxy writes -> visible to the bulk copy engine
  __syncthreads();
  if ((int)threadIdx.x < M && (int)threadIdx.x < TOK) {
    bulk_reduce_add_f32(y + (int64_t)threadIdx.x * ldy + rb * QG_ROWS, sbase + (uint32_t)threadIdx.x * (QG_D_STRIDE * 4),
                        QG_ROWS * 4);
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");       // shared memory must outlive the engine's read, and the sums
                                                                    // are performed before the grid counts as complete
  }
}

// ------------------------------------------------------------------------------------------------ host
bool qmm_mma_supported(int64_t M, int64_t N, int64_t K, int gs, int bits, int dtype, int epi, const void* y, int64_t ldy) {
  if (bits != 4 || gs != 64) return false;
  if (dtype != B200P_F16 && dtype != B200P_BF16) return false;
  if (M < 1 || M > 64) return false;
  if (K % QG_STEP_K != 0 || N % QG_ROWS != 0) return false;
  if (!aligned(y, 16) || (ldy * 4) % 16 != 0) return false;          // bulk reduce: 16-byte aligned token lines
  return epi == B200P_EPI_ADD_RESIDUAL;
}

template <typename T, int NT>
static int launch_qmm_mma_nt(const void* x, const uint32_t* w, const void* s, const void* b, void* y, int64_t M, int64_t N,
                             int64_t K, int64_t ldx, int64_t ldy, bool xperm, cudaStream_t st) {
  const int64_t row_blocks = N / QG_ROWS;
  const int64_t total_steps = K / QG_STEP_K;
  // K splits: fill the resident CTA slots (2 per SM) once, at least two steps per CTA (one to hide the other's loads)
  const int64_t capacity = (int64_t)sm_count() * 2;
  int64_t splits = capacity / row_blocks;
  if (splits > total_steps / 2) splits = total_steps / 2;
  if (splits < 1) splits = 1;
  const int64_t sps = (total_steps + splits - 1) / splits;
  splits = (total_steps + sps - 1) / sps;
  constexpr int SMEM = 2 * NT * 8 * QG_ROW_BYTES > NT * 8 * QG_D_STRIDE * 4 ? 2 * NT * 8 * QG_ROW_BYTES : NT * 8 * QG_D_STRIDE * 4;
  dim3 grid((unsigned)(row_blocks * splits)), block(QG_WARPS * 32);
#define GO(XP_)                                                                                                            \
  do {                                                                                                                     \
    auto kern = qmm_mma_kernel<T, NT, XP_>;                                                                                \
    static bool attr_set = false;                                                                                          \
    if (!attr_set) {                                                                                                       \
      cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);                       \
      if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(qmm_mma_kernel)");                                   \
      attr_set = true;                                                                                                     \
    }                                                                                                                      \
    launch_pdl_opt(pdl_enabled_gemm(), kern, grid, block, SMEM, st, (const T*)x, w, (const T*)s, (const T*)b, (float*)y, (int)M, N, K,   \
                   ldx, ldy, (int)splits, (int)sps);                                                                       \
  } while (0)
  if (xperm) GO(true); else GO(false);
#undef GO
  B200P_LAUNCH_CHECK("qmm_mma_kernel");
  return B200P_OK;
}

template <typename T>
static int launch_qmm_mma(const void* x, const uint32_t* w, const void* s, const void* b, void* y, int64_t M, int64_t N,
                          int64_t K, int64_t ldx, int64_t ldy, bool xperm, cudaStream_t st) {
  if (M <= 8) return launch_qmm_mma_nt<T, 1>(x, w, s, b, y, M, N, K, ldx, ldy, xperm, st);
  if (M <= 16) return launch_qmm_mma_nt<T, 2>(x, w, s, b, y, M, N, K, ldx, ldy, xperm, st);
  if (M <= 32) return launch_qmm_mma_nt<T, 4>(x, w, s, b, y, M, N, K, ldx, ldy, xperm, st);
  return launch_qmm_mma_nt<T, 8>(x, w, s, b, y, M, N, K, ldx, ldy, xperm, st);
}

int qmm_mma_launch(const void* x, const uint32_t* w, const void* s, const void* b, void* y, int64_t M, int64_t N, int64_t K,
                   int64_t ldx, int64_t ldy, int dtype, int flags, cudaStream_t st) {
  B200P_REQUIRE(aligned(s, 2) && aligned(b, 2) && aligned(w, 32), B200P_ERR_ALIGN, "bad alignment (weights: 32 bytes)");
  const bool xperm = (flags & B200P_X_PERM8) != 0;
  if (dtype == B200P_F16) return launch_qmm_mma<__half>(x, w, s, b, y, M, N, K, ldx, ldy, xperm, st);
  return launch_qmm_mma<__nv_bfloat16>(x, w, s, b, y, M, N, K, ldx, ldy, xperm, st);
}

}  // namespace b200p
