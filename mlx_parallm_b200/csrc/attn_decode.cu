// Split-K GQA decode attention (L == 1) for 16-bit KV with head_dim 128 -- the hot attention path.
//
// One CTA = (split, kv head, batch row); its 8 warps interleave 16-key tiles of the CTA's key range.
// Each warp streams its K and V tiles global -> shared with cp.async (3 stages, XOR-swizzled 16-byte
// chunks), forms S = Q K^T for all G = Hq/Hkv query heads of the kv head at once with mma.sync
// m16n8k16 (rows = query heads, padded to 16; K/V are therefore read ONCE per kv head, not once per
// q head), keeps an online softmax per row (quad-shuffle row max, fp32), and accumulates O += P V with
// mma.sync again.  Warps are merged through shared memory; if the sequence is split across CTAs the
// (m, l, O) partials go to a float workspace and attn_reduce_kernel merges them.
//
// Why legacy mma.sync and not tcgen05 here: M = G <= 16 rows of real work per kv head; the kernel is
// bound by streaming K/V (2*T*D*2 bytes per kv head) and the tensor pipe is ~10 % busy either way.
// Algorithmic HBM bytes per launch: sum_b len_b * Hkv * D * 2 (K and V) * sizeof(T).
#include <mutex>
#include <type_traits>

#include <stdlib.h>

#include "common.cuh"
#include "mma_util.cuh"

namespace b200p {

int sm_count();

// warps per CTA: 8 (one 192 KB CTA per SM) for long key ranges; 2 (four 48 KB CTAs per SM) when a CTA's key range is
// short -- then the step is a few waves of fixed per-CTA latency and smaller CTAs cut the number of waves
constexpr int AD_TILE = 16;                    // keys per warp tile
constexpr int AD_STAGES = 3;
constexpr int AD_D = 128;
constexpr int AD_TILE_BYTES = AD_TILE * AD_D * 2;                 // 4 KB (K) ; same for V
constexpr int AD_WARP_SMEM = AD_STAGES * 2 * AD_TILE_BYTES;       // 24 KB per warp
constexpr int AD_SHORT_KEYS = 1024;                               // key range per CTA up to which the 2-warp variant is used

// workspace record per (b, q head, split): [m, l, pad, pad, O[128]] floats
constexpr int AD_REC = 4 + AD_D;
constexpr int AD_MAX_SPLITS = 32;                                 // at most one warp's lanes (attn_reduce_kernel)

template <typename T, int AD_WARPS>
__global__ void __launch_bounds__(AD_WARPS * 32, AD_WARPS == 8 ? 1 : 4)
attn_decode_kernel(const T* __restrict__ q, const T* __restrict__ kc, const T* __restrict__ vc,
                   const int32_t* __restrict__ lengths, T* __restrict__ out, float* __restrict__ ws, int Hq, int Hkv,
                   int64_t Tcap, float scale_log2e, int splits, int out_perm) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int split = blockIdx.x, kvh = blockIdx.y, b = blockIdx.z;
  const int G = Hq / Hkv;
  const int quad_row = lane >> 2, qc = lane & 3;
  pdl_wait();
  int klen = lengths[b] + 1;                       // L == 1: the token appended by this step is key klen-1
  if (klen > Tcap) klen = (int)Tcap;
  // key range of this split, in whole 16-key tiles
  const int n_tiles = (klen + AD_TILE - 1) / AD_TILE;
  const int tile_begin = (int)(((int64_t)n_tiles * split) / splits);
  const int tile_end = (int)(((int64_t)n_tiles * (split + 1)) / splits);

  // ---- Q fragments (A operand, rows = query heads of this kv head, padded to 16 with zeros) ----------
  uint32_t qf[AD_D / 16][4];
  {
    const T* qb = q + ((int64_t)b * Hq + (int64_t)kvh * G) * AD_D;
    const uint32_t* q0 = reinterpret_cast<const uint32_t*>(qb + (int64_t)quad_row * AD_D);
    const uint32_t* q1 = reinterpret_cast<const uint32_t*>(qb + (int64_t)(quad_row + 8) * AD_D);
    const bool r0 = quad_row < G, r1 = quad_row + 8 < G;
#pragma unroll
    for (int ks = 0; ks < AD_D / 16; ++ks) {
      qf[ks][0] = r0 ? q0[ks * 8 + qc] : 0u;          // (row, k = 16ks + 2qc, +1)
      qf[ks][1] = r1 ? q1[ks * 8 + qc] : 0u;          // (row + 8, same k)
      qf[ks][2] = r0 ? q0[ks * 8 + 4 + qc] : 0u;      // (row, k + 8)
      qf[ks][3] = r1 ? q1[ks * 8 + 4 + qc] : 0u;
    }
  }

  const uint32_t wbase = (uint32_t)__cvta_generic_to_shared(smem) + warp * AD_WARP_SMEM;
  const T* kbase = kc + ((int64_t)b * Hkv + kvh) * Tcap * AD_D;
  const T* vbase = vc + ((int64_t)b * Hkv + kvh) * Tcap * AD_D;

  // this warp's tiles: tile_begin + warp, + AD_WARPS, ...
  auto issue_tile = [&](int tile, int stage) {
    const uint32_t ks = wbase + stage * 2 * AD_TILE_BYTES;
    const uint32_t vs = ks + AD_TILE_BYTES;
    const int key0 = tile * AD_TILE;
#pragma unroll
    for (int i = 0; i < 8; ++i) {                    // 16 keys x 16 chunks = 256 chunks / 32 lanes
      const int idx = i * 32 + lane;
      const int key = idx >> 4, ch = idx & 15;
      const bool ok = key0 + key < klen;
      const int64_t goff = (int64_t)(key0 + (ok ? key : 0)) * AD_D + ch * 8;
      const uint32_t soff = (uint32_t)(key * 256 + ((ch ^ (key & 7)) << 4));
      cp_async16(ks + soff, kbase + goff, ok);
      cp_async16(vs + soff, vbase + goff, ok);
    }
  };

  float o[AD_D / 8][4];
#pragma unroll
  for (int j = 0; j < AD_D / 8; ++j) o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f;
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;     // rows quad_row and quad_row + 8

  const int first = tile_begin + warp;
  int n_my = first < tile_end ? (tile_end - first + AD_WARPS - 1) / AD_WARPS : 0;
  // prologue: fill STAGES-1 stages
#pragma unroll
  for (int s = 0; s < AD_STAGES - 1; ++s) {
    if (s < n_my) issue_tile(first + s * AD_WARPS, s);
    cp_async_commit();
  }
  for (int it = 0; it < n_my; ++it) {
    const int stage = it % AD_STAGES;
    if (it + AD_STAGES - 1 < n_my) issue_tile(first + (it + AD_STAGES - 1) * AD_WARPS, (it + AD_STAGES - 1) % AD_STAGES);
    cp_async_commit();
    cp_async_wait<AD_STAGES - 1>();
    __syncwarp();
    const uint32_t ks = wbase + stage * 2 * AD_TILE_BYTES;
    const uint32_t vs = ks + AD_TILE_BYTES;
    const int key0 = (first + it * AD_WARPS) * AD_TILE;

    // ---- S = Q K^T : two n-tiles (keys 0-7, 8-15) ------------------------------------------------
    float s0[4] = {0.f, 0.f, 0.f, 0.f}, s1[4] = {0.f, 0.f, 0.f, 0.f};
    {
      const int mat = lane >> 3, r = lane & 7;
      const int key = r + 8 * (mat >> 1);
#pragma unroll
      for (int kstep = 0; kstep < AD_D / 16; ++kstep) {
        const int ch = 2 * kstep + (mat & 1);
        uint32_t bf[4];
        ldsm_x4(ks + (uint32_t)(key * 256 + ((ch ^ (key & 7)) << 4)), bf);
        mma16816<T>(s0, qf[kstep], bf[0], bf[1]);
        mma16816<T>(s1, qf[kstep], bf[2], bf[3]);
      }
    }
    // scale, mask keys beyond klen, row max
    const int kcol = key0 + 2 * qc;                 // this thread's columns: kcol, kcol+1 (tile 0) and +8 (tile 1)
    float v[8] = {s0[0], s0[1], s1[0], s1[1], s0[2], s0[3], s1[2], s1[3]};   // [row0: 4 cols][row1: 4 cols]
    const bool c0 = kcol < klen, c1 = kcol + 1 < klen, c2 = kcol + 8 < klen, c3 = kcol + 9 < klen;
    v[0] = c0 ? v[0] * scale_log2e : -INFINITY; v[1] = c1 ? v[1] * scale_log2e : -INFINITY;
    v[2] = c2 ? v[2] * scale_log2e : -INFINITY; v[3] = c3 ? v[3] * scale_log2e : -INFINITY;
    v[4] = c0 ? v[4] * scale_log2e : -INFINITY; v[5] = c1 ? v[5] * scale_log2e : -INFINITY;
    v[6] = c2 ? v[6] * scale_log2e : -INFINITY; v[7] = c3 ? v[7] * scale_log2e : -INFINITY;
    float t0 = fmaxf(fmaxf(v[0], v[1]), fmaxf(v[2], v[3]));
    float t1 = fmaxf(fmaxf(v[4], v[5]), fmaxf(v[6], v[7]));
    t0 = fmaxf(t0, __shfl_xor_sync(0xffffffffu, t0, 1)); t0 = fmaxf(t0, __shfl_xor_sync(0xffffffffu, t0, 2));
    t1 = fmaxf(t1, __shfl_xor_sync(0xffffffffu, t1, 1)); t1 = fmaxf(t1, __shfl_xor_sync(0xffffffffu, t1, 2));
    const float mn0 = fmaxf(m0, t0), mn1 = fmaxf(m1, t1);      // finite: key0 < klen so column 0 of the tile is valid
    const float corr0 = exp2f(m0 - mn0), corr1 = exp2f(m1 - mn1);
    m0 = mn0; m1 = mn1;
    float p[8];
#pragma unroll
    for (int i = 0; i < 4; ++i) { p[i] = exp2f(v[i] - mn0); p[4 + i] = exp2f(v[4 + i] - mn1); }
    l0 = l0 * corr0 + (p[0] + p[1] + p[2] + p[3]);
    l1 = l1 * corr1 + (p[4] + p[5] + p[6] + p[7]);
#pragma unroll
    for (int j = 0; j < AD_D / 8; ++j) { o[j][0] *= corr0; o[j][1] *= corr0; o[j][2] *= corr1; o[j][3] *= corr1; }
    // P as A fragment (16 rows x 16 keys)
    uint32_t pf[4];
    pf[0] = pack2<T>(p[0], p[1]);     // (row0, keys 2qc, 2qc+1)
    pf[1] = pack2<T>(p[4], p[5]);     // (row1, same keys)
    pf[2] = pack2<T>(p[2], p[3]);     // (row0, keys + 8)
    pf[3] = pack2<T>(p[6], p[7]);     // (row1, keys + 8)
    // ---- O += P V : 16 n-tiles of 8 head dims ------------------------------------------------------
    {
      const int mat = lane >> 3, r = lane & 7;
      const int key = r + 8 * (mat & 1);
#pragma unroll
      for (int jp = 0; jp < AD_D / 16; ++jp) {
        const int ch = 2 * jp + (mat >> 1);
        uint32_t bf[4];
        ldsm_x4_trans(vs + (uint32_t)(key * 256 + ((ch ^ (key & 7)) << 4)), bf);
        mma16816<T>(o[2 * jp], pf, bf[0], bf[1]);
        mma16816<T>(o[2 * jp + 1], pf, bf[2], bf[3]);
      }
    }
    __syncwarp();                                   // all lanes done reading this stage before it is refilled
  }
  cp_async_wait<0>();
  // ---- finish row sums across the quad, publish this warp's partial ---------------------------------
  l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
  pdl_launch_dependents();
  __syncthreads();                                  // everyone is past its K/V smem: reuse it for the merge
  float* wo = reinterpret_cast<float*>(smem) + warp * (16 * (AD_D + 4));   // [16 rows][m, l, pad, pad, O[128]]
  {
    float* r0p = wo + quad_row * (AD_D + 4);
    float* r1p = wo + (quad_row + 8) * (AD_D + 4);
    if (qc == 0) { r0p[0] = m0; r0p[1] = l0; r1p[0] = m1; r1p[1] = l1; }
#pragma unroll
    for (int j = 0; j < AD_D / 8; ++j) {
      r0p[4 + 8 * j + 2 * qc] = o[j][0]; r0p[4 + 8 * j + 2 * qc + 1] = o[j][1];
      r1p[4 + 8 * j + 2 * qc] = o[j][2]; r1p[4 + 8 * j + 2 * qc + 1] = o[j][3];
    }
  }
  __syncthreads();
  // merge the 8 warps: thread t -> (row = t / 16 ... ) : 256 threads cover G rows x 128 dims in passes
  const float* all = reinterpret_cast<const float*>(smem);
  for (int idx = threadIdx.x; idx < G * AD_D; idx += AD_WARPS * 32) {
    const int row = idx / AD_D, d = idx - row * AD_D;
    float M = -INFINITY;
#pragma unroll
    for (int w = 0; w < AD_WARPS; ++w) M = fmaxf(M, all[w * (16 * (AD_D + 4)) + row * (AD_D + 4)]);
    float Lsum = 0.f, acc = 0.f;
#pragma unroll
    for (int w = 0; w < AD_WARPS; ++w) {
      const float* rp = all + w * (16 * (AD_D + 4)) + row * (AD_D + 4);
      const float f = (rp[0] == -INFINITY) ? 0.f : exp2f(rp[0] - M);
      Lsum += rp[1] * f;
      acc += rp[4 + d] * f;
    }
    const int hq = kvh * G + row;
    if (splits == 1) {
      out[((int64_t)b * Hq + hq) * AD_D + (out_perm ? perm8_pos(d) : d)] = Tr<T>::from_f(acc / Lsum);
    } else {
      float* rec = ws + (((int64_t)b * Hq + hq) * splits + split) * AD_REC;
      rec[4 + d] = acc;
      if (d == 0) { rec[0] = M; rec[1] = Lsum; }
    }
  }
}

// merge the per-split partials: grid = B * Hq, block = 128 (one thread per head dim).
// Two round trips to L2 instead of a serial walk over the splits: lane s of every warp loads (m, l) of split s (splits <= 32 =
// one warp), the softmax factors come out of two warp reductions and stay in the lanes; then each thread issues ALL its
// O[s][d] loads before it uses any (the loop over AD_MAX_SPLITS is unrolled, loads of absent splits are predicated off) and
// picks factor s up with a shuffle.  (The serial form took 5.5 us for 32 splits at B = 1, T = 8192.)
template <typename T>
__global__ void __launch_bounds__(AD_D)
attn_reduce_kernel(const float* __restrict__ ws, T* __restrict__ out, int splits, int out_perm) {
  pdl_wait();
  const int64_t bh = blockIdx.x;
  const int d = threadIdx.x, lane = threadIdx.x & 31;
  const float* base = ws + bh * splits * AD_REC;
  const bool have = lane < splits;
  const float2 ml = have ? *reinterpret_cast<const float2*>(base + lane * AD_REC) : make_float2(-INFINITY, 0.f);
  float o[AD_MAX_SPLITS];
#pragma unroll
  for (int s = 0; s < AD_MAX_SPLITS; ++s) o[s] = s < splits ? base[s * AD_REC + 4 + d] : 0.f;
  float M = ml.x;
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) M = fmaxf(M, __shfl_xor_sync(0xffffffffu, M, off));
  const float f = (ml.x == -INFINITY) ? 0.f : exp2f(ml.x - M);     // factor of split `lane`
  float L = ml.y * f;
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) L += __shfl_xor_sync(0xffffffffu, L, off);
  float acc = 0.f;
#pragma unroll
  for (int s = 0; s < AD_MAX_SPLITS; ++s) acc += o[s] * __shfl_sync(0xffffffffu, f, s);
  pdl_launch_dependents();
  out[bh * AD_D + (out_perm ? perm8_pos(d) : d)] = Tr<T>::from_f(acc / L);
}

// ------------------------------------------------------------------------------------------------ host
static int pick_splits(int64_t B, int Hkv, int64_t Tcap, int ctas_per_sm) {
  // enough CTAs to fill the machine for >= ~2 waves, but >= 256 keys per CTA
  const int64_t base = B * Hkv;
  const int64_t slots = (int64_t)sm_count() * ctas_per_sm;
  int best = 1;
  double best_eff = 0.0;
  for (int s = 1; s <= AD_MAX_SPLITS; ++s) {
    if (s > 1 && Tcap / s < 256) break;
    const int64_t ctas = base * s;
    const int64_t waves = (ctas + slots - 1) / slots;
    const double eff = (double)ctas / (double)(waves * slots);
    if (eff > best_eff + 0.02) { best_eff = eff; best = s; }
  }
  return best;
}

bool attn_decode_supported(int64_t L, int Hq, int Hkv, int D, int dtype) {
  return L == 1 && D == AD_D && (dtype == B200P_F16 || dtype == B200P_BF16) && Hq % Hkv == 0 && Hq / Hkv <= 16;
}

size_t attn_decode_workspace_bytes(int64_t B, int64_t L, int Hq, int D) {
  if (L != 1 || D != AD_D) return 0;
  return (size_t)B * Hq * AD_MAX_SPLITS * AD_REC * sizeof(float);
}

template <typename T, int WARPS>
static int launch_attn_decode_w(const void* q, const void* kc, const void* vc, const int32_t* lengths, void* out, int64_t B,
                                int Hq, int Hkv, int64_t Tcap, float scale, int out_perm, void* ws, int splits, cudaStream_t st) {
  auto kern = attn_decode_kernel<T, WARPS>;
  constexpr int SMEM = WARPS * AD_WARP_SMEM;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(attn_decode_kernel)");
    attr_set = true;
  }
  dim3 grid((unsigned)splits, (unsigned)Hkv, (unsigned)B);
  launch_pdl(kern, grid, WARPS * 32, SMEM, st, (const T*)q, (const T*)kc, (const T*)vc, lengths, (T*)out, (float*)ws, Hq, Hkv,
             Tcap, scale * 1.4426950408889634f, splits, out_perm);
  B200P_LAUNCH_CHECK("attn_decode_kernel");
  if (splits > 1) {
    launch_pdl(attn_reduce_kernel<T>, (unsigned)(B * Hq), AD_D, 0, st, (const float*)ws, (T*)out, splits, out_perm);
    B200P_LAUNCH_CHECK("attn_reduce_kernel");
  }
  return B200P_OK;
}

static int g_ad_warps = [] { const char* e = getenv("B200P_ATTN_WARPS"); return e ? atoi(e) : 0; }();   // 0 = auto

template <typename T>
static int launch_attn_decode(const void* q, const void* kc, const void* vc, const int32_t* lengths, void* out, int64_t B,
                              int Hq, int Hkv, int64_t Tcap, float scale, int out_perm, void* ws, cudaStream_t st) {
  // the variant and the split count depend on capacity, not on the (device-side) lengths, so a captured graph stays valid
  const int s8 = pick_splits(B, Hkv, Tcap, 1);
  const bool small = g_ad_warps ? g_ad_warps == 2 : (Tcap / s8 <= AD_SHORT_KEYS);
  if (small) {
    const int s2 = pick_splits(B, Hkv, Tcap, 4);
    return launch_attn_decode_w<T, 2>(q, kc, vc, lengths, out, B, Hq, Hkv, Tcap, scale, out_perm, ws, s2, st);
  }
  return launch_attn_decode_w<T, 8>(q, kc, vc, lengths, out, B, Hq, Hkv, Tcap, scale, out_perm, ws, s8, st);
}

int attn_decode_launch(const void* q, const void* kc, const void* vc, const int32_t* lengths, void* out, int64_t B,
                       int64_t L, int Hq, int Hkv, int D, int64_t Tcap, float scale, int dtype, int out_perm, void* ws,
                       size_t ws_bytes, cudaStream_t st) {
  B200P_REQUIRE(attn_decode_supported(L, Hq, Hkv, D, dtype), B200P_ERR_ARG, "tensor decode attention: unsupported shape");
  B200P_REQUIRE(B < 65536 && Hkv < 65536, B200P_ERR_ARG, "grid too large");
  B200P_REQUIRE(ws && ws_bytes >= attn_decode_workspace_bytes(B, L, Hq, D) && aligned(ws, 16), B200P_ERR_ARG,
                "attention workspace too small");
  B200P_REQUIRE(aligned(q, 16) && aligned(kc, 16) && aligned(vc, 16), B200P_ERR_ALIGN, "q / caches must be 16-byte aligned");
  if (dtype == B200P_F16) return launch_attn_decode<__half>(q, kc, vc, lengths, out, B, Hq, Hkv, Tcap, scale, out_perm, ws, st);
  return launch_attn_decode<__nv_bfloat16>(q, kc, vc, lengths, out, B, Hq, Hkv, Tcap, scale, out_perm, ws, st);
}

}  // namespace b200p

// ------------------------------------------------------------------------------------------------ no-allocation variant
// b200p_attn_decode_static: the split-K decode attention above with its split partials in a FILE-SCOPE device array
// instead of a caller-allocated workspace -- for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
// serving loops, harnesses that forbid it).  No counterpart in the reference.
//
// The array is part of the library image: every GPU that loads the library holds its 16 MiB until the process exits.
// In HOST code the name `g_attn_static_scratch` is only a stand-in for the symbol, not a device address (a kernel handed
// it would fault or, where the GPU can reach host memory, silently use the wrong memory): the device address is resolved
// once per device with cudaGetSymbolAddress, with that device current, and that pointer is what the kernels receive.
// Every call shares the array and it keeps the last call's contents.  Nothing has to start at zero: attn_decode_kernel writes
// the (m, l, O) record of EVERY (row, query head, split) of the call -- an empty split writes m = -inf, l = 0 -- before
// attn_reduce_kernel of the same call reads them, and no record outside the call's B*Hq*splits is read.  Two calls that
// overlap on the GPU (different streams, or concurrent graph branches) race on the array: serialise them.
constexpr size_t ATTN_STATIC_SCRATCH_BYTES = B200P_ATTN_STATIC_SCRATCH_BYTES;
static __device__ __align__(256) float g_attn_static_scratch[ATTN_STATIC_SCRATCH_BYTES / sizeof(float)];

namespace b200p {
constexpr int ATTN_STATIC_MAX_DEVICES = 64;
static void* g_attn_static_ptr[ATTN_STATIC_MAX_DEVICES] = {};
static std::mutex g_attn_static_mu;

// device address of the scratch array on the CURRENT device (resolved on the first call per device; no allocation)
static int attn_static_scratch_ptr(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return cuda_fail(e, "cudaGetDevice");
  B200P_REQUIRE(dev >= 0 && dev < ATTN_STATIC_MAX_DEVICES, B200P_ERR_ARG, "device index %d beyond %d", dev,
                ATTN_STATIC_MAX_DEVICES);
  std::lock_guard<std::mutex> lock(g_attn_static_mu);
  if (!g_attn_static_ptr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_attn_static_scratch);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(g_attn_static_scratch)");
    B200P_REQUIRE(p != nullptr && aligned(p, 256), B200P_ERR_CUDA, "static attention scratch resolved to a bad address");
    g_attn_static_ptr[dev] = p;
  }
  *out = g_attn_static_ptr[dev];
  return B200P_OK;
}
}  // namespace b200p

extern "C" int b200p_attn_decode_static(const void* q, const void* kcache, const void* vcache, const int32_t* lengths,
                                        void* out, int64_t B, int Hq, int Hkv, int D, int64_t Tcap, float scale, int dtype,
                                        int flags, void* stream) {
  using namespace b200p;
  B200P_REQUIRE(q && kcache && vcache && lengths && out, B200P_ERR_ARG, "null pointer");
  B200P_REQUIRE((flags & ~B200P_OUT_PERM8) == 0, B200P_ERR_ARG, "unknown flags 0x%x", flags);
  B200P_REQUIRE(Hq > 0 && Hkv > 0 && Hq % Hkv == 0, B200P_ERR_ARG, "bad head counts %d/%d", Hq, Hkv);
  B200P_REQUIRE(B >= 0 && Tcap > 0, B200P_ERR_ARG, "bad shape");
  B200P_REQUIRE(attn_decode_supported(1, Hq, Hkv, D, dtype), B200P_ERR_ARG,
                "b200p_attn_decode_static: 16-bit caches, head_dim 128 and at most 16 query heads per kv head only");
  const size_t need = attn_decode_workspace_bytes(B, 1, Hq, D);
  B200P_REQUIRE(need <= ATTN_STATIC_SCRATCH_BYTES, B200P_ERR_ARG,
                "b200p_attn_decode_static: B * Hq = %lld needs %zu bytes of scratch, the static array holds %zu "
                "(B * Hq <= %zu); use b200p_attn with a caller-allocated workspace",
                (long long)(B * Hq), need, ATTN_STATIC_SCRATCH_BYTES,
                ATTN_STATIC_SCRATCH_BYTES / (size_t)(AD_MAX_SPLITS * AD_REC * sizeof(float)));
  if (B == 0) return B200P_OK;
  void* ws = nullptr;
  const int rc = attn_static_scratch_ptr(&ws);
  if (rc != B200P_OK) return rc;
  PdlClass pdl_cls(8);
  return attn_decode_launch(q, kcache, vcache, lengths, out, B, 1, Hq, Hkv, D, Tcap, scale, dtype,
                            (flags & B200P_OUT_PERM8) ? 1 : 0, ws, ATTN_STATIC_SCRATCH_BYTES, (cudaStream_t)stream);
}
