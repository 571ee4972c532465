// Logits -> token: greedy argmax and temperature / nucleus (top-p) sampling.
// One CTA per batch row; the row (V floats, <= 600 KB) is re-read from L2 on each pass.
//
// A row is split over the CTAs of a thread-block cluster (1, 2, 4 or 8, chosen so that batch x cluster fills the SMs once):
// one CTA streams its row out of L2 at ~25 GB/s (measured: 513 KB in 20 us, whatever the loads in flight), so small batches
// need several SMs per row.  Cross-CTA reductions are ALL-TO-ALL through distributed shared memory: every CTA stores its
// partial result into slot [rank] of every CTA of the cluster, one barrier.cluster, then every CTA combines the slots in the
// same order -- so all CTAs take the same decisions without a broadcast, and nothing goes through global memory.
#include <cooperative_groups.h>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace b200p {

constexpr int SAMP_THREADS = 1024;
constexpr int SAMP_MAX_CLUSTER = 8;

// [lo, hi) of the row for this CTA of the cluster; lo is a multiple of 4 so that 16-byte loads stay aligned
__device__ __forceinline__ void row_segment(int64_t V, unsigned rank, unsigned S, int64_t& lo, int64_t& hi) {
  const int64_t per = (((V + S - 1) / S) + 3) & ~(int64_t)3;
  lo = min((int64_t)rank * per, V);
  hi = min(lo + per, V);
}

// ---- argmax (lowest index among ties, like mx.argmax) ----------------------------------------
__device__ __forceinline__ void argmax_combine(float& v, int& i, float ov, int oi) {
  if (ov > v || (ov == v && oi < i)) { v = ov; i = oi; }
}

__device__ __forceinline__ void block_argmax(float& v, int& i, float* shv, int* shi) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float ov = __shfl_xor_sync(0xffffffffu, v, o);
    int oi = __shfl_xor_sync(0xffffffffu, i, o);
    argmax_combine(v, i, ov, oi);
  }
  int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) { shv[wid] = v; shi[wid] = i; }
  __syncthreads();
  if (wid == 0) {
    int nw = blockDim.x >> 5;
    v = lane < nw ? shv[lane] : -INFINITY;
    i = lane < nw ? shi[lane] : 0x7fffffff;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      float ov = __shfl_xor_sync(0xffffffffu, v, o);
      int oi = __shfl_xor_sync(0xffffffffu, i, o);
      argmax_combine(v, i, ov, oi);
    }
    if (lane == 0) { shv[0] = v; shi[0] = i; }
  }
  __syncthreads();
  v = shv[0];
  i = shi[0];
  __syncthreads();
}

// online (max, sum exp) pair so that prob(argmax) = 1 / sum_i exp(l_i - max) comes out of the same pass
__device__ __forceinline__ void osm_push(float& m, float& s, float v) {
  if (v > m) { s = s * __expf(m - v) + 1.0f; m = v; }
  else if (v > -INFINITY) s += __expf(v - m);
}
__device__ __forceinline__ void osm_merge(float& m, float& s, float om, float os) {
  float nm = fmaxf(m, om);
  if (nm == -INFINITY) return;
  s = s * __expf(m - nm) + os * __expf(om - nm);
  m = nm;
}

// every element of row[lo, hi) once, f(value, index): 16-byte loads when the row allows (lo % 4 == 0), four in flight per thread
template <typename F>
__device__ __forceinline__ void row_visit(const float* __restrict__ row, int64_t lo, int64_t hi, bool vec, F&& f) {
  const int64_t nt = blockDim.x;
  const int64_t n = hi - lo;
  const int64_t V4 = vec ? n / 4 : 0;
  const float4* r4 = reinterpret_cast<const float4*>(row + lo);
  int64_t i = threadIdx.x;
  for (; i + 3 * nt < V4; i += 4 * nt) {
    float4 v[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = r4[i + j * nt];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t e = lo + 4 * (i + j * nt);
      f(v[j].x, e); f(v[j].y, e + 1); f(v[j].z, e + 2); f(v[j].w, e + 3);
    }
  }
  for (; i < V4; i += nt) {
    const float4 a = r4[i];
    const int64_t e = lo + 4 * i;
    f(a.x, e); f(a.y, e + 1); f(a.z, e + 2); f(a.w, e + 3);
  }
  for (int64_t j = lo + V4 * 4 + threadIdx.x; j < hi; j += nt) f(row[j], j);
}

// grid (S, B), cluster (S, 1, 1): CTA `rank` of the cluster takes segment `rank` of row blockIdx.y
__global__ void __launch_bounds__(SAMP_THREADS)
argmax_kernel(const float* __restrict__ logits, int32_t* __restrict__ tokens, float* __restrict__ probs_out,
              int64_t V, int64_t ldl, float4* __restrict__ rec_out, int vocab_offset) {
  __shared__ float shv[32];
  __shared__ int shi[32];
  __shared__ float shm[32], shs[32];
  __shared__ float4 s_rec[SAMP_MAX_CLUSTER];       // (max, online-softmax max, sum, index) of every CTA of the cluster
  cg::cluster_group cluster = cg::this_cluster();
  const unsigned rank = cluster.block_rank(), S = cluster.num_blocks();
  pdl_wait();
  const float* row = logits + (int64_t)blockIdx.y * ldl;
  float best = -INFINITY;
  int bi = 0x7fffffff;
  float om = -INFINITY, os = 0.f;
  const bool want_p = probs_out != nullptr || rec_out != nullptr;
  const bool vec = (ldl % 4 == 0) && ((reinterpret_cast<uintptr_t>(logits) & 15) == 0);
  int64_t lo, hi;
  row_segment(V, rank, S, lo, hi);
  // NaN never wins (comparisons false), matching a total order that ignores NaN; a thread sees its indices in ascending
  // order, so the strict compare keeps the lowest index among its ties
  row_visit(row, lo, hi, vec, [&](float v, int64_t i) {
    if (v > best) { best = v; bi = (int)i; }
    if (want_p) osm_push(om, os, v);
  });
  if (want_p) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      float m2 = __shfl_xor_sync(0xffffffffu, om, o), s2 = __shfl_xor_sync(0xffffffffu, os, o);
      osm_merge(om, os, m2, s2);
    }
    if ((threadIdx.x & 31) == 0) { shm[threadIdx.x >> 5] = om; shs[threadIdx.x >> 5] = os; }
  }
  block_argmax(best, bi, shv, shi);  // contains __syncthreads, publishes shm/shs too
  if (threadIdx.x == 0) {
    float m = -INFINITY, sm = 0.f;
    if (want_p)
      for (int w = 0; w < (int)(blockDim.x >> 5); ++w) osm_merge(m, sm, shm[w], shs[w]);
    *cluster.map_shared_rank(&s_rec[rank], 0) = make_float4(best, m, sm, __int_as_float(bi));
  }
  cluster.sync();
  if (rank == 0 && threadIdx.x == 0) {
    float gb = -INFINITY, m = -INFINITY, sm = 0.f;
    int gi = 0x7fffffff;
    for (unsigned r = 0; r < S; ++r) {             // ascending segments: ties keep the lowest index
      const float4 rec = s_rec[r];
      argmax_combine(gb, gi, rec.x, __float_as_int(rec.w));
      osm_merge(m, sm, rec.y, rec.z);
    }
    const int tok = (gi == 0x7fffffff) ? 0 : gi;
    if (rec_out != nullptr) {   // vocab-parallel shard: publish (max, sum exp(l - max), global index)
      rec_out[blockIdx.y] = make_float4(gb, sm, __int_as_float(tok + vocab_offset), 0.f);
    } else {
      tokens[blockIdx.y] = tok;
      if (probs_out != nullptr) probs_out[blockIdx.y] = (sm > 0.f) ? __expf(gb - m) / sm : 0.f;
    }
  }
}

// merge W per-shard records per row: highest max wins, lowest global index among ties (shards hold
// ascending vocab ranges, so this is mx.argmax's lowest-index rule); prob = 1 / sum_r s_r exp(m_r - M).
__global__ void argmax_merge_kernel(const float4* __restrict__ recs, int32_t* __restrict__ tokens,
                                    float* __restrict__ probs_out, int W, int64_t B) {
  pdl_wait();
  const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  float best = -INFINITY, m = -INFINITY, s = 0.f;
  int bi = 0x7fffffff;
  for (int r = 0; r < W; ++r) {
    const float4 rec = recs[(int64_t)r * B + b];
    argmax_combine(best, bi, rec.x, __float_as_int(rec.z));
    osm_merge(m, s, rec.x, rec.y);
  }
  tokens[b] = (bi == 0x7fffffff) ? 0 : bi;
  if (probs_out != nullptr) probs_out[b] = (s > 0.f) ? __expf(best - m) / s : 0.f;
}

// ---- Philox4x32-10 -------------------------------------------------------------------------------
__device__ __forceinline__ void philox_round(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
  uint32_t hi0 = __umulhi(M0, c[0]), lo0 = M0 * c[0];
  uint32_t hi1 = __umulhi(M1, c[2]), lo1 = M1 * c[2];
  uint32_t n0 = hi1 ^ c[1] ^ k0, n1 = lo1, n2 = hi0 ^ c[3] ^ k1, n3 = lo0;
  c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
}
__device__ __forceinline__ float philox_uniform(uint64_t seed, uint64_t offset, uint32_t row) {
  uint32_t c[4] = {(uint32_t)offset, (uint32_t)(offset >> 32), row, 0u};
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    philox_round(c, k0, k1);
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  return ((c[0] >> 8) + 0.5f) * (1.0f / 16777216.0f);  // (0,1)
}

__device__ __forceinline__ float block_reduce(float v, float* sh, bool is_max) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float ov = __shfl_xor_sync(0xffffffffu, v, o);
    v = is_max ? fmaxf(v, ov) : v + ov;
  }
  int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) sh[wid] = v;
  __syncthreads();
  if (wid == 0) {
    int nw = blockDim.x >> 5;
    v = lane < nw ? sh[lane] : (is_max ? -INFINITY : 0.f);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      float ov = __shfl_xor_sync(0xffffffffu, v, o);
      v = is_max ? fmaxf(v, ov) : v + ov;
    }
    if (lane == 0) sh[0] = v;
  }
  __syncthreads();
  v = sh[0];
  __syncthreads();
  return v;
}

// Nucleus sampling, one CTA per row; the row (V floats, <= 600 KB) stays in L2 and is re-read on each pass.
//  1. mx = max(logit).
//  2. e_i = exp((logit_i - mx) / temp) in (0, 1], Z = sum e_i.  The nucleus is found by radix-select over the float bits of
//     e_i (monotone for e >= 0; bits 31..30 are 0 because e <= 1), most significant digit first, 6 bits per pass: the MASS of
//     the elements carrying each digit value (under the prefix chosen so far) is accumulated, then the digits are walked
//     from the top until the inclusive cumulative mass would exceed top_p * Z.  After 5 passes the prefix is the value v* of
//     the first sorted element that does not fit; the nucleus is {e_i > v*}, its mass `kept`.  (Selection is by VALUE: when
//     the boundary falls inside a run of exactly equal probabilities the whole run is dropped, where the reference keeps as
//     many of the run as fit in the unspecified order its argsort leaves ties.)
//     The histograms are LANE-PRIVATE ([64 digits][512 threads] floats in shared memory, bank = thread, summed per digit at
//     the end of a pass): no atomics.  Shared-memory float atomics are CAS loops and thousands of tokens share the leading
//     digits -- the first version of this kernel (one shared 256-bin histogram) took 600 us per step at V = 128 256.
//  3. u ~ U(0, kept): pick the token where the running mass over the nucleus crosses u.  The enumeration order is
//     thread-major (thread t owns the 16-byte groups t, t + 512, ...), as valid an order for a categorical draw as any.
constexpr int TP_THREADS = 512;
constexpr int TP_BINS = 64;
constexpr int TP_PASSES = 5;                       // 5 x 6 bits = bits 29..0
constexpr int TP_SMEM = TP_BINS * TP_THREADS * (int)sizeof(float);

// all-to-all slot layout of one exchange buffer: [rank][TP_XCH] floats
constexpr int TP_XCH = TP_BINS + 2;               // 64 digit masses + Z + Z1

// grid (S, B), cluster (S, 1, 1): CTA `rank` takes segment `rank` of row blockIdx.y in every pass
__global__ void __launch_bounds__(TP_THREADS)
top_p_kernel(const float* __restrict__ logits, int32_t* __restrict__ tokens, float* __restrict__ probs_out,
             int64_t V, int64_t ldl, float inv_temp, float top_p, uint64_t seed, uint64_t offset) {
  extern __shared__ float lhist[];                 // [TP_BINS][TP_THREADS]
  __shared__ float sh[32];
  __shared__ float hist[TP_BINS];
  __shared__ float s_xch[2][SAMP_MAX_CLUSTER][TP_XCH];   // written by every CTA of the cluster, double-buffered by pass parity
  __shared__ float s_maxs[SAMP_MAX_CLUSTER];
  __shared__ float s_tot[SAMP_MAX_CLUSTER];
  __shared__ int s_idx[SAMP_MAX_CLUSTER];
  __shared__ uint32_t s_prefix, s_thr;
  __shared__ float s_above;
  __shared__ int s_done;
  __shared__ float s_scan[TP_THREADS];
  __shared__ int s_tok, s_last;
  cg::cluster_group cluster = cg::this_cluster();
  const unsigned rank = cluster.block_rank(), S = cluster.num_blocks();
  pdl_wait();
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t b_row = blockIdx.y;
  const float* row = logits + b_row * ldl;
  const bool vec = (ldl % 4 == 0) && ((reinterpret_cast<uintptr_t>(logits) & 15) == 0);
  int64_t lo, hi;
  row_segment(V, rank, S, lo, hi);

  float mx = -INFINITY;
  row_visit(row, lo, hi, vec, [&](float x, int64_t) { mx = fmaxf(mx, x); });
  mx = block_reduce(mx, sh, true);
  if (tid < (int)S) *cluster.map_shared_rank(&s_maxs[rank], tid) = mx;
  cluster.sync();
  for (unsigned r = 0; r < S; ++r) mx = fmaxf(mx, s_maxs[r]);

  float z = 0.f, z1 = 0.f;
  uint32_t thr_bits = 0u;  // nucleus = {bits(e) > thr_bits}; 0 keeps every e > 0
  bool empty = false;
  if (!(top_p < 1.0f)) {   // plain categorical(logits / temp)
    row_visit(row, lo, hi, vec, [&](float x, int64_t) {
      const float d = x - mx;
      z += __expf(d * inv_temp);
      z1 += __expf(d);
    });
    z1 = block_reduce(z1, sh, false);
    if (tid < (int)S) *cluster.map_shared_rank(&s_xch[0][rank][TP_BINS + 1], tid) = z1;
    cluster.sync();
    z1 = 0.f;
    for (unsigned r = 0; r < S; ++r) z1 += s_xch[0][r][TP_BINS + 1];
  } else {
    if (tid == 0) { s_prefix = 0u; s_thr = 0u; s_above = 0.f; s_done = 0; }
    float target = 0.f;
    for (int pass = 0; pass < TP_PASSES; ++pass) {
      const int shift = 24 - 6 * pass;
      const int buf = pass & 1;
#pragma unroll 8
      for (int b = 0; b < TP_BINS; ++b) lhist[b * TP_THREADS + tid] = 0.f;
      __syncthreads();
      const uint32_t prefix = s_prefix;            // bits(e) >> (shift + 6) of the elements still in play
      const bool done = s_done != 0;
      if (!done) {
        if (pass == 0) {
          row_visit(row, lo, hi, vec, [&](float x, int64_t) {
            const float d = x - mx;
            const float e = __expf(d * inv_temp);
            z += e;
            z1 += __expf(d);
            const uint32_t bits = __float_as_uint(e);
            if ((bits >> 30) == 0u) lhist[(bits >> 24) * TP_THREADS + tid] += e;     // NaN stays out
          });
        } else {
          row_visit(row, lo, hi, vec, [&](float x, int64_t) {
            const float e = __expf((x - mx) * inv_temp);
            const uint32_t bits = __float_as_uint(e);
            if ((bits >> (shift + 6)) == prefix) lhist[((bits >> shift) & 63u) * TP_THREADS + tid] += e;
          });
        }
      }
      if (pass == 0) {
        z = block_reduce(z, sh, false);
        z1 = block_reduce(z1, sh, false);
      }
      __syncthreads();
      // digit totals of this CTA: warp w sums digits 4w .. 4w+3 over the 512 lane-private columns
      for (int b = warp * (TP_BINS / (TP_THREADS / 32)); b < (warp + 1) * (TP_BINS / (TP_THREADS / 32)); ++b) {
        float v = 0.f;
#pragma unroll
        for (int j = 0; j < TP_THREADS / 32; ++j) v += lhist[b * TP_THREADS + lane + 32 * j];
        v = warp_sum(v);
        if (lane == 0) hist[b] = v;
      }
      __syncthreads();
      // all-to-all: this CTA's totals into slot [rank] of every CTA
      for (int k = tid; k < TP_XCH * (int)S; k += TP_THREADS) {
        const int r = k / TP_XCH, j = k - r * TP_XCH;
        const float v = j < TP_BINS ? hist[j] : (j == TP_BINS ? z : z1);
        *cluster.map_shared_rank(&s_xch[buf][rank][j], r) = v;
      }
      cluster.sync();
      if (tid < TP_BINS) {
        float v = 0.f;
        for (unsigned r = 0; r < S; ++r) v += s_xch[buf][r][tid];
        hist[tid] = v;
      }
      if (pass == 0) {
        float zz = 0.f, zz1 = 0.f;
        for (unsigned r = 0; r < S; ++r) { zz += s_xch[buf][r][TP_BINS]; zz1 += s_xch[buf][r][TP_BINS + 1]; }
        z = zz;
        z1 = zz1;
        target = top_p * z;
      }
      __syncthreads();
      if (tid == 0 && !done) {                     // every CTA of the cluster takes the same decision from the same sums
        float above = s_above;
        int d = TP_BINS - 1;
        for (; d >= 0; --d) {
          if (above + hist[d] > target) break;
          above += hist[d];
        }
        s_above = above;
        if (d < 0) {
          // nothing under this prefix crosses top_p (pass 0: the whole row fits; later passes: only through summation-order
          // rounding) -> every element carrying the prefix is kept
          const uint32_t lowest = prefix << (shift + 6);
          s_done = 1;
          s_thr = lowest == 0u ? 0u : lowest - 1u;
        } else {
          s_prefix = (prefix << 6) | (uint32_t)d;
          if (pass == TP_PASSES - 1) s_thr = (prefix << 6) | (uint32_t)d;
        }
      }
      __syncthreads();
    }
    thr_bits = s_thr;
    if (!s_done && s_above <= 0.f) empty = true;   // even the largest p exceeds top_p (reference: 0/0)
  }

  if (empty) {                                     // documented deviation: the argmax (lowest index among ties)
    if (tid == 0) s_tok = 0x7fffffff;
    __syncthreads();
    int bi = 0x7fffffff;
    row_visit(row, lo, hi, vec, [&](float x, int64_t i) { if (x == mx && (int)i < bi) bi = (int)i; });
    if (bi != 0x7fffffff) atomicMin(&s_tok, bi);
    __syncthreads();
    if (tid < (int)S) *cluster.map_shared_rank(&s_idx[rank], tid) = s_tok;
    cluster.sync();
    if (rank == 0 && tid == 0) {
      int tok = 0x7fffffff;
      for (unsigned r = 0; r < S; ++r) tok = min(tok, s_idx[r]);
      tok = tok == 0x7fffffff ? 0 : tok;
      tokens[b_row] = tok;
      if (probs_out) probs_out[b_row] = __expf(row[tok] - mx) / z1;
    }
    return;
  }

  float local = 0.f;
  int mylast = -1;
  row_visit(row, lo, hi, vec, [&](float x, int64_t i) {
    const float e = __expf((x - mx) * inv_temp);
    if (__float_as_uint(e) > thr_bits) { local += e; mylast = max(mylast, (int)i); }
  });
  s_scan[tid] = local;
  if (tid == 0) { s_tok = -1; s_last = -1; }
  __syncthreads();
  // inclusive Hillis-Steele scan over the threads
  for (int o = 1; o < TP_THREADS; o <<= 1) {
    const float add = tid >= o ? s_scan[tid - o] : 0.f;
    __syncthreads();
    s_scan[tid] += add;
    __syncthreads();
  }
  if (mylast >= 0) atomicMax(&s_last, mylast);
  __syncthreads();
  // nucleus mass of every CTA (and whether it holds any nucleus token) to every CTA
  if (tid < (int)S) {
    *cluster.map_shared_rank(&s_tot[rank], tid) = s_scan[TP_THREADS - 1];
    *cluster.map_shared_rank(&s_idx[rank], tid) = s_last;
  }
  cluster.sync();
  float total = 0.f;
  for (unsigned r = 0; r < S; ++r) total += s_tot[r];
  const float u = philox_uniform(seed, offset, (uint32_t)b_row) * total;
  // owner CTA: the first whose inclusive cumulative mass exceeds u; on a rounding seam the last CTA holding a nucleus token
  int owner = -1, last_nonempty = -1;
  float base = 0.f, owner_base = 0.f;
  for (unsigned r = 0; r < S; ++r) {
    if (s_idx[r] >= 0) last_nonempty = (int)r;
    if (owner < 0 && s_tot[r] > 0.f && u < base + s_tot[r]) { owner = (int)r; owner_base = base; }
    base += s_tot[r];
  }
  if (owner < 0) { owner = last_nonempty < 0 ? 0 : last_nonempty; owner_base = INFINITY; }   // INFINITY: nobody claims, take the last
  if ((int)rank != owner) return;
  // position inside this CTA's mass (-inf on the seam path: no thread claims it and the last nucleus token is taken)
  const float ul = owner_base == INFINITY ? -INFINITY : fmaxf(u - owner_base, 0.f);
  const float before = tid ? s_scan[tid - 1] : 0.f;
  if (local > 0.f && ul >= before && ul < s_scan[tid]) {   // the owning thread walks its elements in the same order
    float run = before;
    int pick = -1, last = -1;
    row_visit(row, lo, hi, vec, [&](float x, int64_t i) {
      const float e = __expf((x - mx) * inv_temp);
      if (__float_as_uint(e) > thr_bits) {
        run += e;
        last = (int)i;
        if (pick < 0 && ul < run) pick = (int)i;
      }
    });
    s_tok = pick >= 0 ? pick : last;
  }
  __syncthreads();
  if (tid == 0) {
    int tok = s_tok;
    if (tok < 0) tok = s_last < 0 ? 0 : s_last;    // rounding seam: the last nucleus token of this CTA
    tokens[b_row] = tok;
    if (probs_out) probs_out[b_row] = __expf(row[tok] - mx) / z1;
  }
}

}  // namespace b200p

using namespace b200p;

// cluster size: the largest of 8, 4, 2, 1 with B * S <= SMs * ctas_per_sm (one wave); B200P_SAMPLER_CLUSTER overrides
static int cluster_for(int64_t B, int ctas_per_sm) {
  static int forced = -1;
  if (forced < 0) {
    const char* e = getenv("B200P_SAMPLER_CLUSTER");
    forced = e ? atoi(e) : 0;
    if (forced != 1 && forced != 2 && forced != 4 && forced != 8) forced = 0;
  }
  if (forced) return forced;
  int dev = 0, sms = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  int s = SAMP_MAX_CLUSTER;
  while (s > 1 && B * s > (int64_t)sms * ctas_per_sm) s >>= 1;
  return s;
}

// grid (S, B) with thread-block clusters of S CTAs along x (one cluster per row) + the PDL attribute
template <typename... KArgs, typename... Args>
static cudaError_t launch_cluster(void (*kern)(KArgs...), int S, int64_t B, int threads, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)S, (unsigned)B);
  cfg.blockDim = dim3((unsigned)threads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)S;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

extern "C" int b200p_argmax(const float* logits, int32_t* tokens, float* probs_out, int64_t B, int64_t V,
                            int64_t ldl, void* stream) {
  PdlClass pdl_cls(32);
  B200P_REQUIRE(logits && tokens, B200P_ERR_ARG, "null pointer");
  B200P_REQUIRE(V > 0 && V < (1ll << 31) && ldl >= V, B200P_ERR_ARG, "bad V/ldl");
  B200P_REQUIRE(B <= 65535, B200P_ERR_ARG, "at most 65535 rows per call (grid.y)");
  if (B == 0) return B200P_OK;
  cudaError_t e = launch_cluster(argmax_kernel, cluster_for(B, 2), B, SAMP_THREADS, 0, (cudaStream_t)stream, logits, tokens,
                                 probs_out, V, ldl, (float4*)nullptr, 0);
  if (e != cudaSuccess) return cuda_fail(e, "argmax_kernel");
  return B200P_OK;
}

extern "C" int b200p_argmax_partial(const float* logits, float* rec_out, int64_t B, int64_t V, int64_t ldl,
                                    int32_t vocab_offset, void* stream) {
  PdlClass pdl_cls(32);
  B200P_REQUIRE(logits && rec_out, B200P_ERR_ARG, "null pointer");
  B200P_REQUIRE(V > 0 && V < (1ll << 31) && ldl >= V && vocab_offset >= 0, B200P_ERR_ARG, "bad V/ldl/offset");
  B200P_REQUIRE(B <= 65535, B200P_ERR_ARG, "at most 65535 rows per call (grid.y)");
  B200P_REQUIRE(aligned(rec_out, 16), B200P_ERR_ALIGN, "rec_out must be 16-byte aligned");
  if (B == 0) return B200P_OK;
  cudaError_t e = launch_cluster(argmax_kernel, cluster_for(B, 2), B, SAMP_THREADS, 0, (cudaStream_t)stream, logits,
                                 (int32_t*)nullptr, (float*)nullptr, V, ldl, reinterpret_cast<float4*>(rec_out), (int)vocab_offset);
  if (e != cudaSuccess) return cuda_fail(e, "argmax_kernel(partial)");
  return B200P_OK;
}

extern "C" int b200p_argmax_merge(const float* recs, int32_t* tokens, float* probs_out, int W, int64_t B,
                                  void* stream) {
  PdlClass pdl_cls(32);
  B200P_REQUIRE(recs && tokens, B200P_ERR_ARG, "null pointer");
  B200P_REQUIRE(W > 0 && aligned(recs, 16), B200P_ERR_ARG, "bad shard count / alignment");
  if (B == 0) return B200P_OK;
  launch_pdl(argmax_merge_kernel, (unsigned)((B + 127) / 128), 128, 0, (cudaStream_t)stream, 
      reinterpret_cast<const float4*>(recs), tokens, probs_out, W, B);
  B200P_LAUNCH_CHECK("argmax_merge_kernel");
  return B200P_OK;
}

extern "C" int b200p_sample_top_p(const float* logits, int32_t* tokens, float* probs_out, int64_t B, int64_t V,
                                  int64_t ldl, float temperature, float top_p, uint64_t seed, uint64_t offset,
                                  void* stream) {
  PdlClass pdl_cls(32);
  B200P_REQUIRE(logits && tokens, B200P_ERR_ARG, "null pointer");
  B200P_REQUIRE(V > 0 && V < (1ll << 31) && ldl >= V, B200P_ERR_ARG, "bad V/ldl");
  B200P_REQUIRE(B <= 65535, B200P_ERR_ARG, "at most 65535 rows per call (grid.y)");
  B200P_REQUIRE(temperature > 0.f, B200P_ERR_ARG, "temperature must be > 0 (use b200p_argmax for greedy)");
  if (B == 0) return B200P_OK;
  if (!(top_p > 0.f)) top_p = 1.0f;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(top_p_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TP_SMEM);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(top_p_kernel)");
    attr_set = true;
  }
  cudaError_t le = launch_cluster(top_p_kernel, cluster_for(B, 1), B, TP_THREADS, TP_SMEM, (cudaStream_t)stream, logits, tokens,
                                  probs_out, V, ldl, 1.0f / temperature, top_p, seed, offset);
  if (le != cudaSuccess) return cuda_fail(le, "top_p_kernel");
  return B200P_OK;
}
