// Logits -> token: greedy argmax and temperature / nucleus (top-p) sampling.
// One CTA per batch row; the row (V floats, <= 600 KB) is re-read from L2 on each pass.
#include "common.cuh"

namespace b200p {

constexpr int SAMP_THREADS = 1024;

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

__global__ void __launch_bounds__(SAMP_THREADS)
argmax_kernel(const float* __restrict__ logits, int32_t* __restrict__ tokens, float* __restrict__ probs_out,
              int64_t V, int64_t ldl, float4* __restrict__ rec_out, int vocab_offset) {
  __shared__ float shv[32];
  __shared__ int shi[32];
  __shared__ float shm[32], shs[32];
  pdl_wait();
  const float* row = logits + (int64_t)blockIdx.x * ldl;
  float best = -INFINITY;
  int bi = 0x7fffffff;
  float om = -INFINITY, os = 0.f;
  const bool want_p = probs_out != nullptr || rec_out != nullptr;
  const bool vec = (ldl % 4 == 0) && ((reinterpret_cast<uintptr_t>(logits) & 15) == 0);
  int64_t V4 = vec ? V / 4 : 0;
  for (int64_t i = threadIdx.x; i < V4; i += blockDim.x) {
    float4 v = reinterpret_cast<const float4*>(row)[i];
    int base = (int)(4 * i);
    // NaN never wins (comparisons false), matching a total order that ignores NaN
    if (v.x > best) { best = v.x; bi = base; }
    if (v.y > best) { best = v.y; bi = base + 1; }
    if (v.z > best) { best = v.z; bi = base + 2; }
    if (v.w > best) { best = v.w; bi = base + 3; }
    if (want_p) { osm_push(om, os, v.x); osm_push(om, os, v.y); osm_push(om, os, v.z); osm_push(om, os, v.w); }
  }
  for (int64_t i = V4 * 4 + threadIdx.x; i < V; i += blockDim.x) {
    float v = row[i];
    if (v > best) { best = v; bi = (int)i; }
    if (want_p) osm_push(om, os, v);
  }
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
    const int tok = (bi == 0x7fffffff) ? 0 : bi;
    float m = -INFINITY, s = 0.f;
    if (want_p)
      for (int w = 0; w < (int)(blockDim.x >> 5); ++w) osm_merge(m, s, shm[w], shs[w]);
    if (rec_out != nullptr) {   // vocab-parallel shard: publish (max, sum exp(l - max), global index)
      rec_out[blockIdx.x] = make_float4(best, s, __int_as_float(tok + vocab_offset), 0.f);
    } else {
      tokens[blockIdx.x] = tok;
      if (probs_out != nullptr) probs_out[blockIdx.x] = (s > 0.f) ? __expf(best - m) / s : 0.f;
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

// Nucleus sampling, one CTA per row.
//  1. mx = max(logit), Z = sum exp((logit-mx)/temp)       -> p_i = exp((logit_i-mx)/temp)/Z
//  2. radix-select (4 x 8-bit digits over the float bits of p_i, which are monotone for p >= 0) the
//     value v* of the first sorted element whose inclusive cumulative mass exceeds top_p; the nucleus
//     is {p_i > v*}; its mass is `kept`.
//  3. u ~ U(0,kept): pick the token where the running mass (in index order) over the nucleus crosses u.
__global__ void __launch_bounds__(SAMP_THREADS)
top_p_kernel(const float* __restrict__ logits, int32_t* __restrict__ tokens, float* __restrict__ probs_out,
             int64_t V, int64_t ldl, float inv_temp, float top_p, uint64_t seed, uint64_t offset) {
  __shared__ float sh[32];
  __shared__ float hist[256];
  __shared__ uint32_t s_prefix;
  __shared__ float s_above;
  __shared__ int s_done;
  __shared__ float s_scan[SAMP_THREADS];
  __shared__ int s_tok;
  pdl_wait();
  const int tid = threadIdx.x;
  const float* row = logits + (int64_t)blockIdx.x * ldl;

  float mx = -INFINITY;
  for (int64_t i = tid; i < V; i += blockDim.x) mx = fmaxf(mx, row[i]);
  mx = block_reduce(mx, sh, true);
  float z = 0.f, z1 = 0.f;
  for (int64_t i = tid; i < V; i += blockDim.x) {
    float d = row[i] - mx;
    z += __expf(d * inv_temp);
    z1 += __expf(d);
  }
  z = block_reduce(z, sh, false);
  z1 = block_reduce(z1, sh, false);
  const float inv_z = 1.0f / z;

  uint32_t thr_bits = 0u;  // nucleus = {bits(p) > thr_bits}; 0 keeps every p > 0
  float kept = 1.0f;
  bool empty = false;
  if (top_p < 1.0f) {
    if (tid == 0) { s_prefix = 0u; s_above = 0.f; s_done = 0; }
    __syncthreads();
    for (int pass = 0; pass < 4; ++pass) {
      const int shift = 24 - 8 * pass;
      if (tid < 256) hist[tid] = 0.f;
      __syncthreads();
      const uint32_t prefix = s_prefix;
      const uint32_t pmask = pass == 0 ? 0u : (0xffffffffu << (shift + 8));
      for (int64_t i = tid; i < V; i += blockDim.x) {
        float p = __expf((row[i] - mx) * inv_temp) * inv_z;
        uint32_t bits = __float_as_uint(p);
        if ((bits & pmask) == prefix) atomicAdd(&hist[(bits >> shift) & 0xffu], p);
      }
      __syncthreads();
      if (tid == 0 && !s_done) {
        float above = s_above;
        int d = 255;
        for (; d >= 0; --d) {
          if (above + hist[d] > top_p) break;
          above += hist[d];
        }
        s_above = above;
        if (d < 0) {
          // nothing in this bucket crosses top_p (pass 0: the whole row fits; later passes: only
          // through atomic-order rounding) -> every element carrying this prefix is kept
          s_done = 1;
          s_prefix = prefix == 0u ? 0u : prefix - 1u;
        } else {
          s_prefix = prefix | ((uint32_t)d << shift);
        }
      }
      __syncthreads();
    }
    thr_bits = s_prefix;
    kept = s_above;
    if (!s_done && kept <= 0.f) empty = true;  // even the largest p exceeds top_p (reference: 0/0)
  }

  int tok = -1;
  if (empty) {
    float best = -INFINITY;
    int bi = 0x7fffffff;
    for (int64_t i = tid; i < V; i += blockDim.x) {
      float v = row[i];
      if (v > best) { best = v; bi = (int)i; }
    }
    __shared__ int shi[32];
    block_argmax(best, bi, sh, shi);
    tok = bi;
  } else {
    // contiguous chunk per thread -> block scan -> owner walks its chunk
    const int64_t chunk = (V + blockDim.x - 1) / blockDim.x;
    const int64_t lo = (int64_t)tid * chunk, hi = min(lo + chunk, V);
    float local = 0.f;
    for (int64_t i = lo; i < hi; ++i) {
      float p = __expf((row[i] - mx) * inv_temp) * inv_z;
      if (__float_as_uint(p) > thr_bits) local += p;
    }
    s_scan[tid] = local;
    __syncthreads();
    // inclusive Hillis-Steele scan (1024 elements, 10 steps)
    for (int o = 1; o < SAMP_THREADS; o <<= 1) {
      float add = tid >= o ? s_scan[tid - o] : 0.f;
      __syncthreads();
      s_scan[tid] += add;
      __syncthreads();
    }
    const float total = s_scan[SAMP_THREADS - 1];
    const float u = philox_uniform(seed, offset, blockIdx.x) * total;
    if (tid == 0) s_tok = -1;
    __syncthreads();
    const float before = tid ? s_scan[tid - 1] : 0.f;
    if (local > 0.f && u >= before && u < s_scan[tid]) {
      float run = before;
      int pick = -1, last = -1;
      for (int64_t i = lo; i < hi; ++i) {
        float p = __expf((row[i] - mx) * inv_temp) * inv_z;
        if (__float_as_uint(p) > thr_bits) {
          run += p;
          last = (int)i;
          if (pick < 0 && u < run) pick = (int)i;
        }
      }
      s_tok = pick >= 0 ? pick : last;
    }
    __syncthreads();
    tok = s_tok;
    if (tok < 0) {  // u landed on a rounding seam: take the last nucleus token
      int cand = -1;
      for (int64_t i = hi - 1; i >= lo; --i) {
        float p = __expf((row[i] - mx) * inv_temp) * inv_z;
        if (__float_as_uint(p) > thr_bits) { cand = (int)i; break; }
      }
      __shared__ int s_last;
      if (tid == 0) s_last = -1;
      __syncthreads();
      if (cand >= 0) atomicMax(&s_last, cand);
      __syncthreads();
      tok = s_last < 0 ? 0 : s_last;
    }
  }
  if (tid == 0) {
    tokens[blockIdx.x] = tok;
    if (probs_out) probs_out[blockIdx.x] = __expf(row[tok] - mx) / z1;
  }
}

}  // namespace b200p

using namespace b200p;

extern "C" int b200p_argmax(const float* logits, int32_t* tokens, float* probs_out, int64_t B, int64_t V,
                            int64_t ldl, void* stream) {
  PdlClass pdl_cls(32);
  B200P_REQUIRE(logits && tokens, B200P_ERR_ARG, "null pointer");
  B200P_REQUIRE(V > 0 && V < (1ll << 31) && ldl >= V, B200P_ERR_ARG, "bad V/ldl");
  if (B == 0) return B200P_OK;
  launch_pdl(argmax_kernel, (unsigned)B, SAMP_THREADS, 0, (cudaStream_t)stream, logits, tokens, probs_out, V, ldl, nullptr, 0);
  B200P_LAUNCH_CHECK("argmax_kernel");
  return B200P_OK;
}

extern "C" int b200p_argmax_partial(const float* logits, float* rec_out, int64_t B, int64_t V, int64_t ldl,
                                    int32_t vocab_offset, void* stream) {
  PdlClass pdl_cls(32);
  B200P_REQUIRE(logits && rec_out, B200P_ERR_ARG, "null pointer");
  B200P_REQUIRE(V > 0 && V < (1ll << 31) && ldl >= V && vocab_offset >= 0, B200P_ERR_ARG, "bad V/ldl/offset");
  B200P_REQUIRE(aligned(rec_out, 16), B200P_ERR_ALIGN, "rec_out must be 16-byte aligned");
  if (B == 0) return B200P_OK;
  launch_pdl(argmax_kernel, (unsigned)B, SAMP_THREADS, 0, (cudaStream_t)stream, logits, nullptr, nullptr, V, ldl,
                                                                       reinterpret_cast<float4*>(rec_out), vocab_offset);
  B200P_LAUNCH_CHECK("argmax_kernel(partial)");
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
  B200P_REQUIRE(temperature > 0.f, B200P_ERR_ARG, "temperature must be > 0 (use b200p_argmax for greedy)");
  if (B == 0) return B200P_OK;
  if (!(top_p > 0.f)) top_p = 1.0f;
  launch_pdl(top_p_kernel, (unsigned)B, SAMP_THREADS, 0, (cudaStream_t)stream, logits, tokens, probs_out, V, ldl,
                                                                      1.0f / temperature, top_p, seed, offset);
  B200P_LAUNCH_CHECK("top_p_kernel");
  return B200P_OK;
}
