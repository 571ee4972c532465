/*
 * b200p.h -- C-ABI of libb200parallm.so: the sm_100a kernels under the batched decode step.
 *
 * The reference (misanthropic-ai/mlx_parallm) has no FFI of its own: its arithmetic is reached
 * through mlx.core (nanobind).  Each entry point below therefore cites the *MLX call site in the
 * reference* that it replaces (paths relative to the reference root).  A maintainer binds these
 * with ctypes; mlx_parallm_b200/_lib.py is that binding.
 *
 * Conventions
 *   - every function returns 0 on success or a negative B200P_ERR_*; b200p_last_error() gives a
 *     thread-local message.
 *   - all pointers are CALLER-OWNED DEVICE pointers unless named host_*; nothing is allocated,
 *     freed or synchronised inside; every launch is ordered on `stream` (a cudaStream_t passed as
 *     void*) and is CUDA-graph capturable.
 *   - `dtype` is the storage type of activations, scales, biases and KV (MLX unifies them through
 *     result_type before quantized_matmul): B200P_F32 / F16 / BF16.
 *   - quantized weights are in MLX layout: w uint32 [N, K*bits/32] (element j of a word in bits
 *     [j*bits,(j+1)*bits)), scales/biases [N, K/group_size]; group_size 64 or 128... (any multiple
 *     of 32 dividing K for the generic path; the tensor path requires 64), bits in {4, 8}.
 *   - there is NO CPU fallback: without an sm_100 device b200p_arch_check() fails and every
 *     launch returns B200P_ERR_ARCH / B200P_ERR_CUDA.
 */
#ifndef B200P_H
#define B200P_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200P_OK          0
#define B200P_ERR_ARG    -1   /* bad shape / null pointer / unsupported combination */
#define B200P_ERR_ALIGN  -2   /* pointer or leading dimension not aligned as required */
#define B200P_ERR_ARCH   -3   /* device is not sm_100 */
#define B200P_ERR_CUDA   -4   /* CUDA runtime error (message in b200p_last_error) */

enum { B200P_F32 = 0, B200P_F16 = 1, B200P_BF16 = 2 };

/* dequantize rounding (oracle/mlx_ops.py MODE_*) */
enum { B200P_DQ_F32_MULADD = 0, B200P_DQ_FMA = 1, B200P_DQ_T_MULADD = 2 };

/* b200p_qmm epilogues */
enum {
  B200P_EPI_STORE = 0,        /* y[M,N] (dtype) = acc                                   */
  B200P_EPI_STORE_F32 = 1,    /* y[M,N] (float) = acc                (lm_head logits)    */
  B200P_EPI_ADD_RESIDUAL = 2, /* y[M,N] (float) += acc               (o_proj, down_proj) */
  B200P_EPI_SILU_MUL = 3      /* y[M,N/2] (dtype) = silu(acc[:, :N/2]) * acc[:, N/2:]
                                 rows [0,N/2) of w are gate_proj, rows [N/2,N) up_proj   */
};

/* Activation-layout flags OR-ed into the `epilogue` argument of b200p_qmm, and passed as `flags` to the producer
 * kernels (b200p_rmsnorm, b200p_attn).  PERM8: inside every aligned group of 8 consecutive elements of a row
 * the values sit in the order [0,4,1,5,2,6,3,7] -- the order in which the nibbles of a 4-bit weight word unpack as
 * packed 16-bit pairs -- so the tensor path skips the re-pairing PRMTs.  Purely internal to a chain of our kernels:
 * a tensor written with Y_PERM8 / OUT_PERM8 must be consumed with X_PERM8. */
#define B200P_X_PERM8   0x100   /* x of b200p_qmm is PERM8 (4-bit weights only)                          */
#define B200P_Y_PERM8   0x200   /* 16-bit y of b200p_qmm (STORE, SILU_MUL) is written PERM8              */
#define B200P_OUT_PERM8 0x1     /* flags of b200p_rmsnorm / b200p_attn: write the 16-bit output PERM8 */
/* Caller's promise about the weights of a b200p_qmm call, OR-ed into `epilogue` like the layout flags: every scale of the
 * matrix satisfies |s| < 0.0625 (checked once at load time by the host: mlx_parallm_b200/models/_transformer.py).  F16
 * calls then dequantise with ONE half2 FMA per code pair on subnormal operands (csrc/dequant.cuh) -- same rounding of
 * q*s+b, two thirds of the instructions.  Ignored for BF16 / F32. */
#define B200P_W_SMALLSCALE 0x400

const char* b200p_version(void);
const char* b200p_last_error(void);
/* 0 iff the current device is compute capability 10.x */
int b200p_arch_check(void);
int b200p_sm_count(void);
/* Programmatic dependent launch for every kernel of the library (default on): kernel N+1's prologue and weight
 * prefetch overlap kernel N's tail; each kernel executes griddepcontrol.wait before touching its inputs. */
int b200p_set_pdl(int enabled);

/* mx.dequantize(w, scales, biases, group_size, bits)            [nn.QuantizedEmbedding.__call__,
 * models/llama.py:212, models/qwen3.py:166; bit-exactness target of the north star]
 * out[N,K] in `dtype`. */
int b200p_dequantize(const uint32_t* w, const void* scales, const void* biases, void* out,
                     int64_t N, int64_t K, int group_size, int bits, int dtype, int mode,
                     void* stream);

/* mx.quantize(w, group_size, bits)                                  [nn.quantize, utils.py:886-905; convert, utils.py:934-981]
 * w `dtype` [N,K] -> wq uint32 [N, K*bits/32], scales / biases `dtype` [N, K/group_size]; every step rounded to `dtype`
 * where MLX's composed ops round it: bit-exact against oracle/mlx_ops.py quantize.
 * group_size 32, 64 or 128. */
int b200p_quantize(const void* w, uint32_t* wq, void* scales, void* biases, int64_t N, int64_t K, int group_size,
                   int bits, int dtype, void* stream);

/* embed_tokens(ids): gather rows + dequantize                    [models/llama.py:212]
 * ids int32 [M]; out float [M,K] if out_f32 else `dtype` [M,K] (values are the dtype-rounded
 * dequantised weights either way). */
int b200p_embed(const int32_t* ids, const uint32_t* w, const void* scales, const void* biases,
                void* out, int64_t M, int64_t V, int64_t K, int group_size, int bits, int dtype,
                int mode, int out_f32, void* stream);

/* mx.quantized_matmul(x, w, scales, biases, transpose=True)       [nn.QuantizedLinear.__call__:
 * models/llama.py:93,143,165,250-252; models/qwen3.py:63,115]
 * x `dtype` [M,K] row-major (ldx elements between rows), y per `epilogue` (ldy elements between
 * rows).  fp32 accumulation.  Internally: the mma.sync register GEMV (4-bit g64, 16-bit, M <= 16, fp32 outputs), the
 * tcgen05 tile path (F16/BF16, group 64) or the generic CUDA-core GEMV (any dtype / shape).  `workspace` (>= b200p_qmm_workspace_bytes) is
 * required by the split-K tensor path; may be NULL for the GEMV path. */
int b200p_qmm(const void* x, const uint32_t* w, const void* scales, const void* biases, void* y,
              int64_t M, int64_t N, int64_t K, int64_t ldx, int64_t ldy, int group_size, int bits,
              int dtype, int epilogue, void* workspace, size_t workspace_bytes, void* stream);
size_t b200p_qmm_workspace_bytes(int64_t M, int64_t N, int64_t K);
/* 1 if the tcgen05 tile path can run this call, else 0 */
int b200p_qmm_tensor_supported(int64_t M, int64_t N, int64_t K, int group_size, int bits, int dtype,
                               int epilogue);
/* force a path for tests/benchmarks: 0 = auto, 1 = CUDA-core GEMV only, 2 = tcgen05 tile kernels only, 3 = mma.sync
 * GEMV only, 4 = mma.sync small-matrix GEMM only (csrc/qmm_mma.cu: ADD_RESIDUAL, M <= 64, N % 128 == 0, K % 256 == 0)
 * (2, 3 and 4: error if the call is unsupported on that path) */
int b200p_qmm_set_path(int path);
/* smallest M routed to the tensor path in auto mode (default 1: measured, scripts/qmm_sweep.py, profiles/r1_run11_qmm_sweep_gemv_vs_tensor.txt) */
int b200p_qmm_set_min_tensor_m(int m);

/* mx.fast.rms_norm(x, w, eps)                                      [nn.RMSNorm: models/llama.py:187,
 * 189,231]  x float [M,H] (the fp32 residual stream) -> y `dtype` [M,H]; w `dtype` [H]. */
int b200p_rmsnorm(const float* x, const void* w, void* y, int64_t M, int64_t H, float eps,
                  int dtype, int flags, void* stream);

/* per-head q/k RMSNorm (qwen3 only) + mx.fast.rope at per-row offsets + cache append
 *                      [models/qwen3.py:65-70; models/llama.py:107-117; models/base.py:119-140]
 * qkv `dtype` [B*L, (Hq+2*Hkv)*D] (q heads, then k heads, then v heads);
 * q_out `dtype` [B, L, Hq, D]; kcache/vcache `dtype` [B, Hkv, Tcap, D];
 * lengths int32 [B] = tokens already in the cache per row (position of token l is lengths[b]+l);
 * q_norm_w / k_norm_w `dtype` [D] or NULL.  Does NOT advance lengths.
 * qkv_f32 = 1: qkv is instead the float accumulation buffer of the fused projection (b200p_qmm with
 * B200P_EPI_ADD_RESIDUAL into a zeroed buffer); values are rounded to `dtype` on read and the buffer is re-zeroed. */
int b200p_rope_kv_append(void* qkv, void* q_out, void* kcache, void* vcache,
                         const int32_t* lengths, const void* q_norm_w, const void* k_norm_w,
                         int64_t B, int64_t L, int Hq, int Hkv, int D, int64_t Tcap,
                         float rope_base, float rope_scale, float norm_eps, int dtype,
                         int qkv_f32, void* stream);
/* nn.silu(gate) * up                                                  [models/llama.py:165]
 * gu float [M, 2*I] (ldg floats between rows): columns [0,I) = gate_proj, [I,2I) = up_proj sums, as accumulated by
 * b200p_qmm(..., B200P_EPI_ADD_RESIDUAL) of the fused gate|up matrix into a ZEROED buffer (stream-K partial tiles
 * then need no second pass); out `dtype` [M, I] (ldo), optionally PERM8 (flags = B200P_OUT_PERM8).
 * The kernel re-zeroes gu behind its read, so the buffer is ready for the next step. */
int b200p_silu_mul(float* gu, void* out, int64_t M, int64_t I, int64_t ldg, int64_t ldo, int dtype, int flags,
                   void* stream);
/* LoRA low-rank update (mlx_lm LoRALinear: y = linear(x) + (scale * (x @ lora_a) @ lora_b).astype(x.dtype))
 *                                       [adapters on q_proj / v_proj: lora_init.py:63-73; SURVEY A.9, row f4]
 * x `dtype` [M,K] (ldx); At `dtype` [16, K] = lora_a transposed, rank padded to 16 with zero rows (k axis permuted like x
 * when x is PERM8); Bt `dtype` [N, 16] = lora_b transposed, padded likewise; y[m, col0 + n] += update for n < N:
 * out_f32 = 1 -> y is the float accumulation buffer of the projection (ldy floats), else its `dtype` output (the update
 * is rounded to `dtype` first, as MLX's astype does).  16-bit types only. */
int b200p_lora_add(const void* x, const void* At, const void* Bt, void* y, int64_t M, int64_t N, int64_t K, int64_t ldx,
                   int64_t ldy, int64_t col0, float scale, int dtype, int out_f32, void* stream);
/* Rows that b200p_rope_kv_append could NOT append because lengths[b] + l reached Tcap (their K / V are dropped and the
 * attention of that step reads a stale slot): *count = number of (token, V head) pairs since the last reset on the current
 * device; reset != 0 clears it.  Synchronises `stream`; not capturable.  The Python surface (models/base.py) grows the cache
 * before every step and checks this at generation boundaries; direct C-ABI callers should do the same.
 *                      [no MLX counterpart: the reference's caches grow inside update_and_fetch, models/base.py:104-117] */
int b200p_kv_overflow_count(unsigned int* count, int reset, void* stream);
/* lengths[b] += delta (device side, keeps the decode graph replayable) */
int b200p_advance_lengths(int32_t* lengths, int64_t B, int32_t delta, void* stream);

/* mx.fast.scaled_dot_product_attention for cached decode / chunked prefill
 *                      [models/llama.py:139-141, models/qwen3.py:111-113; causal rule of
 *                       models/base.py:17-40 applied by bound, not by an additive -1e9 mask]
 * q `dtype` [B, L, Hq, D]; caches as above; query (b,l) attends to keys [0, lengths[b]+l+1)
 * (lengths = tokens in cache BEFORE this call's L tokens were appended).
 * out `dtype` [B, L, Hq*D].  workspace: float, >= b200p_attn_workspace_bytes. */
int b200p_attn(const void* q, const void* kcache, const void* vcache, const int32_t* lengths,
               void* out, int64_t B, int64_t L, int Hq, int Hkv, int D, int64_t Tcap,
               float scale, int dtype, int flags, void* workspace, size_t workspace_bytes, void* stream);
size_t b200p_attn_workspace_bytes(int64_t B, int64_t L, int Hq, int D);
/* Decode attention (L == 1) that allocates NOTHING: b200p_attn's split-K decode kernel with its split partials in a
 * file-scope device array of the library instead of a caller-allocated workspace.
 *                      [no counterpart in the reference: added for callers that
 *                       cannot call cudaMalloc -- code under CUDA graph capture, tight serving loops, harnesses that forbid it]
 * Same arguments and result as b200p_attn with L = 1 (bit-identical: same kernels, same split count); 16-bit caches,
 * D == 128, Hq / Hkv <= 16 only.  The scratch is B200P_ATTN_STATIC_SCRATCH_BYTES (16 MiB, held by every GPU that loads the
 * library until the process exits), i.e. B * Hq <= 992; larger calls are rejected with B200P_ERR_ARG -- use b200p_attn.
 * The call makes no cudaMalloc / cudaFree and calls no allocating routine; its first use on a device resolves the array's
 * device address (cudaGetSymbolAddress), so make that first call outside a stream capture.
 * Every call shares the array, which keeps the last call's contents (nothing in it has to start at zero: each call writes
 * every record it later reads).  Calls that overlap on the GPU -- two streams, concurrent graph branches -- RACE on it:
 * serialise them. */
#define B200P_ATTN_STATIC_SCRATCH_BYTES (16u << 20)
int b200p_attn_decode_static(const void* q, const void* kcache, const void* vcache, const int32_t* lengths,
                             void* out, int64_t B, int Hq, int Hkv, int D, int64_t Tcap,
                             float scale, int dtype, int flags, void* stream);
/* 0 = auto, 1 = generic CUDA-core kernel only, 2 = tensor (mma) kernels only: split-K decode for L == 1, causal flash
 * prefill for L > 1 (16-bit, head_dim 128) */
int b200p_attn_set_path(int path);

/* mx.argmax(logits, axis=-1, keepdims=True)                         [utils.py:350-352,363]
 * logits float [B, V] (ldl floats between rows) -> tokens int32 [B]; lowest index among ties.
 * probs_out (optional, float [B]) receives softmax(logits)[b, token] computed in the same pass.
 * B <= 65535.  Each row is split over a thread-block cluster of 8 / 4 / 2 / 1 CTAs so that B x cluster fills the SMs once
 * (env B200P_SAMPLER_CLUSTER forces a size); the same holds for b200p_argmax_partial and b200p_sample_top_p. */
int b200p_argmax(const float* logits, int32_t* tokens, float* probs_out, int64_t B, int64_t V,
                 int64_t ldl, void* stream);

/* Vocab-parallel greedy head (tensor parallelism; new -- the reference is single-device): each rank
 * reduces ITS logits shard [B, V_local] to a 16-byte record per row {max, sum exp(l - max), global
 * argmax index (int bits), 0}; after an all-gather of the records b200p_argmax_merge picks the token
 * (highest max, lowest global index among ties = mx.argmax's rule) and softmax(logits)[token]. */
int b200p_argmax_partial(const float* logits, float* rec_out, int64_t B, int64_t V, int64_t ldl,
                         int32_t vocab_offset, void* stream);
int b200p_argmax_merge(const float* recs /* [W, B, 4] */, int32_t* tokens, float* probs_out, int W,
                       int64_t B, void* stream);

/* top_p_sampling(logits, top_p, temperature) / mx.random.categorical(logits/temp)
 *                                                        [sample_utils.py:3-39; utils.py:354-361]
 * Draws from the renormalised nucleus {tokens whose inclusive sorted cumulative mass <= top_p}
 * (top_p >= 1 -> plain categorical).  Philox counter RNG keyed by (seed, offset, row).
 * If the nucleus is empty (p_max > top_p; the reference divides 0/0) the argmax is returned.
 * Selection is by value: a nucleus that ends inside a run of exactly equal probabilities drops the run.
 * probs_out (optional, float [B]) receives softmax(logits)[b, token] (untempered, utils.py:363). */
int b200p_sample_top_p(const float* logits, int32_t* tokens, float* probs_out, int64_t B,
                       int64_t V, int64_t ldl, float temperature, float top_p, uint64_t seed,
                       uint64_t offset, void* stream);

/* ---- tensor-parallel exchange over NVLink peer memory (new; the reference is single-device) -----------------
 * Setup calls (host side, synchronous, NOT graph-capturable; the only entry points that allocate):
 *   b200p_symm_alloc   cudaMalloc + zero a buffer peers may map;   b200p_symm_free releases it
 *   b200p_ipc_export   64-byte cudaIpcMemHandle of such a buffer (ship it to the peers with any host transport)
 *   b200p_ipc_import   map a peer's buffer into this process (peer access enabled lazily); b200p_ipc_close unmaps it
 * Each rank owns one flag array of b200p_allreduce_flag_bytes() bytes (zeroed) mapped by every peer. */
#define B200P_AR_MAX_CTAS 1024
int b200p_symm_alloc(void** ptr, size_t bytes);
int b200p_symm_free(void* ptr);
int b200p_ipc_export(void* ptr, void* handle64);
int b200p_ipc_import(const void* handle64, void** ptr);
int b200p_ipc_close(void* ptr);
size_t b200p_allreduce_flag_bytes(void);
/* cap on the exchange grid (0 = auto: the number of CTAs the device can hold at once); tests / tuning */
int b200p_allreduce_set_max_ctas(int n);

/* One-shot all-reduce of the row-parallel partial sums + residual add + the next RMSNorm, one kernel per rank
 *                      [replaces, under TP, the residual adds of models/llama.py:188,190 and the nn.RMSNorm that follows
 *                       (models/llama.py:187,189,231) around a sum over ranks the single-device reference does not need]
 * peer_partial[p] / peer_flags[p]: HOST arrays of `world` device pointers -- rank p's fp32 partial buffer [M, ld_partial]
 * and flag array as mapped in this process (entry `rank` is the local buffer).  h float [M,H] += sum_p partial_p (rank
 * order: bit-identical on all ranks); y `dtype` [M,H] = rms_norm(h, norm_w, eps) (both NULL to skip), optionally PERM8
 * (flags = B200P_OUT_PERM8); zero_buf (local fp32 [M, ld_partial], may be NULL) is zeroed behind the handshake: pass the
 * OTHER partial buffer of a two-buffer rotation (see csrc/allreduce.cu for the protocol and why two buffers suffice).
 * Every rank must issue the same sequence of exchange calls with the same M. */
int b200p_allreduce_residual_rmsnorm(const void* const* peer_partial, void* const* peer_flags, int world, int rank,
                                     float* h, const void* norm_w, void* y, float* zero_buf, int64_t M, int64_t H,
                                     int64_t ld_partial, float eps, int dtype, int flags, void* stream);
/* Two-shot variant with a ROW-SHARDED residual stream (csrc/allreduce.cu): rank r owns rows [r*rpr, (r+1)*rpr), rpr =
 * ceil(M/world); it pulls only its rows of the partials, updates its rows of h (other rows of h go stale on this rank) and
 * pushes rms_norm(h rows) into EVERY rank's x buffer (peer_x[p]: `dtype` [M,H] as mapped here).  1/world of the one-shot
 * traffic; two handshakes.  norm_w is required (x is the product).  Same calling discipline as above. */
int b200p_allreduce_rows_rmsnorm(const void* const* peer_partial, void* const* peer_x, void* const* peer_flags,
                                 int world, int rank, float* h, const void* norm_w, float* zero_buf, int64_t M,
                                 int64_t H, int64_t ld_partial, float eps, int dtype, int flags, void* stream);
/* dst[p * bytes_per_rank ...] = peer p's src block, through peer memory (vocab-parallel greedy head: the W records of
 * b200p_argmax_partial, then b200p_argmax_merge).  bytes_per_rank % 16 == 0. */
int b200p_peer_allgather(const void* const* peer_src, void* const* peer_flags, int world, int rank, void* dst,
                         int64_t bytes_per_rank, void* stream);
/* host_out[0] = 1 if a handshake of this rank ever timed out (a peer stopped), else 0; synchronises `stream` */
int b200p_peer_status(const void* flags_local, uint32_t* host_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200P_H */
