// C-ABI glue: error plumbing, device checks and the path dispatchers of b200p_qmm / b200p_attn.
#include <stdarg.h>
#include <stdlib.h>

#include "common.cuh"

namespace b200p {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
  set_error("%s: %s", what, cudaGetErrorString(e));
  return B200P_ERR_CUDA;
}

// implemented in the kernel translation units
int qmv_launch(const void* x, const uint32_t* w, const void* s, const void* b, void* y, int64_t M, int64_t N,
               int64_t K, int64_t ldx, int64_t ldy, int gs, int bits, int dtype, int epi, int flags, cudaStream_t st);
bool qmm_tc_supported(int64_t M, int64_t N, int64_t K, int gs, int bits, int dtype, int epi);
int qmm_tc_launch(const void* x, const uint32_t* w, const void* s, const void* b, void* y, int64_t M, int64_t N,
                  int64_t K, int64_t ldx, int64_t ldy, int bits, int dtype, int epi, int flags, void* ws, size_t ws_bytes,
                  cudaStream_t st);
size_t qmm_tc_workspace_bytes(int64_t M, int64_t N, int64_t K);
bool qmm_sk_supported(int64_t M, int64_t N, int64_t K, int bits, int epi, const void* y, int64_t ldy);
int qmm_sk_launch(const void* x, const uint32_t* w, const void* s, const void* b, void* y, int64_t M, int64_t N,
                  int64_t K, int64_t ldx, int64_t ldy, int bits, int dtype, int flags, cudaStream_t st);
void qmm_tc_set_min_m(int m);
bool qmv_mma_supported(int64_t M, int64_t N, int64_t K, int gs, int bits, int dtype, int epi);
bool qmm_mma_supported(int64_t M, int64_t N, int64_t K, int gs, int bits, int dtype, int epi, const void* y, int64_t ldy);
int qmm_mma_launch(const void* x, const uint32_t* w, const void* s, const void* b, void* y, int64_t M, int64_t N, int64_t K,
                   int64_t ldx, int64_t ldy, int dtype, int flags, cudaStream_t st);
int qmv_mma_launch(const void* x, const uint32_t* w, const void* s, const void* b, void* y, int64_t M, int64_t N, int64_t K,
                   int64_t ldx, int64_t ldy, int dtype, int epi, int flags, cudaStream_t st);
int attn_generic_launch(const void* q, const void* kc, const void* vc, const int32_t* lengths, void* out,
                        int64_t B, int64_t L, int Hq, int Hkv, int D, int64_t Tcap, float scale, int dtype,
                        int out_perm, cudaStream_t st);
bool attn_decode_supported(int64_t L, int Hq, int Hkv, int D, int dtype);
int attn_decode_launch(const void* q, const void* kc, const void* vc, const int32_t* lengths, void* out,
                       int64_t B, int64_t L, int Hq, int Hkv, int D, int64_t Tcap, float scale, int dtype,
                       int out_perm, void* ws, size_t ws_bytes, cudaStream_t st);
size_t attn_decode_workspace_bytes(int64_t B, int64_t L, int Hq, int D);
bool attn_prefill_supported(int64_t L, int Hq, int Hkv, int D, int dtype);
int attn_prefill_launch(const void* q, const void* kc, const void* vc, const int32_t* lengths, void* out, int64_t B, int64_t L,
                        int Hq, int Hkv, int D, int64_t Tcap, float scale, int dtype, int out_perm, cudaStream_t st);

static const bool g_use_sk = [] { const char* e = getenv("B200P_QMM_SK"); return !e || atoi(e) != 0; }();
static int g_qmm_path = 0;   // 0 auto, 1 CUDA-core gemv, 2 tcgen05 tiles, 3 mma.sync gemv, 4 mma.sync small-matrix GEMM
// pipeline-free mma.sync GEMM (qmm_mma.cu) for the accumulate epilogue above the GEMV's rows: matrices up to this many weights
// (q|k|v 25 M and o_proj 17 M of Llama-3-8B, every tensor-parallel shard; down_proj 59 M and gate|up stay on the tile kernels)
static long long g_mma_max_weights = [] { const char* e = getenv("B200P_MMA_MAX_WEIGHTS"); return e ? atoll(e) : (1ll << 25); }();
// ... and up to this many rows: measured against the tile kernels (profiles/r2_run35_small_gemm_paths.txt) it wins at 8 and 16
// rows (q|k|v 6.2 vs 7.2 us, o_proj 4.3 vs 5.7, TP=8 shards 2.6-4.1 vs 3.8-5.3), ties at 32 and loses at 64, where it is bound
// by the mma.sync rate (q|k|v 14.6 vs 8.9 us)
static int g_gemv_silu_max_m = [] { const char* e = getenv("B200P_GEMV_SILU_MAX_M"); return e ? atoi(e) : 0; }();
static int g_mma_max_m = [] { const char* e = getenv("B200P_MMA_MAX_M"); return e ? atoi(e) : 16; }();
// Auto rule for the mma.sync GEMV, from the measured three-way sweep (profiles/r1_run12_qmm_sweep_three_paths.txt, us per
// launch in a CUDA graph): it wins where the tile kernel's fixed cost dominates -- M <= 4 and matrices up to ~32 M weights
// (q|k|v 6.4 vs 8.4 us, o_proj 5.2 vs 6.5 us at M = 1) -- and loses on the big streams (gate|up 23.9 vs 22.4 us) and as M
// grows (its activation loads are per 16-row warp, not shared through shared memory).  The CUDA-core GEMV loses everywhere.
static int g_gemv_max_m = [] { const char* e = getenv("B200P_GEMV_MAX_M"); return e ? atoi(e) : 4; }();
static long long g_gemv_max_weights = [] { const char* e = getenv("B200P_GEMV_MAX_WEIGHTS"); return e ? atoll(e) : (1ll << 25); }();
static int g_attn_path = 0;  // 0 auto, 1 generic, 2 tensor
// 0 off, 1 every kernel, 2 only the tensor-path GEMM (B200P_PDL overrides the default at load time)
static int g_pdl = [] { const char* e = getenv("B200P_PDL"); return e ? atoi(e) : 2; }();
static int g_sm_count = -1;
static int g_cc_major = -1;

static int query_device() {
  if (g_cc_major >= 0) return B200P_OK;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return cuda_fail(e, "cudaGetDevice");
  cudaDeviceProp p;
  e = cudaGetDeviceProperties(&p, dev);
  if (e != cudaSuccess) return cuda_fail(e, "cudaGetDeviceProperties");
  g_cc_major = p.major;
  g_sm_count = p.multiProcessorCount;
  return B200P_OK;
}

// per-class PDL mask for the non-GEMM kernels (see common.cuh); B200P_PDL=1 means every class
static int g_pdl_mask = [] { const char* e = getenv("B200P_PDL_MASK"); return e ? atoi(e) : 0; }();
static thread_local int g_pdl_class = 0;
void pdl_set_class(int cls) { g_pdl_class = cls; }
bool pdl_enabled() { return g_pdl == 1 || (g_pdl_mask & g_pdl_class) != 0; }
bool pdl_enabled_gemm() { return g_pdl != 0; }

int sm_count() {
  query_device();
  return g_sm_count > 0 ? g_sm_count : 148;
}

}  // namespace b200p

using namespace b200p;

// 0.2: round-2 ABI (PERM8 flags, stream-K / small-matrix GEMM paths, b200p_attn_decode_static, b200p_kv_overflow_count)
extern "C" const char* b200p_version(void) { return "b200parallm 0.2 (sm_100a)"; }
extern "C" const char* b200p_last_error(void) { return g_err; }

extern "C" int b200p_arch_check(void) {
  int rc = query_device();
  if (rc) return rc;
  B200P_REQUIRE(g_cc_major == 10, B200P_ERR_ARCH, "device compute capability %d.x is not sm_100", g_cc_major);
  return B200P_OK;
}

extern "C" int b200p_sm_count(void) {
  int rc = query_device();
  if (rc) return rc;
  return g_sm_count;
}

extern "C" int b200p_set_pdl(int enabled) {
  g_pdl = enabled;
  return B200P_OK;
}

extern "C" int b200p_qmm_set_min_tensor_m(int m) {
  B200P_REQUIRE(m >= 1, B200P_ERR_ARG, "min tensor-path M must be >= 1");
  qmm_tc_set_min_m(m);
  return B200P_OK;
}

extern "C" int b200p_qmm_set_path(int path) {
  B200P_REQUIRE(path >= 0 && path <= 4, B200P_ERR_ARG, "path must be 0..4");
  g_qmm_path = path;
  return B200P_OK;
}

extern "C" int b200p_attn_set_path(int path) {
  B200P_REQUIRE(path >= 0 && path <= 3, B200P_ERR_ARG, "path must be 0..3");
  g_attn_path = path;
  return B200P_OK;
}

extern "C" size_t b200p_qmm_workspace_bytes(int64_t M, int64_t N, int64_t K) {
  return qmm_tc_workspace_bytes(M, N, K);
}

extern "C" int b200p_qmm_tensor_supported(int64_t M, int64_t N, int64_t K, int group_size, int bits, int dtype,
                                          int epilogue) {
  return qmm_tc_supported(M, N, K, group_size, bits, dtype, epilogue) ? 1 : 0;
}

extern "C" int b200p_qmm(const void* x, const uint32_t* w, const void* scales, const void* biases, void* y,
                         int64_t M, int64_t N, int64_t K, int64_t ldx, int64_t ldy, int group_size, int bits,
                         int dtype, int epilogue_flags, void* workspace, size_t workspace_bytes, void* stream) {
  const int epilogue = epilogue_flags & 0xff, flags = epilogue_flags & ~0xff;
  B200P_REQUIRE((flags & ~(B200P_X_PERM8 | B200P_Y_PERM8 | B200P_W_SMALLSCALE)) == 0, B200P_ERR_ARG, "unknown layout flags 0x%x", flags);
  B200P_REQUIRE(!(flags & B200P_X_PERM8) || (bits == 4 && dtype != B200P_F32), B200P_ERR_ARG,
                "B200P_X_PERM8 is the layout of 16-bit activations for 4-bit weights");
  B200P_REQUIRE(!(flags & B200P_Y_PERM8) || ((epilogue == B200P_EPI_STORE || epilogue == B200P_EPI_SILU_MUL) && dtype != B200P_F32),
                B200P_ERR_ARG, "B200P_Y_PERM8 applies to 16-bit outputs (STORE, SILU_MUL)");
  B200P_REQUIRE(x && w && scales && biases && y, B200P_ERR_ARG, "null pointer");
  B200P_REQUIRE(bits == 4 || bits == 8, B200P_ERR_ARG, "bits must be 4 or 8, got %d", bits);
  B200P_REQUIRE(M >= 0 && N > 0 && K > 0 && N % 2 == 0, B200P_ERR_ARG, "bad shape M=%lld N=%lld K=%lld (N must be even)",
                (long long)M, (long long)N, (long long)K);
  B200P_REQUIRE(group_size >= 32 && group_size % 32 == 0 && K % group_size == 0, B200P_ERR_ARG,
                "group_size %d must be a multiple of 32 dividing K=%lld", group_size, (long long)K);
  B200P_REQUIRE(epilogue >= 0 && epilogue <= 3, B200P_ERR_ARG, "bad epilogue %d", epilogue);
  B200P_REQUIRE(dtype >= 0 && dtype <= 2, B200P_ERR_ARG, "bad dtype %d", dtype);
  const size_t xs = dtype == B200P_F32 ? 4 : 2;
  B200P_REQUIRE(aligned(x, 16) && aligned(w, 16) && (ldx * xs) % 16 == 0 && ldx >= K, B200P_ERR_ALIGN,
                "x rows and w must be 16-byte aligned (ldx=%lld)", (long long)ldx);
  B200P_REQUIRE((K * bits / 8) % 16 == 0, B200P_ERR_ALIGN, "packed weight rows must be 16-byte multiples (K=%lld)",
                (long long)K);
  B200P_REQUIRE(ldy >= (epilogue == B200P_EPI_SILU_MUL ? N / 2 : N), B200P_ERR_ARG, "ldy too small");
  if (M == 0) return B200P_OK;
  PdlClass pdl_cls(64);
  cudaStream_t st = (cudaStream_t)stream;
  // (SILU_MUL through the GEMV: 16-bit output, PERM8 allowed; the other GEMV epilogues write fp32 in the natural order)
  const bool mma_ok = qmv_mma_supported(M, N, K, group_size, bits, dtype, epilogue) && aligned(w, 32) &&
                      (epilogue == B200P_EPI_SILU_MUL ? aligned(y, 2) : (aligned(y, 4) && !(flags & B200P_Y_PERM8)));
  if (g_qmm_path == 3) B200P_REQUIRE(mma_ok, B200P_ERR_ARG, "mma.sync GEMV forced but unsupported for this call");
  // up to two rows the GEMV also wins on matrices of twice that size (down_proj 4096 x 14336: 12.8 vs 14.1 us at M = 1,
  // profiles/r1_run34_gemv_splits.txt); gate|up (117 M weights) stays on the tile kernel
  // gate|up with the fused silu(gate) * up epilogue through the GEMV (rows paired gate / up, no K split): up to
  // B200P_GEMV_SILU_MAX_M rows, any size; 0 = never (the default: see profiles/r2_run39_*)
  if (mma_ok && epilogue == B200P_EPI_SILU_MUL && g_qmm_path == 0 && M <= g_gemv_silu_max_m)
    return qmv_mma_launch(x, w, scales, biases, y, M, N, K, ldx, ldy, dtype, epilogue, flags, st);
  if (mma_ok && (g_qmm_path == 3 || (g_qmm_path == 0 && epilogue != B200P_EPI_SILU_MUL && M <= g_gemv_max_m &&
                                     N * K <= (M <= 2 ? 2 * g_gemv_max_weights : g_gemv_max_weights))))
    return qmv_mma_launch(x, w, scales, biases, y, M, N, K, ldx, ldy, dtype, epilogue, flags, st);
  const bool gm_ok = qmm_mma_supported(M, N, K, group_size, bits, dtype, epilogue, y, ldy) && aligned(w, 32);
  if (g_qmm_path == 4) B200P_REQUIRE(gm_ok, B200P_ERR_ARG, "mma.sync small-matrix GEMM forced but unsupported for this call");
  if (gm_ok && (g_qmm_path == 4 || (g_qmm_path == 0 && M > g_gemv_max_m && M <= g_mma_max_m && N * K <= g_mma_max_weights)))
    return qmm_mma_launch(x, w, scales, biases, y, M, N, K, ldx, ldy, dtype, flags, st);
  const bool tc_ok = qmm_tc_supported(M, N, K, group_size, bits, dtype, epilogue);
  if (g_qmm_path == 2) {
    B200P_REQUIRE(tc_ok, B200P_ERR_ARG, "tensor path forced but unsupported for this call");
  }
  // fp32 accumulate epilogue: the stream-K kernel (partial tiles summed by TMA reduce-adds, no workspace, no tail);
  // B200P_QMM_SK=0 keeps everything on the whole-tile kernel (A/B measurements)
  if (tc_ok && g_qmm_path != 1 && g_use_sk && qmm_sk_supported(M, N, K, bits, epilogue, y, ldy))
    return qmm_sk_launch(x, w, scales, biases, y, M, N, K, ldx, ldy, bits, dtype, flags, st);
  if (tc_ok && g_qmm_path != 1) {
    B200P_REQUIRE(workspace && workspace_bytes >= qmm_tc_workspace_bytes(M, N, K), B200P_ERR_ARG,
                  "tensor path needs a workspace of %zu bytes", qmm_tc_workspace_bytes(M, N, K));
    return qmm_tc_launch(x, w, scales, biases, y, M, N, K, ldx, ldy, bits, dtype, epilogue, flags, workspace,
                         workspace_bytes, st);
  }
  return qmv_launch(x, w, scales, biases, y, M, N, K, ldx, ldy, group_size, bits, dtype, epilogue, flags, st);
}

extern "C" size_t b200p_attn_workspace_bytes(int64_t B, int64_t L, int Hq, int D) {
  return attn_decode_workspace_bytes(B, L, Hq, D);
}

extern "C" int b200p_attn(const void* q, const void* kcache, const void* vcache, const int32_t* lengths, void* out,
                          int64_t B, int64_t L, int Hq, int Hkv, int D, int64_t Tcap, float scale, int dtype,
                          int flags, void* workspace, size_t workspace_bytes, void* stream) {
  B200P_REQUIRE(q && kcache && vcache && lengths && out, B200P_ERR_ARG, "null pointer");
  B200P_REQUIRE((flags & ~B200P_OUT_PERM8) == 0, B200P_ERR_ARG, "unknown flags 0x%x", flags);
  const int out_perm = (flags & B200P_OUT_PERM8) ? 1 : 0;
  B200P_REQUIRE(!out_perm || dtype != B200P_F32, B200P_ERR_ARG, "B200P_OUT_PERM8 is a 16-bit layout");
  B200P_REQUIRE(Hq > 0 && Hkv > 0 && Hq % Hkv == 0, B200P_ERR_ARG, "bad head counts %d/%d", Hq, Hkv);
  B200P_REQUIRE(B >= 0 && L >= 0 && Tcap > 0, B200P_ERR_ARG, "bad shape");
  B200P_REQUIRE(dtype >= 0 && dtype <= 2, B200P_ERR_ARG, "bad dtype %d", dtype);
  if (B * L == 0) return B200P_OK;
  PdlClass pdl_cls(8);
  cudaStream_t st = (cudaStream_t)stream;
  const bool tc_ok = attn_decode_supported(L, Hq, Hkv, D, dtype);
  const bool pf_ok = attn_prefill_supported(L, Hq, Hkv, D, dtype);
  if (g_attn_path == 2) B200P_REQUIRE(tc_ok || pf_ok, B200P_ERR_ARG, "tensor attention forced but unsupported for this call");
  if (pf_ok && g_attn_path != 1)
    return attn_prefill_launch(q, kcache, vcache, lengths, out, B, L, Hq, Hkv, D, Tcap, scale, dtype, out_perm, st);
  if (tc_ok && g_attn_path != 1) {
    B200P_REQUIRE(workspace && workspace_bytes >= attn_decode_workspace_bytes(B, L, Hq, D), B200P_ERR_ARG,
                  "attention needs a workspace of %zu bytes", attn_decode_workspace_bytes(B, L, Hq, D));
    return attn_decode_launch(q, kcache, vcache, lengths, out, B, L, Hq, Hkv, D, Tcap, scale, dtype, out_perm, workspace,
                              workspace_bytes, st);
  }
  return attn_generic_launch(q, kcache, vcache, lengths, out, B, L, Hq, Hkv, D, Tcap, scale, dtype, out_perm, st);
}
