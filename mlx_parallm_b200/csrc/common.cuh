// Shared helpers for libb200parallm (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <utility>

#include "../../include/b200p.h"

namespace b200p {

// ---- error plumbing (thread-local message, negative return codes) -------------------------
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);

#define B200P_REQUIRE(cond, code, ...)            \
  do {                                            \
    if (!(cond)) {                                \
      ::b200p::set_error(__VA_ARGS__);            \
      return (code);                              \
    }                                             \
  } while (0)

#define B200P_LAUNCH_CHECK(what)                                   \
  do {                                                             \
    cudaError_t e__ = cudaGetLastError();                          \
    if (e__ != cudaSuccess) return ::b200p::cuda_fail(e__, what);  \
  } while (0)

static inline bool aligned(const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; }

// ---- storage-type traits ---------------------------------------------------------------------
template <typename T> struct Tr;
template <> struct Tr<float> {
  static __device__ __forceinline__ float to_f(float v) { return v; }
  static __device__ __forceinline__ float from_f(float v) { return v; }
};
template <> struct Tr<__half> {
  static __device__ __forceinline__ float to_f(__half v) { return __half2float(v); }
  static __device__ __forceinline__ __half from_f(float v) { return __float2half_rn(v); }
};
template <> struct Tr<__nv_bfloat16> {
  static __device__ __forceinline__ float to_f(__nv_bfloat16 v) { return __bfloat162float(v); }
  static __device__ __forceinline__ __nv_bfloat16 from_f(float v) { return __float2bfloat16_rn(v); }
};

// dispatch a dtype enum to a template parameter
#define B200P_DISPATCH_DTYPE(dtype, T, ...)                                              \
  switch (dtype) {                                                                       \
    case B200P_F32: { using T = float; __VA_ARGS__; break; }                             \
    case B200P_F16: { using T = __half; __VA_ARGS__; break; }                            \
    case B200P_BF16: { using T = __nv_bfloat16; __VA_ARGS__; break; }                    \
    default: ::b200p::set_error("bad dtype %d", (int)(dtype)); return B200P_ERR_ARG;     \
  }

// ---- small device utilities -------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// streaming 128-bit load that does not pollute L1 (weights / KV are read once)
__device__ __forceinline__ uint4 ldg_stream(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

// streaming 256-bit load (sm_100: LDG.E.256): one whole 32-byte sector per lane.  Two 128-bit loads of the two halves of a
// sector by the same lane make the warp request every sector twice (each instruction touches half of it).
__device__ __forceinline__ void ldg_stream256(const void* p, uint4& lo, uint4& hi) {
  asm volatile("ld.global.nc.L1::no_allocate.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(lo.x), "=r"(lo.y), "=r"(lo.z), "=r"(lo.w), "=r"(hi.x), "=r"(hi.y), "=r"(hi.z), "=r"(hi.w)
               : "l"(p));
}

// exact small-integer -> float via the 2^23 magic (avoids I2F)
__device__ __forceinline__ float u8_to_float(uint32_t q) {
  return __uint_as_float(0x4B000000u | q) - 8388608.0f;
}

// PERM8 activation layout (include/b200p.h): position of element n inside its aligned group of 8 ([0,4,1,5,2,6,3,7])
__host__ __device__ __forceinline__ int perm8_pos(int n) { return (n & ~7) | (((n & 3) << 1) | ((n >> 2) & 1)); }

// Programmatic dependent launch: wait for the producer grid's memory to be visible / allow the
// consumer grid to start its prologue.  No-ops when the launch carries no PDL attribute.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// Every kernel of the library is launched with programmatic stream serialization (unless b200p_set_pdl(0)): the
// next kernel's CTAs may be scheduled as soon as this one's have all passed griddepcontrol.launch_dependents or
// exited, and each kernel executes griddepcontrol.wait before it touches anything a predecessor reads or writes.
bool pdl_enabled();
// kernel class of the C-ABI call being executed (B200P_PDL_MASK selects, per class, which launches carry the PDL attribute):
// 2 rmsnorm, 4 rope+append, 8 attention, 16 silu_mul, 32 embed / dequantize / advance / samplers, 64 GEMV paths
void pdl_set_class(int cls);
struct PdlClass {
  explicit PdlClass(int c) { pdl_set_class(c); }
  ~PdlClass() { pdl_set_class(0); }
};
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl_opt(bool pdl, void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

}  // namespace b200p
