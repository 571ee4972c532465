// split-K tensor decode attention (placeholder until the kernel lands).
#include "common.cuh"
namespace b200p {
bool attn_decode_supported(int64_t, int, int, int, int) { return false; }
size_t attn_decode_workspace_bytes(int64_t, int64_t, int, int) { return 0; }
int attn_decode_launch(const void*, const void*, const void*, const int32_t*, void*, int64_t, int64_t, int, int, int,
                       int64_t, float, int, void*, size_t, cudaStream_t) {
  set_error("tensor attention not built");
  return B200P_ERR_ARG;
}
}  // namespace b200p
