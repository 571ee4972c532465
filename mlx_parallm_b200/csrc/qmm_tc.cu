// tcgen05 / TMEM tile path of b200p_qmm (placeholder until the tile kernel lands).
#include "common.cuh"
namespace b200p {
bool qmm_tc_supported(int64_t, int64_t, int64_t, int, int, int, int) { return false; }
size_t qmm_tc_workspace_bytes(int64_t, int64_t, int64_t) { return 0; }
int qmm_tc_launch(const void*, const uint32_t*, const void*, const void*, void*, int64_t, int64_t, int64_t, int64_t,
                  int64_t, int, int, int, void*, size_t, cudaStream_t) {
  set_error("tensor path not built");
  return B200P_ERR_ARG;
}
}  // namespace b200p
