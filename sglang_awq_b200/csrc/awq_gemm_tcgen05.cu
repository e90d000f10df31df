#include "common.cuh"
#include "host_common.h"
extern "C" int awq_b200_gemm(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K,
                             int64_t N, int dtype, void* workspace, int64_t workspace_bytes, void* stream) {
  return awq::set_error(4, "awq_b200_gemm: not built yet");
}
extern "C" int awq_b200_linear_route(int64_t M, int64_t K, int64_t N) { return M <= 64 ? 0 : 1; }
extern "C" int awq_b200_linear(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K,
                               int64_t N, int dtype, void* workspace, int64_t workspace_bytes, void* stream) {
  if (awq_b200_linear_route(M, K, N) == 0)
    return awq_b200_gemv(x, tiles, bias, y, M, K, N, dtype, workspace, workspace_bytes, stream);
  return awq_b200_gemm(x, tiles, bias, y, M, K, N, dtype, workspace, workspace_bytes, stream);
}
