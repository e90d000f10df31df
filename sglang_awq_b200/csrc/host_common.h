// host_common.h -- error plumbing shared by the extern "C" entry points.
#pragma once

#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>

#include "../../include/awq_b200.h"

namespace awq {

// sets the thread-local error message; returns the code to hand back to the caller
int set_error(int code, const char* fmt, ...);
void clear_error();
// SM count of the current device (cached per device)
int sm_count();
bool pdl_enabled();
// bytes of a weight array the decode kernels prefetch into L2 in their prologue (0 = off)
int64_t l2_prefetch_bytes(int64_t total_bytes);
// scratch the tcgen05 kernel needs for split-K at problem sizes up to (max_M, max_N)
int64_t gemm_workspace_bytes(int64_t max_M, int64_t max_N);

}  // namespace awq

#define AWQ_CHECK(cond, ...)                         \
  do {                                               \
    if (!(cond)) return awq::set_error(1, __VA_ARGS__); \
  } while (0)

#define AWQ_CUDA_CHECK(expr, what)                                                             \
  do {                                                                                         \
    cudaError_t e__ = (expr);                                                                  \
    if (e__ != cudaSuccess) return awq::set_error(2, "%s: %s", what, cudaGetErrorString(e__)); \
  } while (0)

// after a <<<>>> launch: report launch-configuration errors without synchronising
#define AWQ_CUDA_LAUNCH_CHECK(what) AWQ_CUDA_CHECK(cudaPeekAtLastError(), what)
