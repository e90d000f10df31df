// api.cu -- library-level entry points of the C ABI (include/awq_b200.h): errors, device
// check, routing.  The kernels live next to their own extern "C" launchers.
#include <atomic>
#include <cstring>

#include "host_common.h"

namespace awq {

static thread_local char g_err[512] = "";
static std::atomic<int> g_pdl{1};

int set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}
void clear_error() { g_err[0] = 0; }

bool pdl_enabled() { return g_pdl.load(std::memory_order_relaxed) != 0; }

static std::atomic<long long> g_l2_prefetch_cap{0};
int64_t l2_prefetch_bytes(int64_t total_bytes) {
  const int64_t cap = g_l2_prefetch_cap.load(std::memory_order_relaxed);
  return total_bytes < cap ? total_bytes : cap;
}

int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

}  // namespace awq

extern "C" int awq_b200_abi_version(void) { return AWQ_B200_ABI_VERSION; }

extern "C" const char* awq_b200_last_error(void) { return awq::g_err; }

extern "C" void awq_b200_set_pdl(int enabled) { awq::g_pdl.store(enabled ? 1 : 0); }

extern "C" void awq_b200_set_l2_prefetch_mb(int mb) { awq::g_l2_prefetch_cap.store((long long)(mb < 0 ? 0 : mb) << 20); }

extern "C" int awq_b200_check_device(void) {
  static int ok[64] = {0};  // 0 unknown, 1 ok, -1 bad
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess)
    return awq::set_error(3, "awq_b200: no CUDA device (%s); this library has no CPU fallback", cudaGetErrorString(e));
  if (dev >= 0 && dev < 64 && ok[dev] == 1) return 0;
  int major = 0, minor = 0;
  cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
  cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev);
  if (major != 10)
    return awq::set_error(3, "awq_b200: device %d is sm_%d%d; the kernels are built for sm_100a only", dev, major, minor);
  if (dev >= 0 && dev < 64) ok[dev] = 1;
  return 0;
}

extern "C" int64_t awq_b200_num_tiles(int64_t K, int64_t N) {
  if (K < 0 || N < 0 || K % AWQ_B200_TILE_K != 0) return -1;
  return ((N + AWQ_B200_TILE_N - 1) / AWQ_B200_TILE_N) * (K / AWQ_B200_TILE_K);
}
