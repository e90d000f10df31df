// awq_gemv_direct.cu -- kernel 2a: decode path for very small batches (M <= kDirectMaxM).
//
//   y[M,N] = x[M,K] @ W (+ bias)
//
// Replaces, for M = 1..4, the three launches of the reference's AWQLinearMethod.apply
// (python/sglang/srt/layers/quantization/awq.py:434-451).  HBM-bound: what matters is that
// every SM has enough weight bytes in flight from the first cycle of the launch to the last,
// that co-resident CTAs read neighbouring memory (HBM wants a sweep, see awq_gemv_cluster.cu),
// and that the instruction stream per byte stays under the issue budget (22.5 B/clk/SM at
// 6.5 TB/s => ~180 issue slots per SM sub-partition per KB of weights).
//
//   * one CTA = one 16-row slab of the output over the WHOLE K: no cross-CTA reduction, no
//     workspace, no semaphores, deterministic; the finest work granule the MMA shape allows
//     (32 KB of weights at K = 4096), so the hardware CTA scheduler balances the SMs.  The
//     CTA's 8 warps split K in contiguous group ranges and are reduced through 4 KB of shared
//     memory in warp order;
//   * weights go HBM -> registers directly (LDG.128, no shared-memory ring, no cp.async, no
//     barrier in the stream loop).  The tile records were laid out so that the
//     lane (g, c) of a warp needs exactly one 16-byte granule of row g and one of row g+8 per
//     128-k group: a warp-wide load is 512 contiguous bytes;
//   * kDepth groups (kDepth KB per warp) are requested before anything else, BEFORE
//     griddepcontrol.wait: weights never depend on the previous kernel, so under programmatic
//     dependent launch the next layer's stream starts while this layer drains;
//   * every warp stages ITS k-range of x once in a private shared-memory region (conflict-free
//     fragment layout) -- an L2 round trip inside the stream loop would sit behind outstanding
//     DRAM requests in the SM's return path;
//   * fp16: subnormal tensor-core operands + group-sum correction (gemv_common.cuh), 5 ALU ops
//     per 8 weights; bf16: exact per-weight (w - z) * s path.
// x traffic is K*M*2 bytes per 16 rows, which is why this kernel is only used for M <= 4;
// larger M share x between the 8 slabs of an n-tile (awq_gemv_cluster.cu).
#include "common.cuh"
#include "gemv_common.cuh"
#include "host_common.h"

namespace awq {

constexpr int kDirWarps = 8;
constexpr int kDirThreads = 32 * kDirWarps;

// smem: [warp][ngmax * M * 320 B x blocks][ngmax * 8 floats of group sums] | [8][128] floats of partials
__host__ __device__ constexpr int dir_warp_bytes(int ngmax, int M) { return ((ngmax * M * kXBlock + ngmax * 32) + 127) & ~127; }
__host__ __device__ constexpr int dir_smem_bytes(int ngmax, int M) { return kDirWarps * dir_warp_bytes(ngmax, M) + kDirWarps * 512; }

template <typename T, int M, int kDepth, int kMinB, bool kPush>
__global__ void __launch_bounds__(kDirThreads, kMinB)
gemv_direct_kernel(const T* __restrict__ x, const uint8_t* __restrict__ tiles, const T* __restrict__ bias,
                   T* __restrict__ y, int K, int N, int KT, int ngmax, long long pf_bytes, int pf_ctas, const ArPush push) {
  constexpr bool kFast = (Elem<T>::kDtype == 0);
  extern __shared__ __align__(128) uint8_t smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, c = lane & 3;
  if (pf_bytes > 0 && lane == 0 && int(blockIdx.x) < pf_ctas)
    l2_prefetch_sweep(tiles, pf_bytes, int(blockIdx.x), pf_ctas, warp, kDirWarps);
  const int slab = blockIdx.x, nt = slab >> 3, sl = slab & 7;
  // this warp's groups [kb, ke)
  const int kb = (warp * KT) >> 3, ke = ((warp + 1) * KT) >> 3;
  const int ng = ke - kb;

  const uint8_t* pw = tiles + (int64_t(nt) * KT + kb) * kTileBytes + (sl * 1024 + g * 64 + 16 * (c ^ ((g >> 1) & 3)));
  const uint8_t* pm = tiles + (int64_t(nt) * KT + kb) * kTileBytes + (kTileMetaOff + sl * kTileMetaBlock + 4 * g);
  const int dz = 32 - 2 * g;

  uint4 wa[kDepth], wb[kDepth];
  uint32_t s2[kDepth], z2[kDepth];
#pragma unroll
  for (int s = 0; s < kDepth; ++s)
    if (s < ng) {
      wa[s] = ldg128_stream(pw + s * kTileBytes);
      wb[s] = ldg128_stream(pw + s * kTileBytes + 512);
      s2[s] = ldg32_stream(pm + s * kTileBytes);
      z2[s] = ldg16_stream(pm + s * kTileBytes + dz);
    }

  pdl_launch_dependents();
  pdl_wait();  // x and y belong to the previous kernel until here

  const uint32_t xs = smem_u32(smem) + uint32_t(warp * dir_warp_bytes(ngmax, M));
  const uint32_t sxs = xs + uint32_t(ngmax * M * kXBlock);
  stage_x<T, M, 8, kFast>(x, K, kb * kTileK, ng, xs, sxs, 0, 1, lane);
  __syncwarp();

  // lane (g, c) feeds token g (clamped: columns >= M are never stored)
  uint32_t xq = xs + uint32_t((g < M ? g : M - 1) * kXBlock + 16 * c);
  uint32_t sxq = sxs + 8u * uint32_t(c);

  float acc[1][4] = {{0.f, 0.f, 0.f, 0.f}};
  for (int i = 0; i < ng; i += kDepth) {
#pragma unroll
    for (int s = 0; s < kDepth; ++s) {
      if (i + s < ng) {
        uint4 xf[1][4];
#pragma unroll
        for (int sp = 0; sp < 4; ++sp) xf[0][sp] = lds128(xq + uint32_t(s * M * kXBlock + 64 * sp));
        if constexpr (kFast) {
          const uint2 v = lds64(sxq + uint32_t(s * 32));
          const float2 nsx[1] = {make_float2(__uint_as_float(v.x), __uint_as_float(v.y))};
          group_mma_subnormal<1>(wa[s], wb[s], s2[s], z2[s], xf, nsx, acc);
        } else {
          gemv_group_mma<T, 1>(wa[s], wb[s], s2[s], z2[s], xf, acc);
        }
        if (i + s + kDepth < ng) {
          wa[s] = ldg128_stream(pw + (s + kDepth) * kTileBytes);
          wb[s] = ldg128_stream(pw + (s + kDepth) * kTileBytes + 512);
          s2[s] = ldg32_stream(pm + (s + kDepth) * kTileBytes);
          z2[s] = ldg16_stream(pm + (s + kDepth) * kTileBytes + dz);
        }
      }
    }
    pw += kDepth * kTileBytes;
    pm += kDepth * kTileBytes;
    xq += kDepth * M * kXBlock;
    sxq += kDepth * 32;
  }

  // cross-warp reduction in warp order (deterministic)
  float* red = reinterpret_cast<float*>(smem + kDirWarps * dir_warp_bytes(ngmax, M));
  *reinterpret_cast<float4*>(&red[warp * 128 + 4 * lane]) = make_float4(acc[0][0], acc[0][1], acc[0][2], acc[0][3]);
  __syncthreads();
  if (tid < 128) {
    // element i of lane L: i = 0: (row g, tok 2c)  1: (row g, tok 2c+1)  2: (row g+8, tok 2c)  3: (row g+8, tok 2c+1)
    const int L = tid >> 2, i = tid & 3;
    const int row = (L >> 2) + 8 * (i >> 1), t = 2 * (L & 3) + (i & 1);
    const int n = slab * 16 + row;
    const bool valid = t < M && n < N;
    float s = 0.f;
    if (valid) {
#pragma unroll
      for (int w = 0; w < kDirWarps; ++w) s += red[w * 128 + tid];
      if (bias != nullptr) s += Elem<T>::to_float(bias[n]);
      y[int64_t(t) * N + n] = Elem<T>::from_float(s);
    }
    // row-parallel layers (kPush instantiations only: the plain kernel pays nothing): the all-reduce's packets leave from here
    if constexpr (kPush) ar_push_pair<T>(push, t, n, N, valid, s);
  }
}

template <typename T, int M, int kDepth, int kMinB, bool kPush>
static int gemv_direct_launch_p(const void* x, const void* tiles, const void* bias, void* y, int64_t K, int64_t N,
                              cudaStream_t st, const ArPush& push) {
  const int KT = int(K / kTileK);
  const int ngmax = (KT + kDirWarps - 1) / kDirWarps;
  const int smem = dir_smem_bytes(ngmax, M);
  static int configured[64] = {0};  // largest dynamic smem size opted into so far, per device (per instantiation)
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (smem > configured[dev]) {
    AWQ_CHECK(smem <= 227 * 1024, "awq_b200_gemv: K * M too large for the direct decode kernel");
    AWQ_CUDA_CHECK(cudaFuncSetAttribute(gemv_direct_kernel<T, M, kDepth, kMinB, kPush>,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, smem),
                   "awq_b200_gemv: smem opt-in (direct)");
    configured[dev] = smem;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(unsigned((N + 15) / 16));
  cfg.blockDim = dim3(kDirThreads);
  cfg.dynamicSmemBytes = size_t(smem);
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  AWQ_CUDA_CHECK(cudaLaunchKernelEx(&cfg, gemv_direct_kernel<T, M, kDepth, kMinB, kPush>, (const T*)x, (const uint8_t*)tiles,
                                    (const T*)bias, (T*)y, int(K), int(N), KT, ngmax,
                                    (long long)l2_prefetch_bytes(int64_t((N + kTileN - 1) / kTileN) * KT * kTileBytes),
                                    int((N + 15) / 16) < sm_count() ? int((N + 15) / 16) : sm_count(), push),
                 "awq_b200_gemv: launch (direct)");
  return 0;
}

template <typename T, int M, int kDepth, int kMinB>
static int gemv_direct_launch(const void* x, const void* tiles, const void* bias, void* y, int64_t K, int64_t N,
                              cudaStream_t st, const ArPush& push) {
  if (push.peers != nullptr) return gemv_direct_launch_p<T, M, kDepth, kMinB, true>(x, tiles, bias, y, K, N, st, push);
  return gemv_direct_launch_p<T, M, kDepth, kMinB, false>(x, tiles, bias, y, K, N, st, push);
}

static int g_direct_variant = -1;  // -1 = pick from the shape, else forced instantiation (debug)
void gemv_direct_set_variant(int v) { g_direct_variant = v; }

// largest M the direct kernel is instantiated for
constexpr int kDirectMaxM = 4;

bool gemv_direct_fits(int64_t M, int64_t K) {
  if (M < 1 || M > kDirectMaxM) return false;
  const int64_t ngmax = (K / kTileK + kDirWarps - 1) / kDirWarps;
  return dir_smem_bytes(int(ngmax), int(M)) <= 72 * 1024;  // keep >= 3 CTAs per SM resident
}

template <typename T, int M>
static int gemv_direct_m(const void* x, const void* tiles, const void* bias, void* y, int64_t K, int64_t N,
                         cudaStream_t st, const ArPush& push) {
  if constexpr (Elem<T>::kDtype == 1) {
    return gemv_direct_launch<T, M, 4, 3>(x, tiles, bias, y, K, N, st, push);
  } else {
    int variant = g_direct_variant;
    if (variant < 0) {
      const int64_t ctas = (N + 15) / 16, sms = sm_count();
      variant = ctas <= 2 * sms ? 1 : (ctas <= 3 * sms ? 0 : 3);
    }
    switch (variant) {
      case 1: return gemv_direct_launch<T, M, 4, 2>(x, tiles, bias, y, K, N, st, push);
      case 2: return gemv_direct_launch<T, M, 3, 4>(x, tiles, bias, y, K, N, st, push);
      case 3: return gemv_direct_launch<T, M, 2, 4>(x, tiles, bias, y, K, N, st, push);
      default: return gemv_direct_launch<T, M, 4, 3>(x, tiles, bias, y, K, N, st, push);
    }
  }
}

int gemv_direct(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K, int64_t N,
                int dtype, cudaStream_t st, const ArPush* push_or_null) {
  const ArPush none = {nullptr, 0, 0, 0, 0, 0, 0};
  const ArPush& push = push_or_null != nullptr ? *push_or_null : none;
#define AWQ_DIR_CASE(MM)                                                                       \
  case MM:                                                                                     \
    return dtype == AWQ_B200_F16 ? gemv_direct_m<__half, MM>(x, tiles, bias, y, K, N, st, push)       \
                                 : gemv_direct_m<__nv_bfloat16, MM>(x, tiles, bias, y, K, N, st, push)
  switch (int(M)) {
    AWQ_DIR_CASE(1);
    AWQ_DIR_CASE(2);
    AWQ_DIR_CASE(3);
    AWQ_DIR_CASE(4);
    default:
      return set_error(1, "awq_b200_gemv: direct kernel supports M <= %d", kDirectMaxM);
  }
#undef AWQ_DIR_CASE
}

}  // namespace awq
