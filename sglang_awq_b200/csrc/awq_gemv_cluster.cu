// awq_gemv_cluster.cu -- kernel 2: decode path  y[M,N] = x[M,K] @ W (+ bias),  1 <= M <= 16.
//
// Replaces, for small M, the three launches of the reference's AWQLinearMethod.apply
// (python/sglang/srt/layers/quantization/awq.py:434-451: awq_dequantize -> torch.matmul ->
// add_), which move 4.5 B per weight through HBM; this kernel moves the algorithmic 0.53 B.
//
// What bounds it on B200 (measured, profiles/ + scripts/micro/stream_bw.cu):
//   (1) HBM wants a SWEEP: CTAs that are resident together must read neighbouring memory.
//       ~2400 independent far-apart streams reach 4.2 TB/s, the same loads issued so that
//       co-resident CTAs touch adjacent KBs reach 6.3-7.0 TB/s;
//   (2) the instruction stream: at 22.5 B/clk/SM a 16-row x 128-k slice (1 KB) may cost at most
//       ~180 issue slots per SM sub-partition, everything included.
//
// Structure
//   * work unit = (n-tile of 128 rows, 1/S of K); the S units of an n-tile form a thread-block
//     CLUSTER (S = 1, 2, 4, 8).  blockIdx order == memory order of the tile records, so the
//     resident CTAs sweep the weight array front to back.
//   * the CTA's 8 warps take the 8 16-row slabs of the n-tile; a warp pulls its slab's bytes
//     HBM -> registers directly (two coalesced 512-byte LDG.128 per group, kDepth groups in
//     flight, first requests issued BEFORE griddepcontrol.wait -- weights never depend on the
//     previous kernel).  No shared-memory ring, no barrier in the stream loop.
//   * the CTA's x chunk [M, K/S] is staged ONCE in shared memory (conflict-free fragment
//     layout) and shared by all 8 slabs: x is re-read N/128 times, not N/16 times.
//   * fp16: nibbles become tensor-core operands with ONE logic op per register: a nibble that
//     sits in the low mantissa bits of an fp16 SUBNORMAL is the exact value w * 2^-24 (or
//     16 w * 2^-24 one nibble up); mma.sync consumes subnormals exactly (checked on B200,
//     scripts/micro/mma_subnormal.cu).  The zero point leaves the inner loop:
//         sum_k (w - z) x = sum_k w x  -  z * Sx,     Sx = sum of x over the group,
//     and Sx is produced while staging x by one extra MMA against a matrix of ones.
//     5 ALU ops per 8 weights instead of 9 (and 22 in the reference's dequant).
//     bf16 keeps the exact (w - z) * s per-weight path (reference rounding).
//   * split-K closes inside the cluster through DISTRIBUTED SHARED MEMORY: every warp pushes
//     its fp32 partial into the shared memory of the CTA that owns the slab, one cluster
//     barrier, the owner sums the S partials in rank order -> deterministic, no workspace, no
//     global semaphores, no atomics.
#include "common.cuh"
#include "gemv_common.cuh"
#include "host_common.h"

namespace awq {

constexpr int kCluWarps = 8;
constexpr int kCluThreads = 32 * kCluWarps;

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_arrive_relaxed() { asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_arrive_release() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait_acquire() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
__device__ __forceinline__ uint32_t map_to_cta(uint32_t smem_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void st_cluster_f4(uint32_t addr, float a, float b, float c, float d) {
  asm volatile("st.shared::cluster.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
#ifdef AWQ_TRACE
__device__ unsigned long long* g_trace = nullptr;
__device__ __forceinline__ unsigned long long gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#define TRACE(k)                                                                  \
  do {                                                                            \
    if (g_trace != nullptr && lane == 0 && (warp == 0 || warp == 7))              \
      g_trace[(size_t(blockIdx.x) * 2 + (warp == 7)) * 16 + (k)] = gtime();       \
  } while (0)
#else
#define TRACE(k) do {} while (0)
#endif

template <typename T>
struct CluRaw {
  uint4 wa, wb;     // granule of row g / row g+8: 4 words = 32 consecutive k each
  uint32_t s2, z2;  // {scale g, scale g+8} as 2 x 16 bit; {zero g, zero g+8} as 2 bytes
};

// smem: [x blocks: ngmax * M * 320 B][nsx: ngmax * NT8*8 floats][incoming partials: 8 * NT8 * 512 B]
__host__ __device__ constexpr int clu_sx_off(int ngmax, int M) { return (ngmax * M * kXBlock + 127) & ~127; }
__host__ __device__ constexpr int clu_part_off(int ngmax, int M, int nt8) {
  return clu_sx_off(ngmax, M) + ((ngmax * nt8 * 32 + 127) & ~127);
}
__host__ __device__ constexpr int clu_smem_bytes(int ngmax, int M, int nt8) {
  return clu_part_off(ngmax, M, nt8) + kCluWarps * nt8 * 512;
}

// M is a template parameter: every shared-memory offset of the stream loop becomes an
// immediate and the staging loops resolve at compile time (the loop is issue-bound).
template <typename T, int M, int kDepth, int kMinB, bool kPush>
__global__ void __launch_bounds__(kCluThreads, kMinB)
gemv_cluster_kernel(const T* __restrict__ x, const uint8_t* __restrict__ tiles, const T* __restrict__ bias,
                    T* __restrict__ y, int K, int N, int KT, int lgS, int ngmax, long long pf_bytes, int pf_ctas,
                    const ArPush push) {
  constexpr bool kFast = (Elem<T>::kDtype == 0);  // fp16: subnormal operands + Sx correction
  constexpr int NT8 = (M + 7) / 8;
  extern __shared__ __align__(128) uint8_t smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, c = lane & 3;
  if (pf_bytes > 0 && lane == 0 && int(blockIdx.x) < pf_ctas)
    l2_prefetch_sweep(tiles, pf_bytes, int(blockIdx.x), pf_ctas, warp, kCluWarps);
  const int rank = int(cluster_ctarank());
  const int nt = int(blockIdx.x) >> lgS;
  const int kb = (rank * KT) >> lgS, ke = ((rank + 1) * KT) >> lgS;
  const int ng = ke - kb;

  // weights of group q: pw + q * kTileBytes (+512 for row g+8); scales pw + dm; zeros pw + dz
  const uint8_t* pw = tiles + (int64_t(nt) * KT + kb) * kTileBytes + (warp * 1024 + g * 64 + 16 * (c ^ ((g >> 1) & 3)));
  const uint8_t* pm = tiles + (int64_t(nt) * KT + kb) * kTileBytes + (kTileMetaOff + warp * kTileMetaBlock + 4 * g);
  const int dz = 32 - 2 * g;

  TRACE(0);
  CluRaw<T> buf[kDepth];
#pragma unroll
  for (int s = 0; s < kDepth; ++s)
    if (s < ng) {
      buf[s].wa = ldg128_stream(pw + s * kTileBytes);
      buf[s].wb = ldg128_stream(pw + s * kTileBytes + 512);
      buf[s].s2 = ldg32_stream(pm + s * kTileBytes);
      buf[s].z2 = ldg16_stream(pm + s * kTileBytes + dz);
    }

  cluster_arrive_relaxed();  // "this CTA is running": peers wait for it before they touch our smem
  pdl_launch_dependents();
  pdl_wait();  // x and y belong to the previous kernel until here
  TRACE(1);

  // ---- stage the x chunk [M tokens][ng groups] (+ Sx) --------------------------------------
  const uint32_t xs = smem_u32(smem);
  const uint32_t sxs = xs + uint32_t(clu_sx_off(ngmax, M));
  stage_x<T, M, NT8 * 8, kFast>(x, K, kb * kTileK, ng, xs, sxs, warp, kCluWarps, lane);
  TRACE(2);
  __syncthreads();
  TRACE(3);

  // ---- stream loop ----------------------------------------------------------------------------
  uint32_t xq[NT8];
#pragma unroll
  for (int t = 0; t < NT8; ++t) {
    const int tok = 8 * t + g;
    xq[t] = xs + uint32_t((tok < M ? tok : M - 1) * kXBlock + 16 * c);
  }
  uint32_t sxq = sxs + 8u * uint32_t(c);

  float acc[NT8][4];
#pragma unroll
  for (int t = 0; t < NT8; ++t)
#pragma unroll
    for (int i = 0; i < 4; ++i) acc[t][i] = 0.f;

  for (int i = 0; i < ng; i += kDepth) {
#pragma unroll
    for (int s = 0; s < kDepth; ++s) {
      if (i + s < ng) {
        uint4 xf[NT8][4];
#pragma unroll
        for (int t = 0; t < NT8; ++t)
#pragma unroll
          for (int sp = 0; sp < 4; ++sp) xf[t][sp] = lds128(xq[t] + uint32_t(s * M * kXBlock + 64 * sp));
        if constexpr (kFast) {
          float2 nsx[NT8];
#pragma unroll
          for (int t = 0; t < NT8; ++t) {
            const uint2 v = lds64(sxq + uint32_t(s * NT8 * 32 + 32 * t));
            nsx[t] = make_float2(__uint_as_float(v.x), __uint_as_float(v.y));
          }
          group_mma_subnormal<NT8>(buf[s].wa, buf[s].wb, buf[s].s2, buf[s].z2, xf, nsx, acc);
        } else {
          gemv_group_mma<T, NT8>(buf[s].wa, buf[s].wb, buf[s].s2, buf[s].z2, xf, acc);
        }
        if (i == 0 && s == 0) TRACE(4);
        if (i + s + kDepth < ng) {
          buf[s].wa = ldg128_stream(pw + (s + kDepth) * kTileBytes);
          buf[s].wb = ldg128_stream(pw + (s + kDepth) * kTileBytes + 512);
          buf[s].s2 = ldg32_stream(pm + (s + kDepth) * kTileBytes);
          buf[s].z2 = ldg16_stream(pm + (s + kDepth) * kTileBytes + dz);
        }
      }
    }
    pw += kDepth * kTileBytes;
    pm += kDepth * kTileBytes;
#pragma unroll
    for (int t = 0; t < NT8; ++t) xq[t] += kDepth * M * kXBlock;
    sxq += kDepth * NT8 * 32;
  }

  // ---- close split-K inside the cluster --------------------------------------------------------
  // slab `warp` is owned by CTA (warp % S); its partial lands in that CTA's smem at
  // [source rank][warp / S][token block][lane] (float4 each)
  const uint32_t part = xs + uint32_t(clu_part_off(ngmax, M, NT8));
  const int S = 1 << lgS, own_slabs = kCluWarps >> lgS;  // this CTA owns slabs rank, rank + S, ...
  TRACE(5);
  cluster_wait_acquire();                                // every CTA of the cluster is running
  TRACE(6);
  {
    const uint32_t dst = map_to_cta(part, uint32_t(warp & (S - 1))) +
                         uint32_t(((rank * own_slabs + (warp >> lgS)) * NT8) * 512 + 16 * lane);
#pragma unroll
    for (int t = 0; t < NT8; ++t) st_cluster_f4(dst + 512u * t, acc[t][0], acc[t][1], acc[t][2], acc[t][3]);
  }
  cluster_arrive_release();
  cluster_wait_acquire();
  TRACE(7);

  const float* pf = reinterpret_cast<const float*>(smem + clu_part_off(ngmax, M, NT8));
  const int own_elems = own_slabs * NT8 * 128;
  for (int e = tid; e < own_elems; e += kCluThreads) {
    float sum = 0.f;
    for (int r = 0; r < S; ++r) sum += pf[r * own_elems + e];  // rank order: deterministic
    const int sl = e / (NT8 * 128), rem = e % (NT8 * 128), t = rem >> 7, idx = rem & 127;
    // element i of lane L: i = 0: (row g, tok 2c)  1: (row g, tok 2c+1)  2: (row g+8, tok 2c)  3: (row g+8, tok 2c+1)
    const int L = idx >> 2, i = idx & 3;
    const int row = (L >> 2) + 8 * (i >> 1), tok = 8 * t + 2 * (L & 3) + (i & 1);
    const int n = nt * kTileN + (rank + sl * S) * 16 + row;
    const bool valid = tok < M && n < N;
    if (valid) {
      if (bias != nullptr) sum += Elem<T>::to_float(bias[n]);
      y[int64_t(tok) * N + n] = Elem<T>::from_float(sum);
    }
    // row-parallel layers (kPush instantiations only: the plain kernel pays nothing): the all-reduce's packets leave from here
    if constexpr (kPush) ar_push_pair<T>(push, tok, n, N, valid, sum);
  }
  TRACE(8);
#ifdef AWQ_TRACE
  if (g_trace != nullptr && lane == 0 && warp == 0) {
    unsigned smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    g_trace[size_t(blockIdx.x) * 32 + 15] = smid;
  }
#endif
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
static int g_clu_split = 0;    // 0 = heuristic, else forced cluster size (1, 2, 4, 8)
static int g_clu_variant = -1;  // -1 = pick from the shape, else forced instantiation (debug)
void gemv_cluster_debug(int split, int variant) {
  g_clu_split = split;
  g_clu_variant = variant;
}

static int clu_pick_lgS(int64_t KT, int64_t NT) {
  int s = 8;
  if (g_clu_split == 1 || g_clu_split == 2 || g_clu_split == 4 || g_clu_split == 8) {
    s = g_clu_split;
  } else {
    // as many CTAs as it takes to keep every SM streaming, but >= 4 groups per warp when K allows
    const int64_t sms = sm_count();
    while (s > 1 && (KT / s < 4 || (KT / s < 8 && NT * (s / 2) >= 6 * sms))) s >>= 1;
  }
  while (s > 1 && KT < s) s >>= 1;
  return s == 8 ? 3 : (s == 4 ? 2 : (s == 2 ? 1 : 0));
}

template <typename T, int M, int kDepth, int kMinB, bool kPush>
static int gemv_cluster_launch_p(const void* x, const void* tiles, const void* bias, void* y, int64_t K, int64_t N,
                               cudaStream_t st, const ArPush& push) {
  constexpr int NT8 = (M + 7) / 8;
  const int KT = int(K / kTileK), NT = int((N + kTileN - 1) / kTileN);
  const int lgS = clu_pick_lgS(KT, NT), S = 1 << lgS;
  const int ngmax = (KT + S - 1) / S;
  const int smem = clu_smem_bytes(ngmax, M, NT8);
  AWQ_CHECK(smem <= 227 * 1024, "awq_b200_gemv: K * M too large for the decode kernel (%d B of shared memory)", smem);
  static int configured[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (smem > configured[dev]) {
    AWQ_CUDA_CHECK(cudaFuncSetAttribute(gemv_cluster_kernel<T, M, kDepth, kMinB, kPush>,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, smem),
                   "awq_b200_gemv: smem opt-in");
    configured[dev] = smem;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(unsigned(NT * S));
  cfg.blockDim = dim3(kCluThreads);
  cfg.dynamicSmemBytes = size_t(smem);
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = unsigned(S);
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  AWQ_CUDA_CHECK(cudaLaunchKernelEx(&cfg, gemv_cluster_kernel<T, M, kDepth, kMinB, kPush>, (const T*)x, (const uint8_t*)tiles,
                                    (const T*)bias, (T*)y, int(K), int(N), KT, lgS, ngmax,
                                    (long long)l2_prefetch_bytes(int64_t(NT) * KT * kTileBytes),
                                    NT * S < sm_count() ? NT * S : sm_count(), push),
                 "awq_b200_gemv: launch");
  return 0;
}

template <typename T, int M, int kDepth, int kMinB>
static int gemv_cluster_launch(const void* x, const void* tiles, const void* bias, void* y, int64_t K, int64_t N,
                               cudaStream_t st, const ArPush& push) {
  if (push.peers != nullptr) return gemv_cluster_launch_p<T, M, kDepth, kMinB, true>(x, tiles, bias, y, K, N, st, push);
  return gemv_cluster_launch_p<T, M, kDepth, kMinB, false>(x, tiles, bias, y, K, N, st, push);
}

#ifdef AWQ_TRACE
extern "C" void awq_b200_debug_set_trace(void* buf) {
  unsigned long long* p = (unsigned long long*)buf;
  cudaMemcpyToSymbol(g_trace, &p, sizeof(p));
}
#endif

bool gemv_cluster_fits(int64_t M, int64_t K) {
  if (M < 1 || M > 16) return false;
  const int64_t KT = K / kTileK;
  const int64_t ngmax = (KT + 7) / 8;  // at the largest split
  return clu_smem_bytes(int(ngmax), int(M), M <= 8 ? 1 : 2) <= 100 * 1024;
}

template <typename T, int M>
static int gemv_cluster_m(const void* x, const void* tiles, const void* bias, void* y, int64_t K, int64_t N,
                          cudaStream_t st, const ArPush& push) {
  if constexpr (Elem<T>::kDtype == 1) {
    return gemv_cluster_launch<T, M, (M <= 8 ? 4 : 3), (M <= 8 ? 3 : 2)>(x, tiles, bias, y, K, N, st, push);
  } else {
    int variant = g_clu_variant;
    if (variant < 0) {
      // everything resident in one wave -> fewest CTAs per SM, all registers; else more CTAs per SM
      const int64_t KT = K / kTileK, NT = (N + kTileN - 1) / kTileN;
      const int64_t ctas = NT << clu_pick_lgS(KT, NT);
      variant = ctas <= 2 * sm_count() ? 1 : (M <= 8 ? 0 : 3);
    }
    switch (variant) {
      case 1: return gemv_cluster_launch<T, M, 4, 2>(x, tiles, bias, y, K, N, st, push);
      case 3: return gemv_cluster_launch<T, M, (M <= 8 ? 2 : 3), (M <= 8 ? 4 : 3)>(x, tiles, bias, y, K, N, st, push);
      default: return gemv_cluster_launch<T, M, 4, 3>(x, tiles, bias, y, K, N, st, push);
    }
  }
}

template <typename T>
static int gemv_cluster_t(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K,
                          int64_t N, cudaStream_t st, const ArPush& push) {
  switch (int(M)) {
#define AWQ_CLU_CASE(MM) \
  case MM: return gemv_cluster_m<T, MM>(x, tiles, bias, y, K, N, st, push)
    AWQ_CLU_CASE(1); AWQ_CLU_CASE(2); AWQ_CLU_CASE(3); AWQ_CLU_CASE(4);
    AWQ_CLU_CASE(5); AWQ_CLU_CASE(6); AWQ_CLU_CASE(7); AWQ_CLU_CASE(8);
    AWQ_CLU_CASE(9); AWQ_CLU_CASE(10); AWQ_CLU_CASE(11); AWQ_CLU_CASE(12);
    AWQ_CLU_CASE(13); AWQ_CLU_CASE(14); AWQ_CLU_CASE(15); AWQ_CLU_CASE(16);
#undef AWQ_CLU_CASE
    default: return set_error(1, "awq_b200_gemv: cluster kernel supports 1 <= M <= 16");
  }
}

int gemv_cluster(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K, int64_t N,
                 int dtype, cudaStream_t st, const ArPush* push_or_null) {
  const ArPush none = {nullptr, 0, 0, 0, 0, 0, 0};
  const ArPush& push = push_or_null != nullptr ? *push_or_null : none;
  if (dtype == AWQ_B200_F16) return gemv_cluster_t<__half>(x, tiles, bias, y, M, K, N, st, push);
  return gemv_cluster_t<__nv_bfloat16>(x, tiles, bias, y, M, K, N, st, push);
}

}  // namespace awq
