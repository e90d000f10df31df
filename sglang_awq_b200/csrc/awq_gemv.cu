// awq_gemv.cu -- kernel 2: decode path  y[M,N] = x[M,K] @ W (+ bias),  1 <= M <= 64.
//
// Replaces, for small M, the three launches of the reference's AWQLinearMethod.apply
// (python/sglang/srt/layers/quantization/awq.py:434-451: awq_dequantize -> torch.matmul ->
// add_), which move 4.5 B per weight through HBM.  This kernel moves the algorithmic 0.53 B.
//
// The op is HBM-bound, so the design is about keeping HBM saturated, not about FLOPs:
//   * weights come in the tile layout (awq_repack.cu): one pipeline stage = one contiguous
//     8 KB tile + 256 B scales + 64 B zeros, fetched by ONE elected thread with cp.async.bulk
//     (TMA 1-D, UBLKCP) into an 8-deep shared-memory ring guarded by mbarriers -> ~100 KB in
//     flight per CTA, 2 CTAs per SM, no register staging, no per-thread address arithmetic;
//   * x rows for the same 128-wide k slice ride in the same stage (M bulk copies of 256 B);
//   * programmatic dependent launch: the first ring's worth of WEIGHT copies is issued before
//     griddepcontrol.wait (weights never depend on the previous kernel; x does), so the HBM
//     stream of layer i+1 starts while layer i drains;
//   * each of 8 consumer warps owns 16 output rows: per tile it loads 2x16 B of packed
//     weights, turns them into exact (w - z) fp16/bf16 register fragments with 9 ALU ops per
//     8 weights, and reduces over k with mma.sync.m16n8k16 (fp32 accumulate).  CUDA-core FMAs
//     cannot do this at HBM rate: >= 25 issue slots per 8 weights at M = 1 against a budget of
//     ~22 (4 schedulers / (46 weights/clk/SM)); with the tensor pipe it is 10-11 for M <= 16.
//     The group scale is applied in fp32 once per (row, group) when the group's partial sum is
//     folded into the running accumulator -- more accurate than scaling each weight in fp16;
//   * split-K across CTAs (small-N layers cannot fill 148 SMs otherwise): fp32 partials go to
//     an L2-resident workspace, the last CTA to arrive (per n-tile semaphore) sums them in a
//     FIXED order -> deterministic output, no float atomics.
#include "common.cuh"
#include "host_common.h"

namespace awq {

constexpr int kGemvConsumerWarps = 8;
constexpr int kGemvThreads = 32 * (kGemvConsumerWarps + 1);
constexpr int kXPitch = 272;  // 256 B of k + 16 B: ldmatrix rows land in distinct bank groups
constexpr int kXOff = 8576;   // w 8192 | scales 256 | zeros 64 | pad 64
constexpr int kSemInts = 4096;

__host__ __device__ constexpr int gemv_stage_bytes(int nt8) { return kXOff + 8 * nt8 * kXPitch; }
__host__ __device__ constexpr int gemv_stages(int nt8) { return nt8 <= 2 ? 8 : (nt8 == 4 ? 6 : 4); }
constexpr int kOutPitch = 132;  // floats; 132 % 32 == 4 -> conflict-free fragment scatter

template <typename T>
__device__ __forceinline__ void mma_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1);
template <>
__device__ __forceinline__ void mma_16816<__half>(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
template <>
__device__ __forceinline__ void mma_16816<__nv_bfloat16>(float (&d)[4], const uint32_t (&a)[4], uint32_t b0,
                                                        uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__device__ __forceinline__ void ldmatrix_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}

template <typename T>
__device__ __forceinline__ float bits16_to_float(uint32_t b) {
  const uint16_t h = uint16_t(b);
  return Elem<T>::to_float(*reinterpret_cast<const T*>(&h));
}

template <typename T, int NT8>
__global__ void __launch_bounds__(kGemvThreads, NT8 <= 2 ? 2 : 1)
gemv_kernel(const T* __restrict__ x, const uint32_t* __restrict__ wq, const T* __restrict__ sc,
            const uint8_t* __restrict__ zq, const T* __restrict__ bias, T* __restrict__ y,
            float* __restrict__ ws_partial, int* __restrict__ ws_sem, int M, int K, int N, int KT, int splits) {
  constexpr int kStages = gemv_stages(NT8);
  constexpr int kStageBytes = gemv_stage_bytes(NT8);
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ int s_last;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int nt = blockIdx.x / splits, split = blockIdx.x - nt * splits;
  const int kt_begin = int(int64_t(KT) * split / splits);
  const int kt_end = int(int64_t(KT) * (split + 1) / splits);
  const int nk = kt_end - kt_begin;

  const uint32_t smem_base = smem_u32(smem);
  const uint32_t bar_full = smem_base + kStages * kStageBytes;
  const uint32_t bar_empty = bar_full + 8 * kStages;

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < kStages; ++s) {
      mbar_init(bar_full + 8 * s, 1);
      mbar_init(bar_empty + 8 * s, kGemvConsumerWarps);
    }
    mbar_fence_init();
  }
  __syncthreads();
  pdl_launch_dependents();

  if (warp == kGemvConsumerWarps) {
    // ------------------------------ producer: one elected thread ------------------------------
    if (lane == 0) {
      const int64_t tile0 = int64_t(nt) * KT + kt_begin;
      const uint32_t tx = kTileWBytes + kTileSBytes + kTileZBytes + uint32_t(M) * 256u;
      const uint64_t pol_w = policy_evict_first();  // weights are streamed once per call
      auto issue_w = [&](int it, int s) {
        const uint32_t st = smem_base + s * kStageBytes, bar = bar_full + 8 * s;
        const int64_t t = tile0 + it;
        bulk_g2s_hint(st, wq + t * kTileWWords, kTileWBytes, bar, pol_w);
        bulk_g2s_hint(st + kTileWBytes, sc + t * kTileN, kTileSBytes, bar, pol_w);
        bulk_g2s_hint(st + kTileWBytes + kTileSBytes, zq + t * kTileZBytes, kTileZBytes, bar, pol_w);
      };
      auto issue_x = [&](int it, int s) {
        const uint32_t st = smem_base + s * kStageBytes + kXOff, bar = bar_full + 8 * s;
        const T* src = x + int64_t(kt_begin + it) * kTileK;
        for (int m = 0; m < M; ++m) bulk_g2s(st + m * kXPitch, src + int64_t(m) * K, 256, bar);
      };
      const int pre = nk < kStages ? nk : kStages;
      for (int it = 0; it < pre; ++it) {
        mbar_arrive_expect_tx(bar_full + 8 * it, tx);
        issue_w(it, it);
      }
      pdl_wait();  // x (and the workspace) belong to the previous kernel until here
      for (int it = 0; it < pre; ++it) issue_x(it, it);
      for (int it = pre; it < nk; ++it) {
        const int s = it % kStages;
        mbar_wait(bar_empty + 8 * s, ((it / kStages) & 1) ^ 1);
        mbar_arrive_expect_tx(bar_full + 8 * s, tx);
        issue_w(it, s);
        issue_x(it, s);
      }
    }
    return;
  }

  // ---------------------------------- consumers: 8 warps x 16 rows ----------------------------------
  const int g = lane >> 2, c = lane & 3;
  const uint32_t w_off_a = uint32_t(16 * warp + g) * 64u + 16u * uint32_t(c ^ ((g >> 1) & 3));
  const uint32_t w_off_b = w_off_a + 8u * 64u;
  const uint32_t s_off = kTileWBytes + uint32_t(16 * warp + 2 * g) * 2u;
  const uint32_t z_off = kTileWBytes + kTileSBytes + uint32_t(8 * warp + g);
  const uint32_t x_off = kXOff + uint32_t(lane & 7) * kXPitch + 16u * uint32_t(lane >> 3);

  float acc[NT8][4];
#pragma unroll
  for (int t = 0; t < NT8; ++t)
#pragma unroll
    for (int i = 0; i < 4; ++i) acc[t][i] = 0.f;

  for (int it = 0; it < nk; ++it) {
    const int s = it % kStages;
    mbar_wait(bar_full + 8 * s, (it / kStages) & 1);
    const uint32_t st = smem_base + s * kStageBytes;

    const uint4 wa4 = lds128(st + w_off_a);
    const uint4 wb4 = lds128(st + w_off_b);
    const uint32_t s2 = lds32(st + s_off);
    const uint32_t zb = lds8(st + z_off);
    const uint32_t wa[4] = {wa4.x, wa4.y, wa4.z, wa4.w};
    const uint32_t wb[4] = {wb4.x, wb4.y, wb4.z, wb4.w};
    uint32_t za_e, za_o, zb_e, zb_o;
    Unpack<T>::zero_consts_scalar(zb & 0xFu, za_e, za_o);
    Unpack<T>::zero_consts_scalar(zb >> 4, zb_e, zb_o);

    float gacc[NT8][4];
#pragma unroll
    for (int t = 0; t < NT8; ++t)
#pragma unroll
      for (int i = 0; i < 4; ++i) gacc[t][i] = 0.f;

#pragma unroll
    for (int sp = 0; sp < 4; ++sp) {
      uint32_t ra[4], rb[4];
      Unpack<T>::raw(wa[sp], ra);
      Unpack<T>::raw(wb[sp], rb);
      uint32_t a0[4], a1[4];  // A fragments of k-steps 2sp and 2sp+1
      a0[0] = Unpack<T>::template diff<0>(ra[0], za_e);
      a0[1] = Unpack<T>::template diff<0>(rb[0], zb_e);
      a0[2] = Unpack<T>::template diff<1>(ra[1], za_o);
      a0[3] = Unpack<T>::template diff<1>(rb[1], zb_o);
      a1[0] = Unpack<T>::template diff<2>(ra[2], za_e);
      a1[1] = Unpack<T>::template diff<2>(rb[2], zb_e);
      a1[2] = Unpack<T>::template diff<3>(ra[3], za_o);
      a1[3] = Unpack<T>::template diff<3>(rb[3], zb_o);
#pragma unroll
      for (int t = 0; t < NT8; ++t) {
        uint32_t b0, b1, b2, b3;
        ldmatrix_x4(st + x_off + uint32_t(t) * 8u * kXPitch + 64u * uint32_t(sp), b0, b1, b2, b3);
        mma_16816<T>(gacc[t], a0, b0, b1);
        mma_16816<T>(gacc[t], a1, b2, b3);
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(bar_empty + 8 * s);

    const float sa = bits16_to_float<T>(s2 & 0xFFFFu), sb = bits16_to_float<T>(s2 >> 16);
#pragma unroll
    for (int t = 0; t < NT8; ++t) {
      acc[t][0] = fmaf(sa, gacc[t][0], acc[t][0]);
      acc[t][1] = fmaf(sa, gacc[t][1], acc[t][1]);
      acc[t][2] = fmaf(sb, gacc[t][2], acc[t][2]);
      acc[t][3] = fmaf(sb, gacc[t][3], acc[t][3]);
    }
  }

  // ------------------------------------------ epilogue ------------------------------------------
  pdl_wait();  // y / workspace may still be in use by the previous kernel
  named_bar_sync(1, 32 * kGemvConsumerWarps);  // every warp is done reading the ring
  float* so = reinterpret_cast<float*>(smem);  // [8*NT8][kOutPitch]
#pragma unroll
  for (int t = 0; t < NT8; ++t) {
    const int tok = 8 * t + 2 * c, row = 16 * warp + g;
    so[tok * kOutPitch + row] = acc[t][0];
    so[(tok + 1) * kOutPitch + row] = acc[t][1];
    so[tok * kOutPitch + row + 8] = acc[t][2];
    so[(tok + 1) * kOutPitch + row + 8] = acc[t][3];
  }
  named_bar_sync(1, 32 * kGemvConsumerWarps);

  const int n0 = nt * kTileN;
  const int NPAD = ((N + kTileN - 1) / kTileN) * kTileN;
  auto store_y = [&](int m, int j, float4 v) {
    const int n = n0 + 4 * j;
    if (n >= N) return;  // N % 8 == 0 -> groups of 4 are all-in or all-out
    if (bias != nullptr) {
      v.x += Elem<T>::to_float(bias[n]);
      v.y += Elem<T>::to_float(bias[n + 1]);
      v.z += Elem<T>::to_float(bias[n + 2]);
      v.w += Elem<T>::to_float(bias[n + 3]);
    }
    const T o0 = Elem<T>::from_float(v.x), o1 = Elem<T>::from_float(v.y);
    const T o2 = Elem<T>::from_float(v.z), o3 = Elem<T>::from_float(v.w);
    uint2 pk;
    pk.x = uint32_t(*reinterpret_cast<const uint16_t*>(&o0)) | (uint32_t(*reinterpret_cast<const uint16_t*>(&o1)) << 16);
    pk.y = uint32_t(*reinterpret_cast<const uint16_t*>(&o2)) | (uint32_t(*reinterpret_cast<const uint16_t*>(&o3)) << 16);
    *reinterpret_cast<uint2*>(y + int64_t(m) * N + n) = pk;
  };

  if (splits == 1) {
    for (int idx = tid; idx < M * 32; idx += 32 * kGemvConsumerWarps) {
      const int m = idx >> 5, j = idx & 31;
      store_y(m, j, *reinterpret_cast<const float4*>(so + m * kOutPitch + 4 * j));
    }
    return;
  }

  // split-K: publish the partial, last arriver reduces in split order
  for (int idx = tid; idx < M * 32; idx += 32 * kGemvConsumerWarps) {
    const int m = idx >> 5, j = idx & 31;
    const float4 v = *reinterpret_cast<const float4*>(so + m * kOutPitch + 4 * j);
    __stcg(reinterpret_cast<float4*>(ws_partial + (int64_t(split) * M + m) * NPAD + n0 + 4 * j), v);
  }
  __threadfence();
  named_bar_sync(1, 32 * kGemvConsumerWarps);
  if (tid == 0) {
    const int ticket = atomicAdd(ws_sem + nt, 1);
    s_last = (ticket == splits - 1);
  }
  named_bar_sync(1, 32 * kGemvConsumerWarps);
  if (!s_last) return;
  __threadfence();
  for (int idx = tid; idx < M * 32; idx += 32 * kGemvConsumerWarps) {
    const int m = idx >> 5, j = idx & 31;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int sp = 0; sp < splits; ++sp) {
      const float4 p = __ldcg(reinterpret_cast<const float4*>(ws_partial + (int64_t(sp) * M + m) * NPAD + n0 + 4 * j));
      v.x += p.x;
      v.y += p.y;
      v.z += p.z;
      v.w += p.w;
    }
    store_y(m, j, v);
  }
  if (tid == 0) ws_sem[nt] = 0;  // leave the workspace consistent for the next call
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------

struct GemvPlan {
  int nt8, splits, n_tiles, k_tiles, smem;
};

static GemvPlan gemv_plan(int64_t M, int64_t K, int64_t N) {
  GemvPlan p;
  p.nt8 = M <= 8 ? 1 : (M <= 16 ? 2 : (M <= 32 ? 4 : 8));
  p.n_tiles = int((N + kTileN - 1) / kTileN);
  p.k_tiles = int(K / kTileK);
  const int resident = (p.nt8 <= 2 ? 2 : 1) * sm_count();
  int splits = resident / p.n_tiles;
  if (splits < 1) splits = 1;
  if (splits > p.k_tiles) splits = p.k_tiles;
  if (splits > 32) splits = 32;
  p.splits = splits;
  p.smem = gemv_stages(p.nt8) * gemv_stage_bytes(p.nt8) + 16 * gemv_stages(p.nt8);
  return p;
}

template <typename T, int NT8>
static int gemv_launch(const GemvPlan& p, const void* x, const int32_t* wq, const void* sc, const uint8_t* zq,
                       const void* bias, void* y, int64_t M, int64_t K, int64_t N, void* workspace,
                       cudaStream_t st) {
  auto kern = gemv_kernel<T, NT8>;
  static bool attr_done[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && !attr_done[dev]) {
    AWQ_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, p.smem),
                   "awq_b200_gemv: cudaFuncSetAttribute");
    attr_done[dev] = true;
  }
  int* sem = reinterpret_cast<int*>(workspace);
  float* partial = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(workspace) + kSemInts * sizeof(int));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(unsigned(p.n_tiles * p.splits));
  cfg.blockDim = dim3(kGemvThreads);
  cfg.dynamicSmemBytes = size_t(p.smem);
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  AWQ_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, (const T*)x, (const uint32_t*)wq, (const T*)sc, zq, (const T*)bias,
                                    (T*)y, partial, sem, int(M), int(K), int(N), p.k_tiles, p.splits),
                 "awq_b200_gemv: launch");
  return 0;
}

}  // namespace awq

using namespace awq;

extern "C" int64_t awq_b200_workspace_bytes(int64_t max_M, int64_t max_N) {
  if (max_M < 1) max_M = 1;
  if (max_N < 1) max_N = 1;
  const int64_t nt = (max_N + kTileN - 1) / kTileN;
  // splits * n_tiles <= 2 * SMs + n_tiles; 1024 covers any sm_100 part
  const int64_t rows = (1024 + nt) * kTileN;
  const int64_t m = max_M < 256 ? max_M : 256;  // split-K is only used for M <= 256
  return int64_t(kSemInts) * 4 + rows * m * 4;
}

extern "C" int awq_b200_gemv(const void* x, const int32_t* qweight_t, const void* scales_t,
                             const uint8_t* qzeros_t, const void* bias, void* y, int64_t M, int64_t K,
                             int64_t N, int dtype, void* workspace, int64_t workspace_bytes, void* stream) {
  if (M == 0 || N == 0) return 0;
  AWQ_CHECK(M >= 1 && M <= 64, "awq_b200_gemv: M (%lld) must be in [1, 64]", (long long)M);
  AWQ_CHECK(K > 0 && K % kTileK == 0, "awq_b200_gemv: K (%lld) must be a positive multiple of 128", (long long)K);
  AWQ_CHECK(N > 0 && N % 8 == 0, "awq_b200_gemv: N (%lld) must be a positive multiple of 8", (long long)N);
  AWQ_CHECK(dtype == AWQ_B200_F16 || dtype == AWQ_B200_BF16, "awq_b200_gemv: bad dtype %d", dtype);
  AWQ_CHECK(x && qweight_t && scales_t && qzeros_t && y, "awq_b200_gemv: null pointer");
  AWQ_CHECK(((uintptr_t)x % 16) == 0 && ((uintptr_t)qweight_t % 16) == 0 && ((uintptr_t)scales_t % 16) == 0 &&
                ((uintptr_t)qzeros_t % 16) == 0 && ((uintptr_t)y % 8) == 0,
            "awq_b200_gemv: x/qweight_t/scales_t/qzeros_t must be 16-byte aligned, y 8-byte aligned");
  if (int rc = awq_b200_check_device()) return rc;
  const GemvPlan p = gemv_plan(M, K, N);
  AWQ_CHECK(p.n_tiles <= kSemInts, "awq_b200_gemv: N too large");
  if (p.splits > 1) {
    const int64_t npad = int64_t(p.n_tiles) * kTileN;
    const int64_t need = int64_t(kSemInts) * 4 + int64_t(p.splits) * M * npad * 4;
    AWQ_CHECK(workspace != nullptr && workspace_bytes >= need,
              "awq_b200_gemv: workspace too small (%lld < %lld bytes); size it with awq_b200_workspace_bytes",
              (long long)workspace_bytes, (long long)need);
    AWQ_CHECK(((uintptr_t)workspace % 16) == 0, "awq_b200_gemv: workspace must be 16-byte aligned");
  }
  cudaStream_t st = (cudaStream_t)stream;
#define AWQ_GEMV_CASE(TT, NN) \
  return gemv_launch<TT, NN>(p, x, qweight_t, scales_t, qzeros_t, bias, y, M, K, N, workspace, st)
  if (dtype == AWQ_B200_F16) {
    switch (p.nt8) {
      case 1: AWQ_GEMV_CASE(__half, 1);
      case 2: AWQ_GEMV_CASE(__half, 2);
      case 4: AWQ_GEMV_CASE(__half, 4);
      default: AWQ_GEMV_CASE(__half, 8);
    }
  } else {
    switch (p.nt8) {
      case 1: AWQ_GEMV_CASE(__nv_bfloat16, 1);
      case 2: AWQ_GEMV_CASE(__nv_bfloat16, 2);
      case 4: AWQ_GEMV_CASE(__nv_bfloat16, 4);
      default: AWQ_GEMV_CASE(__nv_bfloat16, 8);
    }
  }
#undef AWQ_GEMV_CASE
}
