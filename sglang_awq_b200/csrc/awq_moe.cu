// awq_moe.cu -- AWQ W4A16 mixture-of-experts linears for decode batches.
//
// Replaces the two grouped GEMMs of the reference's AWQMoEMethod.apply
// (python/sglang/srt/layers/quantization/awq.py:661-852 -> MoeRunner(MARLIN) -> fused_marlin_moe ->
// moe_wna16_marlin_gemm, sgl-kernel/csrc/moe/marlin_moe_wna16/): for every (token, k) PAIR of the router's top-k,
//     h[pair]   = silu(x[token] @ W13_gate[e]) * (x[token] @ W13_up[e])        e = topk_ids[pair]
//     o[pair]   = topk_weight[pair] * (h[pair] @ W2[e])
//     out[token] = sum_k o[token, k]
// At decode the op is a batch of independent GEMVs, each against ONE expert's matrix, and it is HBM-bound: what
// matters is that the selected experts' packed weights stream once, at full rate, with nothing else in the way.
// So the kernel is the one-slab-per-CTA register-streaming decode kernel (awq_gemv_direct.cu) with an expert
// indirection in front of it: grid = (16-row slabs, pairs); block (slab, pair) reads the pair's expert id, offsets
// into the stacked per-expert tile arrays (same 8576-byte records as the dense path; the W13 arrays
// are repacked with gate / up rows interleaved, awq_b200_repack_gate_up, so SiluAndMul is the epilogue) and streams
// that expert's slab over the whole K.  No sorting, no padding to block sizes, no atomics, deterministic, CUDA-graph
// capturable (the expert ids are read on the device).  Pairs routed to the same expert re-read its weights (from
// L2 when they run together); batches with many pairs per expert belong to the grouped tcgen05 path on the host
// side (srt/layers/quantization/awq.py AWQMoEMethod.apply), which gathers the tokens of each expert.
#include "common.cuh"
#include "gemv_common.cuh"
#include "host_common.h"

namespace awq {

constexpr int kMoeWarps = 8;
constexpr int kMoeThreads = 32 * kMoeWarps;

__host__ __device__ constexpr int moe_warp_bytes(int ngmax) { return ((ngmax * kXBlock + ngmax * 32) + 127) & ~127; }
__host__ __device__ constexpr int moe_smem_bytes(int ngmax) { return kMoeWarps * moe_warp_bytes(ngmax) + kMoeWarps * 512; }

struct MoeArgs {
  const int32_t* expert_ids;  // [pairs], -1 (or any id outside [0, E)) = not routed here: the row is zero-filled
  const float* row_scale;     // [pairs] or null: multiplied into the row in fp32 before the single rounding
  int K, N, KT, ngmax, num_experts, x_div, silu;
  long long expert_stride;    // bytes between consecutive experts' tile arrays
};

template <typename T, int kDepth>
__global__ void __launch_bounds__(kMoeThreads, 3)
moe_gemv_kernel(const T* __restrict__ x, const uint8_t* __restrict__ tiles, T* __restrict__ y, const MoeArgs a) {
  constexpr bool kFast = (Elem<T>::kDtype == 0);
  extern __shared__ __align__(128) uint8_t smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, c = lane & 3;
  const int slab = blockIdx.x, nt = slab >> 3, sl = slab & 7, pair = blockIdx.y;
  const int KT = a.KT;
  const int kb = (warp * KT) >> 3, ke = ((warp + 1) * KT) >> 3;
  const int ng = ke - kb;
  const int n_out = a.silu ? a.N >> 1 : a.N;

  pdl_launch_dependents();
  pdl_wait();  // the expert ids, the activations and the row scales come from the previous kernels
  const int expert = a.expert_ids[pair];
  if (expert < 0 || expert >= a.num_experts) {  // block-uniform
    const int per = a.silu ? 8 : 16, col0 = slab * per;
    if (tid < per && col0 + tid < n_out) y[int64_t(pair) * n_out + col0 + tid] = Elem<T>::from_float(0.f);
    return;
  }
  const uint8_t* base = tiles + int64_t(expert) * a.expert_stride;
  const uint8_t* pw = base + (int64_t(nt) * KT + kb) * kTileBytes + (sl * 1024 + g * 64 + 16 * (c ^ ((g >> 1) & 3)));
  const uint8_t* pm = base + (int64_t(nt) * KT + kb) * kTileBytes + (kTileMetaOff + sl * kTileMetaBlock + 4 * g);
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

  // this warp's k-range of the pair's activation row, staged once (fragment layout of gemv_common.cuh)
  const T* xrow = x + int64_t(pair / a.x_div) * a.K;
  const uint32_t xs = smem_u32(smem) + uint32_t(warp * moe_warp_bytes(a.ngmax));
  const uint32_t sxs = xs + uint32_t(a.ngmax * kXBlock);
  stage_x<T, 1, 8, kFast>(xrow, a.K, kb * kTileK, ng, xs, sxs, 0, 1, lane);
  __syncwarp();
  uint32_t xq = xs + uint32_t(16 * c);  // one token: every lane group reads token 0 (columns >= 1 are never stored)
  uint32_t sxq = sxs + 8u * uint32_t(c);

  float acc[1][4] = {{0.f, 0.f, 0.f, 0.f}};
  for (int i = 0; i < ng; i += kDepth) {
#pragma unroll
    for (int s = 0; s < kDepth; ++s) {
      if (i + s < ng) {
        uint4 xf[1][4];
#pragma unroll
        for (int sp = 0; sp < 4; ++sp) xf[0][sp] = lds128(xq + uint32_t(s * kXBlock + 64 * sp));
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
    xq += kDepth * kXBlock;
    sxq += kDepth * 32;
  }

  // cross-warp reduction in warp order (deterministic); token 0 lives in elements i = 0 (row g) and i = 2 (row g + 8)
  // of the lanes with c = 0
  float* red = reinterpret_cast<float*>(smem + kMoeWarps * moe_warp_bytes(a.ngmax));
  *reinterpret_cast<float4*>(&red[warp * 128 + 4 * lane]) = make_float4(acc[0][0], acc[0][1], acc[0][2], acc[0][3]);
  __syncthreads();
  if (tid < 16) {
    const int row = tid;  // row of the slab: lane L = 4 * (row & 7), element 2 * (row >> 3)
    const int e = 4 * (4 * (row & 7)) + 2 * (row >> 3);
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < kMoeWarps; ++w) s += red[w * 128 + e];
    if (a.silu) {
      // rows 0..7 = gate, rows 8..15 = up of the same 8 output columns (awq_b200_repack_gate_up)
      const float gq = Elem<T>::to_float(Elem<T>::from_float(s));  // the unfused sequence rounds the linear's output first
      const float uq = __shfl_down_sync(0x0000ffffu, gq, 8);
      const int col = slab * 8 + row;
      if (row < 8 && col < n_out) y[int64_t(pair) * n_out + col] = Elem<T>::from_float(gq / (1.f + __expf(-gq)) * uq);
    } else {
      const int n = slab * 16 + row;
      if (a.row_scale != nullptr) s *= a.row_scale[pair];
      if (n < a.N) y[int64_t(pair) * a.N + n] = Elem<T>::from_float(s);
    }
  }
}

// out[token, :] = T(sum_k float(o[token, k, :])): the reference sums the top-k rows with torch.sum (fp32 accumulation,
// one rounding; fused_marlin_moe's final reduction over intermediate_cache3)
template <typename T>
__global__ void __launch_bounds__(256)
moe_sum_kernel(const uint4* __restrict__ o, uint4* __restrict__ out, int topk, int vec_per_row, long long nvec) {
  pdl_wait();
  const long long i = blockIdx.x * 256ll + threadIdx.x;
  if (i < nvec) {
    const long long tok = i / vec_per_row, v = i - tok * vec_per_row;
    float s[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int k = 0; k < topk; ++k) {
      const uint4 p = o[(tok * topk + k) * vec_per_row + v];
      const T* h = reinterpret_cast<const T*>(&p);
#pragma unroll
      for (int j = 0; j < 8; ++j) s[j] += Elem<T>::to_float(h[j]);
    }
    uint4 r;
    T* rh = reinterpret_cast<T*>(&r);
#pragma unroll
    for (int j = 0; j < 8; ++j) rh[j] = Elem<T>::from_float(s[j]);
    out[i] = r;
  }
  pdl_launch_dependents();
}

template <typename T>
static int moe_launch(const void* x, const void* tiles, void* y, const MoeArgs& a, int64_t pairs, cudaStream_t st) {
  constexpr int kDepth = 4;
  const int smem = moe_smem_bytes(a.ngmax);
  AWQ_CHECK(smem <= 227 * 1024, "awq_b200_moe_linear: K too large (%d B of shared memory)", smem);
  static int configured[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (smem > configured[dev]) {
    AWQ_CUDA_CHECK(cudaFuncSetAttribute(moe_gemv_kernel<T, kDepth>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem),
                   "awq_b200_moe_linear: smem opt-in");
    configured[dev] = smem;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(unsigned((a.N + 15) / 16), unsigned(pairs));
  cfg.blockDim = dim3(kMoeThreads);
  cfg.dynamicSmemBytes = size_t(smem);
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  AWQ_CUDA_CHECK(cudaLaunchKernelEx(&cfg, moe_gemv_kernel<T, kDepth>, (const T*)x, (const uint8_t*)tiles, (T*)y, a),
                 "awq_b200_moe_linear: launch");
  return 0;
}

}  // namespace awq

using namespace awq;

extern "C" int awq_b200_moe_linear(const void* x, const void* tiles, const int32_t* expert_ids, const float* row_scale, void* y,
                                   int64_t pairs, int64_t x_div, int64_t K, int64_t N, int64_t num_experts, int silu,
                                   int dtype, void* stream) {
  if (pairs == 0 || N == 0) return 0;
  AWQ_CHECK(pairs > 0 && pairs <= 65535, "awq_b200_moe_linear: pairs (%lld) must be in [1, 65535]", (long long)pairs);
  AWQ_CHECK(x_div >= 1, "awq_b200_moe_linear: x_div must be >= 1");
  AWQ_CHECK(K > 0 && K % kTileK == 0, "awq_b200_moe_linear: K (%lld) must be a positive multiple of 128", (long long)K);
  AWQ_CHECK(N > 0 && N % (silu ? 16 : 8) == 0, "awq_b200_moe_linear: bad N (%lld)", (long long)N);
  AWQ_CHECK(num_experts >= 1, "awq_b200_moe_linear: num_experts must be >= 1");
  AWQ_CHECK(!(silu && row_scale), "awq_b200_moe_linear: row_scale is for the down projection only");
  AWQ_CHECK(dtype == AWQ_B200_F16 || dtype == AWQ_B200_BF16, "awq_b200_moe_linear: bad dtype %d", dtype);
  AWQ_CHECK(x && tiles && expert_ids && y, "awq_b200_moe_linear: null pointer");
  AWQ_CHECK(((uintptr_t)x % 16) == 0 && ((uintptr_t)tiles % 16) == 0, "awq_b200_moe_linear: x and tiles must be 16-byte aligned");
  if (int rc = awq_b200_check_device()) return rc;
  MoeArgs a;
  a.expert_ids = expert_ids;
  a.row_scale = row_scale;
  a.K = int(K);
  a.N = int(N);
  a.KT = int(K / kTileK);
  a.ngmax = (a.KT + kMoeWarps - 1) / kMoeWarps;
  a.num_experts = int(num_experts);
  a.x_div = int(x_div);
  a.silu = silu ? 1 : 0;
  a.expert_stride = awq_b200_num_tiles(K, N) * (long long)kTileBytes;
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == AWQ_B200_F16) return moe_launch<__half>(x, tiles, y, a, pairs, st);
  return moe_launch<__nv_bfloat16>(x, tiles, y, a, pairs, st);
}

extern "C" int awq_b200_moe_sum(const void* o, void* out, int64_t tokens, int64_t topk, int64_t hidden, int dtype, void* stream) {
  if (tokens == 0 || hidden == 0) return 0;
  AWQ_CHECK(tokens > 0 && topk >= 1 && hidden > 0 && hidden % 8 == 0, "awq_b200_moe_sum: bad sizes");
  AWQ_CHECK(dtype == AWQ_B200_F16 || dtype == AWQ_B200_BF16, "awq_b200_moe_sum: bad dtype %d", dtype);
  AWQ_CHECK(o && out && ((uintptr_t)o % 16) == 0 && ((uintptr_t)out % 16) == 0, "awq_b200_moe_sum: pointers must be 16-byte aligned");
  if (int rc = awq_b200_check_device()) return rc;
  const long long nvec = tokens * (hidden / 8);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(unsigned((nvec + 255) / 256));
  cfg.blockDim = dim3(256);
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  if (dtype == AWQ_B200_F16)
    AWQ_CUDA_CHECK(cudaLaunchKernelEx(&cfg, moe_sum_kernel<__half>, (const uint4*)o, (uint4*)out, int(topk), int(hidden / 8), nvec),
                   "awq_b200_moe_sum: launch");
  else
    AWQ_CUDA_CHECK(cudaLaunchKernelEx(&cfg, moe_sum_kernel<__nv_bfloat16>, (const uint4*)o, (uint4*)out, int(topk),
                                      int(hidden / 8), nvec),
                   "awq_b200_moe_sum: launch");
  return 0;
}
