// awq_moe.cu -- AWQ W4A16 mixture-of-experts linears for decode batches.
//
// Replaces the two grouped GEMMs of the reference's AWQMoEMethod.apply
// (python/sglang/srt/layers/quantization/awq.py:661-852 -> MoeRunner(MARLIN) -> fused_marlin_moe ->
// moe_wna16_marlin_gemm, sgl-kernel/csrc/moe/marlin_moe_wna16/): for every (token, k) PAIR of the router's top-k,
//     h[pair]   = silu(x[token] @ W13_gate[e]) * (x[token] @ W13_up[e])        e = topk_ids[pair]
//     o[pair]   = topk_weight[pair] * (h[pair] @ W2[e])
//     out[token] = sum_k o[token, k]
// At decode the op is HBM-bound: what matters is that the packed weights of every SELECTED expert stream exactly
// once, at full rate, and that the experts nobody selected cost nothing.  So the kernel is the one-slab-per-CTA
// register-streaming decode kernel (awq_gemv_direct.cu) with an expert dimension: grid = (16-row slabs, min(experts,
// pairs)).  Block (slab, j) takes the j-th distinct expert of the pair list, scans the list for its pairs (device-side: no host routing, CUDA-graph capturable), exits at
// once when nobody chose e, and otherwise streams the expert's slab over the whole K against up to kCap of its pairs
// at a time -- the pairs are the token columns of the MMA, their activation rows are GATHERED while staging.  More
// than kCap pairs per expert re-stream the 16-row slab (32 KB at K = 4096: an L2 hit).  Per-expert tile arrays are
// the dense path's records, stacked; W13 is repacked with gate / up rows interleaved
// (awq_b200_repack_gate_up) so SiluAndMul is the epilogue.  No sorting, no padding to block sizes, no atomics,
// deterministic.  Batches of hundreds of tokens belong to the grouped tcgen05 path on the host side
// (srt/layers/quantization/awq.py AWQMoEMethod.apply), which gathers the tokens of each expert.
#include "common.cuh"
#include "gemv_common.cuh"
#include "host_common.h"

#include <cmath>

namespace awq {

constexpr int kMoeWarps = 8;
constexpr int kMoeThreads = 32 * kMoeWarps;
constexpr int kMoeMaxPairs = 1024;  // pair-list capacity in shared memory

__host__ __device__ constexpr int moe_warp_bytes(int ngmax, int cap) { return ((ngmax * cap * kXBlock + ngmax * 32) + 127) & ~127; }
__host__ __device__ constexpr int moe_smem_bytes(int ngmax, int cap) {
  return kMoeWarps * moe_warp_bytes(ngmax, cap) + kMoeWarps * 512 + kMoeMaxPairs * 4;
}

struct MoeArgs {
  const int32_t* expert_ids;  // [pairs]; ids outside [0, E) belong to nobody: their rows are zero-filled by block (*, 0)
  const float* row_scale;     // [pairs] or null: multiplied into the row in fp32 before the single rounding
  int pairs, K, N, KT, ngmax, num_experts, x_div, silu;
  long long expert_stride;    // bytes between consecutive experts' tile arrays
};

// stage_x of gemv_common.cuh with gathered rows: token t of the chunk is row rows[t] of x (shared-memory table)
template <typename T, int kCap, bool kFast>
__device__ __forceinline__ void moe_stage_x(const T* __restrict__ x, int K, int k0, int ng, uint32_t xs, uint32_t sxs, int lane,
                                            const int* rows, int m) {
  constexpr int kU = 4;
  const int half = lane >> 4, j = lane & 15;
  const uint32_t dst_j = uint32_t(((j & 3) * 4 + (j >> 2)) * 16);
  const int nqp = (ng + 1) >> 1;
  const int nitems = nqp * m;
  const T* xcol = x + k0 + 8 * j;
  for (int it0 = 0; it0 < nitems; it0 += kU) {
    uint4 v[kU];
    int qs[kU], toks[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int it = it0 + u;
      const int qp = it / m;
      toks[u] = it - qp * m;
      qs[u] = 2 * qp + half;
      v[u] = make_uint4(0u, 0u, 0u, 0u);
      if (it < nitems && qs[u] < ng) v[u] = ldg128(xcol + int64_t(rows[toks[u]]) * K + qs[u] * kTileK);
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      if (it0 + u < nitems) {  // warp-uniform
        const bool ok = qs[u] < ng;
        const uint32_t dst = xs + uint32_t((qs[u] * kCap + toks[u]) * kXBlock) + dst_j;
        if constexpr (kFast) {
          if (ok) sts128(dst, make_uint4(v[u].x, v[u].z, v[u].y, v[u].w));
          float d[4] = {0.f, 0.f, 0.f, 0.f};
          const uint32_t a[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
          mma_16816<__half>(d, a, 0x3c003c00u, 0x3c003c00u);
          float sum = d[0] + d[2];
          sum += __shfl_xor_sync(0xffffffffu, sum, 4);
          sum += __shfl_xor_sync(0xffffffffu, sum, 8);
          if (ok && j == 0) sts32(sxs + uint32_t((qs[u] * 8 + toks[u]) * 4), __float_as_uint(sum * -5.9604644775390625e-08f));
        } else {
          if (ok) sts128(dst, v[u]);
        }
      }
    }
  }
}

template <typename T, int kCap, int kDepth>
__global__ void __launch_bounds__(kMoeThreads, 2)
moe_gemv_kernel(const T* __restrict__ x, const uint8_t* __restrict__ tiles, T* __restrict__ y, const MoeArgs a) {
  constexpr bool kFast = (Elem<T>::kDtype == 0);
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ int s_xrow[8];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, c = lane & 3;
  const int KT = a.KT;
  const int kb = (warp * KT) >> 3, ke = ((warp + 1) * KT) >> 3;
  const int ng = ke - kb;
  const int n_out = a.silu ? a.N >> 1 : a.N;
  int* list = reinterpret_cast<int*>(smem + kMoeWarps * moe_warp_bytes(a.ngmax, kCap) + kMoeWarps * 512);

  pdl_launch_dependents();
  pdl_wait();  // the expert ids, the activations and the row scales come from the previous kernels

  // ---- which expert is this block's?  blockIdx.y = j selects the j-th DISTINCT expert of the pair list (in id
  // order): the grid has min(experts, pairs) rows, so a batch of one token does not launch (and park, with their
  // shared memory) the blocks of the experts nobody chose ----
  __shared__ unsigned int s_present[32];
  __shared__ int s_expert;
  if (tid < 32) s_present[tid] = 0u;
  __syncthreads();
  for (int p = tid; p < a.pairs; p += kMoeThreads) {
    const int id = a.expert_ids[p];
    if (id >= 0 && id < a.num_experts) atomicOr(&s_present[id >> 5], 1u << (id & 31));
  }
  __syncthreads();
  if (warp == 0) {
    const unsigned int word = s_present[lane];
    int incl = __popc(word);  // inclusive prefix sum of the set bits over the 32 words
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += v;
    }
    const int excl = incl - __popc(word), j = int(blockIdx.y);
    if (lane == 0) s_expert = -1;
    __syncwarp();
    if (j >= excl && j < incl) s_expert = 32 * lane + int(__fns(word, 0, j - excl + 1));
  }
  __syncthreads();
  const int expert = s_expert;

  // ---- this expert's pairs, in pair order (every warp counts; warp 0 writes the list) ----
  int cnt = 0;
  for (int base = 0; base < a.pairs; base += 32) {
    const int p = base + lane;
    const int id = p < a.pairs ? a.expert_ids[p] : -2;
    const bool mine = id == expert && expert >= 0;
    const unsigned m = __ballot_sync(0xffffffffu, mine);
    if (warp == 0 && mine) list[cnt + __popc(m & ((1u << lane) - 1u))] = p;
    cnt += __popc(m);
    // pairs that belong to no expert of this rank (masked ids): the blocks of row 0 zero their rows
    if (blockIdx.y == 0 && warp == 1 && p < a.pairs && (id < 0 || id >= a.num_experts)) {
      for (int col = int(blockIdx.x); col < n_out; col += int(gridDim.x)) y[int64_t(p) * n_out + col] = Elem<T>::from_float(0.f);
    }
  }
  if (cnt == 0) return;  // fewer distinct experts than grid rows: no weight traffic at all
  __syncthreads();

  // ---- stream: the block walks slabs blockIdx.x, blockIdx.x + gridDim.x, ... of its expert.  With one round of pairs
  // (cnt <= kCap, the decode case) the activations are gathered and staged ONCE for all of them; the first weight
  // requests of the next slab are issued before the cross-warp reduction of the current one ----
  const uint8_t* base_e = tiles + int64_t(expert) * a.expert_stride;
  const int lane_w = g * 64 + 16 * (c ^ ((g >> 1) & 3)), lane_m = kTileMetaOff + 4 * g;
  const int dz = 32 - 2 * g;
  const uint32_t xs = smem_u32(smem) + uint32_t(warp * moe_warp_bytes(a.ngmax, kCap));
  const uint32_t sxs = xs + uint32_t(a.ngmax * kCap * kXBlock);
  float* red = reinterpret_cast<float*>(smem + kMoeWarps * moe_warp_bytes(a.ngmax, kCap));
  const int nslabs = (a.N + 15) >> 4;
  const int nrounds = (cnt + kCap - 1) / kCap;

  uint4 wa[kDepth], wb[kDepth];
  uint32_t s2[kDepth], z2[kDepth];
  auto slab_ptrs = [&](int slab, const uint8_t*& pw, const uint8_t*& pm) {
    const uint8_t* rec = base_e + (int64_t(slab >> 3) * KT + kb) * kTileBytes;
    pw = rec + ((slab & 7) * 1024 + lane_w);
    pm = rec + ((slab & 7) * kTileMetaBlock + lane_m);
  };
  auto request = [&](const uint8_t* pw, const uint8_t* pm) {
#pragma unroll
    for (int s = 0; s < kDepth; ++s)
      if (s < ng) {
        wa[s] = ldg128_stream(pw + s * kTileBytes);
        wb[s] = ldg128_stream(pw + s * kTileBytes + 512);
        s2[s] = ldg32_stream(pm + s * kTileBytes);
        z2[s] = ldg16_stream(pm + s * kTileBytes + dz);
      }
  };
  const uint8_t *pw, *pm;
  int slab = blockIdx.x;
  if (slab >= nslabs) return;
  slab_ptrs(slab, pw, pm);
  request(pw, pm);
  int staged_round = -1;

  while (slab < nslabs) {
    for (int round = 0; round < nrounds; ++round) {
      const int c0 = round * kCap;
      const int m = cnt - c0 < kCap ? cnt - c0 : kCap;  // pairs of this round = token columns of the MMA
      if (staged_round != round) {
        __syncthreads();  // everybody is done with the previous table (and with `red`)
        if (tid < 8) s_xrow[tid] = list[c0 + (tid < m ? tid : m - 1)] / a.x_div;
        __syncthreads();
        moe_stage_x<T, kCap, kFast>(x, a.K, kb * kTileK, ng, xs, sxs, lane, s_xrow, m);
        __syncwarp();
        staged_round = round;
      }
      uint32_t xq = xs + uint32_t((g < m ? g : m - 1) * kXBlock + 16 * c);
      uint32_t sxq = sxs + 8u * uint32_t(c);

      float acc[1][4] = {{0.f, 0.f, 0.f, 0.f}};
      const uint8_t *cw = pw, *cm = pm;
      for (int i = 0; i < ng; i += kDepth) {
#pragma unroll
        for (int s = 0; s < kDepth; ++s) {
          if (i + s < ng) {
            uint4 xf[1][4];
#pragma unroll
            for (int sp = 0; sp < 4; ++sp) xf[0][sp] = lds128(xq + uint32_t(s * kCap * kXBlock + 64 * sp));
            if constexpr (kFast) {
              const uint2 v = lds64(sxq + uint32_t(s * 32));
              const float2 nsx[1] = {make_float2(__uint_as_float(v.x), __uint_as_float(v.y))};
              group_mma_subnormal<1>(wa[s], wb[s], s2[s], z2[s], xf, nsx, acc);
            } else {
              gemv_group_mma<T, 1>(wa[s], wb[s], s2[s], z2[s], xf, acc);
            }
            if (i + s + kDepth < ng) {
              wa[s] = ldg128_stream(cw + (s + kDepth) * kTileBytes);
              wb[s] = ldg128_stream(cw + (s + kDepth) * kTileBytes + 512);
              s2[s] = ldg32_stream(cm + (s + kDepth) * kTileBytes);
              z2[s] = ldg16_stream(cm + (s + kDepth) * kTileBytes + dz);
            }
          }
        }
        cw += kDepth * kTileBytes;
        cm += kDepth * kTileBytes;
        xq += kDepth * kCap * kXBlock;
        sxq += kDepth * 32;
      }
      // the next work item's first requests go out before this one is reduced and stored: the same slab again for
      // the next round of pairs (an L1 / L2 hit), else the block's next slab
      const int out_slab = slab;
      if (round + 1 < nrounds) {
        request(pw, pm);
      } else {
        slab += gridDim.x;
        if (slab < nslabs) {
          slab_ptrs(slab, pw, pm);
          request(pw, pm);
        }
      }

      // cross-warp reduction in warp order (deterministic)
      __syncthreads();  // the previous item's readers of `red` are done
      *reinterpret_cast<float4*>(&red[warp * 128 + 4 * lane]) = make_float4(acc[0][0], acc[0][1], acc[0][2], acc[0][3]);
      __syncthreads();
      if (tid < 128) {
        // element i of lane L: i = 0: (row g, tok 2c)  1: (row g, tok 2c+1)  2: (row g+8, tok 2c)  3: (row g+8, tok 2c+1)
        const int L = tid >> 2, i = tid & 3;
        const int row = (L >> 2) + 8 * (i >> 1), t = 2 * (L & 3) + (i & 1);
        float sum = 0.f;
#pragma unroll
        for (int w = 0; w < kMoeWarps; ++w) sum += red[w * 128 + tid];
        const int p = list[c0 + (t < m ? t : 0)];
        if (a.silu) {
          // rows 0..7 = gate, rows 8..15 = up of the same 8 output columns (awq_b200_repack_gate_up): element i + 2
          const float gq = Elem<T>::to_float(Elem<T>::from_float(sum));  // the unfused sequence rounds the linear's output first
          const float uq = __shfl_down_sync(0xffffffffu, gq, 2);
          const int col = out_slab * 8 + row;
          if (i < 2 && t < m && col < n_out) y[int64_t(p) * n_out + col] = Elem<T>::from_float(gq / (1.f + __expf(-gq)) * uq);
        } else {
          const int n = out_slab * 16 + row;
          if (a.row_scale != nullptr && t < m) sum *= a.row_scale[p];
          if (t < m && n < a.N) y[int64_t(p) * a.N + n] = Elem<T>::from_float(sum);
        }
      }
    }
  }
}

// ---- long-K variant (W2 of Mixtral-sized experts, K = 14336; W13 of hidden-7168 models) ------------------------------
// moe_gemv_kernel stages a warp's WHOLE k-slice of x for every pair of the round, so a long K shrinks the round to 2 or
// 4 pairs and an expert with 8 pairs streams its slabs 4 times (Mixtral W2 at 32 tokens: 0.18 of HBM).  Here the
// warp's slice is cut into CHUNKS of kDepth records: x is staged per chunk for all 8 pairs (10 KB per warp whatever K
// is), and the block keeps the fp32 accumulators of kMoeSB slabs in registers while it walks the chunks -- one pass
// over the weights per 8 pairs at any K.  The kMoeSB slabs of a batch are the 8 slabs of ONE n-tile, so for a fixed
// chunk the block reads whole 8576-byte records.  Piece order: (batch of slabs) > round of 8 pairs > chunk > slab; the
// first requests of the next piece go out while the current one is multiplied.
constexpr int kMoeSB = 8;
__host__ __device__ constexpr int moe_chunk_warp_bytes(int depth) { return ((depth * 8 * kXBlock + depth * 32) + 127) & ~127; }
__host__ __device__ constexpr int moe_chunk_smem_bytes(int depth) {
  return kMoeWarps * moe_chunk_warp_bytes(depth) + 2 * kMoeWarps * 512 + kMoeMaxPairs * 4;
}

template <typename T, int kDepth>
__global__ void __launch_bounds__(kMoeThreads, 2)
moe_gemv_chunk_kernel(const T* __restrict__ x, const uint8_t* __restrict__ tiles, T* __restrict__ y, const MoeArgs a, int sb) {
  constexpr bool kFast = (Elem<T>::kDtype == 0);
  constexpr int kCap = 8;
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ int s_xrow[8];
  __shared__ unsigned int s_present[32];
  __shared__ int s_expert;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, c = lane & 3;
  const int KT = a.KT;
  const int kb = (warp * KT) >> 3, ke = ((warp + 1) * KT) >> 3;
  const int ng = ke - kb;
  const int n_out = a.silu ? a.N >> 1 : a.N;
  float* red = reinterpret_cast<float*>(smem + kMoeWarps * moe_chunk_warp_bytes(kDepth));
  int* list = reinterpret_cast<int*>(smem + kMoeWarps * moe_chunk_warp_bytes(kDepth) + 2 * kMoeWarps * 512);

  pdl_launch_dependents();
  pdl_wait();  // the expert ids, the activations and the row scales come from the previous kernels

  // ---- this block's expert = the blockIdx.y-th distinct one of the pair list, and its pairs (as in moe_gemv_kernel) ----
  if (tid < 32) s_present[tid] = 0u;
  __syncthreads();
  for (int p = tid; p < a.pairs; p += kMoeThreads) {
    const int id = a.expert_ids[p];
    if (id >= 0 && id < a.num_experts) atomicOr(&s_present[id >> 5], 1u << (id & 31));
  }
  __syncthreads();
  if (warp == 0) {
    const unsigned int word = s_present[lane];
    int incl = __popc(word);
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += v;
    }
    const int excl = incl - __popc(word), j = int(blockIdx.y);
    if (lane == 0) s_expert = -1;
    __syncwarp();
    if (j >= excl && j < incl) s_expert = 32 * lane + int(__fns(word, 0, j - excl + 1));
  }
  __syncthreads();
  const int expert = s_expert;
  int cnt = 0;
  for (int base = 0; base < a.pairs; base += 32) {
    const int p = base + lane;
    const int id = p < a.pairs ? a.expert_ids[p] : -2;
    const bool mine = id == expert && expert >= 0;
    const unsigned m = __ballot_sync(0xffffffffu, mine);
    if (warp == 0 && mine) list[cnt + __popc(m & ((1u << lane) - 1u))] = p;
    cnt += __popc(m);
    if (blockIdx.y == 0 && warp == 1 && p < a.pairs && (id < 0 || id >= a.num_experts)) {
      for (int col = int(blockIdx.x); col < n_out; col += int(gridDim.x)) y[int64_t(p) * n_out + col] = Elem<T>::from_float(0.f);
    }
  }
  if (cnt == 0) return;
  __syncthreads();

  const uint8_t* base_e = tiles + int64_t(expert) * a.expert_stride;
  const int lane_w = g * 64 + 16 * (c ^ ((g >> 1) & 3)), lane_m = kTileMetaOff + 4 * g;
  const int dz = 32 - 2 * g;
  const uint32_t xs = smem_u32(smem) + uint32_t(warp * moe_chunk_warp_bytes(kDepth));
  const uint32_t sxs = xs + uint32_t(kDepth * kCap * kXBlock);
  const int nslabs = (a.N + 15) >> 4;
  const int nbatches = (nslabs + sb - 1) / sb;
  const int nrounds = (cnt + kCap - 1) / kCap;
  const int nchunks = (ng + kDepth - 1) / kDepth;

  uint4 wa[kDepth], wb[kDepth];
  uint32_t s2[kDepth], z2[kDepth];
  auto piece_ptrs = [&](int batch, int chunk, int sl, const uint8_t*& pw, const uint8_t*& pm) {
    const int slab = batch * sb + sl;
    const uint8_t* rec = base_e + (int64_t(slab >> 3) * KT + kb + chunk * kDepth) * kTileBytes;
    pw = rec + ((slab & 7) * 1024 + lane_w);
    pm = rec + ((slab & 7) * kTileMetaBlock + lane_m);
  };
  auto load = [&](int s, const uint8_t* pw, const uint8_t* pm) {
    wa[s] = ldg128_stream(pw + s * kTileBytes);
    wb[s] = ldg128_stream(pw + s * kTileBytes + 512);
    s2[s] = ldg32_stream(pm + s * kTileBytes);
    z2[s] = ldg16_stream(pm + s * kTileBytes + dz);
  };
  auto chunk_len = [&](int chunk) { return ng - chunk * kDepth < kDepth ? ng - chunk * kDepth : kDepth; };

  int batch = blockIdx.x;
  if (batch >= nbatches) return;
  if (nchunks > 0) {
    const uint8_t *pw, *pm;
    piece_ptrs(batch, 0, 0, pw, pm);
#pragma unroll
    for (int s = 0; s < kDepth; ++s)
      if (s < chunk_len(0)) load(s, pw, pm);
  }
  int rp = 0;  // parity of the reduction buffer

  for (; batch < nbatches; batch += gridDim.x) {
    const int nsl = nslabs - batch * sb < sb ? nslabs - batch * sb : sb;
    for (int round = 0; round < nrounds; ++round) {
      const int c0 = round * kCap;
      const int m = cnt - c0 < kCap ? cnt - c0 : kCap;  // pairs of this round = token columns of the MMA
      __syncthreads();  // everybody is done with the previous table
      if (tid < 8) s_xrow[tid] = list[c0 + (tid < m ? tid : m - 1)] / a.x_div;
      __syncthreads();
      const uint32_t xq = xs + uint32_t((g < m ? g : m - 1) * kXBlock + 16 * c);
      const uint32_t sxq = sxs + 8u * uint32_t(c);

      float acc[kMoeSB][4];
#pragma unroll
      for (int sl = 0; sl < kMoeSB; ++sl) acc[sl][0] = acc[sl][1] = acc[sl][2] = acc[sl][3] = 0.f;

      for (int chunk = 0; chunk < nchunks; ++chunk) {
        const int ngc = chunk_len(chunk);
        __syncwarp();  // the warp's readers of the previous chunk are done (its x region is private)
        moe_stage_x<T, kCap, kFast>(x, a.K, (kb + chunk * kDepth) * kTileK, ngc, xs, sxs, lane, s_xrow, m);
        __syncwarp();
#pragma unroll
        for (int sl = 0; sl < kMoeSB; ++sl) {
          if (sl < nsl) {
            // the piece after this one: next slab of the batch, else next chunk, else next round, else next batch
            int nb = batch, nc = chunk, ns = sl + 1;
            bool has_next = true;
            if (ns >= nsl) {
              ns = 0;
              if (++nc >= nchunks) {
                nc = 0;
                if (round + 1 >= nrounds) {
                  nb = batch + int(gridDim.x);
                  has_next = nb < nbatches;
                }
              }
            }
            const uint8_t *npw = nullptr, *npm = nullptr;
            int nlen = 0;
            if (has_next) {
              piece_ptrs(nb, nc, ns, npw, npm);
              nlen = chunk_len(nc);
            }
#pragma unroll
            for (int s = 0; s < kDepth; ++s) {
              if (s < ngc) {
                uint4 xf[1][4];
#pragma unroll
                for (int sp = 0; sp < 4; ++sp) xf[0][sp] = lds128(xq + uint32_t(s * kCap * kXBlock + 64 * sp));
                if constexpr (kFast) {
                  const uint2 v = lds64(sxq + uint32_t(s * 32));
                  const float2 nsx[1] = {make_float2(__uint_as_float(v.x), __uint_as_float(v.y))};
                  group_mma_subnormal<1>(wa[s], wb[s], s2[s], z2[s], xf, nsx, &acc[sl]);
                } else {
                  gemv_group_mma<T, 1>(wa[s], wb[s], s2[s], z2[s], xf, &acc[sl]);
                }
              }
              if (s < nlen) load(s, npw, npm);
            }
          }
        }
      }

      // cross-warp reduction in warp order (deterministic), slab by slab, double-buffered: one barrier per slab
#pragma unroll
      for (int sl = 0; sl < kMoeSB; ++sl) {
        if (sl < nsl) {
          float* buf = red + rp * (kMoeWarps * 128);
          rp ^= 1;
          *reinterpret_cast<float4*>(&buf[warp * 128 + 4 * lane]) = make_float4(acc[sl][0], acc[sl][1], acc[sl][2], acc[sl][3]);
          __syncthreads();
          if (tid < 128) {
            const int out_slab = batch * sb + sl;
            const int L = tid >> 2, i = tid & 3;
            const int row = (L >> 2) + 8 * (i >> 1), t = 2 * (L & 3) + (i & 1);
            float sum = 0.f;
#pragma unroll
            for (int w = 0; w < kMoeWarps; ++w) sum += buf[w * 128 + tid];
            const int p = list[c0 + (t < m ? t : 0)];
            if (a.silu) {
              const float gq = Elem<T>::to_float(Elem<T>::from_float(sum));
              const float uq = __shfl_down_sync(0xffffffffu, gq, 2);
              const int col = out_slab * 8 + row;
              if (i < 2 && t < m && col < n_out) y[int64_t(p) * n_out + col] = Elem<T>::from_float(gq / (1.f + __expf(-gq)) * uq);
            } else {
              const int n = out_slab * 16 + row;
              if (a.row_scale != nullptr && t < m) sum *= a.row_scale[p];
              if (t < m && n < a.N) y[int64_t(p) * a.N + n] = Elem<T>::from_float(sum);
            }
          }
        }
      }
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

static int g_moe_blocks_per_sm = 0;  // debug: blocks per SM the grid is sized for (0 = default)

template <typename T, int kCap>
static int moe_launch_cap(const void* x, const void* tiles, void* y, const MoeArgs& a, cudaStream_t st) {
  constexpr int kDepth = 4;
  const int smem = moe_smem_bytes(a.ngmax, kCap);
  static int configured[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (smem > configured[dev]) {
    AWQ_CUDA_CHECK(cudaFuncSetAttribute(moe_gemv_kernel<T, kCap, kDepth>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem),
                   "awq_b200_moe_linear: smem opt-in");
    configured[dev] = smem;
  }
  cudaLaunchConfig_t cfg = {};
  // rows = candidates for distinct experts; columns so that ~4 blocks per SM exist in total (2 resident): a block
  // walks N / 16 / columns slabs
  const int rows = a.num_experts < a.pairs ? a.num_experts : a.pairs;
  const int nslabs = (a.N + 15) / 16;
  // blocks per SM the grid is sized for (2 are resident).  The rows of experts nobody chose exit at once, and how many
  // there are is only known on the device: with E experts and p pairs about E (1 - (1 - 1/E)^p) rows are live.  Measured
  // on the Mixtral-shaped layer (scripts/moe_probe.py, MOE_BLOCKS_PER_SM): one token (2 pairs, both rows live) is 8 %
  // faster as ONE wave; 8 pairs over 8 experts (a third of the rows empty) 9 % faster with twice as many, smaller blocks
  int per_sm = 4;
  if (a.pairs <= 2) {
    per_sm = 2;
  } else {
    const double live = a.num_experts * (1.0 - pow(1.0 - 1.0 / a.num_experts, double(a.pairs)));
    if (live < 0.75 * rows) per_sm = 8;
  }
  if (g_moe_blocks_per_sm > 0) per_sm = g_moe_blocks_per_sm;
  int cols = (per_sm * sm_count() + rows - 1) / rows;
  if (cols > nslabs) cols = nslabs;
  cfg.gridDim = dim3(unsigned(cols), unsigned(rows));
  cfg.blockDim = dim3(kMoeThreads);
  cfg.dynamicSmemBytes = size_t(smem);
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  AWQ_CUDA_CHECK(cudaLaunchKernelEx(&cfg, moe_gemv_kernel<T, kCap, kDepth>, (const T*)x, (const uint8_t*)tiles, (T*)y, a),
                 "awq_b200_moe_linear: launch");
  return 0;
}

static int g_moe_chunk = 1;  // debug: 0 keeps long-K problems on moe_gemv_kernel (rounds of 2 / 4 pairs), 2 sends all of them to the chunked kernel

template <typename T>
static int moe_launch_chunk(const void* x, const void* tiles, void* y, const MoeArgs& a, cudaStream_t st) {
  constexpr int kDepth = 4;
  constexpr int smem = moe_chunk_smem_bytes(kDepth);
  static bool configured[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (!configured[dev]) {
    AWQ_CUDA_CHECK(cudaFuncSetAttribute(moe_gemv_chunk_kernel<T, kDepth>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem),
                   "awq_b200_moe_linear: smem opt-in (long K)");
    configured[dev] = true;
  }
  // slabs per batch: a whole n-tile (8 slabs) unless that leaves fewer than two blocks per SM
  const int rows = a.num_experts < a.pairs ? a.num_experts : a.pairs;
  const int nslabs = (a.N + 15) / 16;
  int sb = kMoeSB;
  while (sb > 1 && ((nslabs + sb - 1) / sb) * rows < 2 * sm_count()) sb >>= 1;
  const int nbatches = (nslabs + sb - 1) / sb;
  int cols = (4 * sm_count() + rows - 1) / rows;
  if (cols > nbatches) cols = nbatches;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(unsigned(cols), unsigned(rows));
  cfg.blockDim = dim3(kMoeThreads);
  cfg.dynamicSmemBytes = size_t(smem);
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  AWQ_CUDA_CHECK(cudaLaunchKernelEx(&cfg, moe_gemv_chunk_kernel<T, kDepth>, (const T*)x, (const uint8_t*)tiles, (T*)y, a, sb),
                 "awq_b200_moe_linear: launch (long K)");
  return 0;
}

// token columns per round: as many of an expert's pairs as fit beside each other in shared memory (<= 8 = the MMA's n);
// when the average expert has more pairs than such a round holds (long K: 2 or 4), the chunked kernel takes over
template <typename T>
static int moe_launch(const void* x, const void* tiles, void* y, const MoeArgs& a, cudaStream_t st) {
  const int budget = 96 * 1024;  // two CTAs per SM
  int cap = 8;
  while (cap > 1 && (moe_smem_bytes(a.ngmax, cap) > budget || cap / 2 >= a.pairs)) cap >>= 1;
  // ... when the AVERAGE expert has more pairs than one round holds (measured, Mixtral W2: 4 pairs over 4 experts or 16
  // over 8 are faster on the whole-K kernel, whose x is staged once per block)
  const int rows = a.num_experts < a.pairs ? a.num_experts : a.pairs;
  if (g_moe_chunk && (a.pairs > cap * rows || g_moe_chunk == 2) && a.KT >= kMoeWarps) return moe_launch_chunk<T>(x, tiles, y, a, st);
  AWQ_CHECK(moe_smem_bytes(a.ngmax, cap) <= 227 * 1024, "awq_b200_moe_linear: K too large (%d B of shared memory)",
            moe_smem_bytes(a.ngmax, cap));
  switch (cap) {
    case 8: return moe_launch_cap<T, 8>(x, tiles, y, a, st);
    case 4: return moe_launch_cap<T, 4>(x, tiles, y, a, st);
    case 2: return moe_launch_cap<T, 2>(x, tiles, y, a, st);
    default: return moe_launch_cap<T, 1>(x, tiles, y, a, st);
  }
}

}  // namespace awq

using namespace awq;

extern "C" int awq_b200_moe_linear(const void* x, const void* tiles, const int32_t* expert_ids, const float* row_scale, void* y,
                                   int64_t pairs, int64_t x_div, int64_t K, int64_t N, int64_t num_experts, int silu,
                                   int dtype, void* stream) {
  if (pairs == 0 || N == 0) return 0;
  AWQ_CHECK(pairs > 0 && pairs <= kMoeMaxPairs, "awq_b200_moe_linear: pairs (%lld) must be in [1, %d]", (long long)pairs, kMoeMaxPairs);
  AWQ_CHECK(num_experts <= 1024, "awq_b200_moe_linear: at most 1024 experts");
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
  a.pairs = int(pairs);
  a.K = int(K);
  a.N = int(N);
  a.KT = int(K / kTileK);
  a.ngmax = (a.KT + kMoeWarps - 1) / kMoeWarps;
  a.num_experts = int(num_experts);
  a.x_div = int(x_div);
  a.silu = silu ? 1 : 0;
  a.expert_stride = awq_b200_num_tiles(K, N) * (long long)kTileBytes;
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == AWQ_B200_F16) return moe_launch<__half>(x, tiles, y, a, st);
  return moe_launch<__nv_bfloat16>(x, tiles, y, a, st);
}

extern "C" void awq_b200_debug_set_moe_chunk(int on) { g_moe_chunk = on; }
extern "C" void awq_b200_debug_set_moe_blocks_per_sm(int n) { g_moe_blocks_per_sm = n; }

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
