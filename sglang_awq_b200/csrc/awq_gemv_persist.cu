// awq_gemv_persist.cu -- kernel 2b: persistent decode kernel, y[M,N] = x[M,K] @ W (+ bias),
// for every (M, K) whose activations fit in one SM's shared memory (K * M * 2.5 B <= ~190 KB).
//
// Replaces the three launches of the reference's AWQLinearMethod.apply
// (python/sglang/srt/layers/quantization/awq.py:434-451) for decode batches.
//
// Why persistent (measured with the per-CTA phase trace, scripts/trace_gemv.py): a short-lived
// CTA has HBM requests in flight only during its first ~2 us (the loaded DRAM latency); while
// it multiplies, reduces and retires, its slot streams nothing, and its prologue, x staging and
// epilogue cost as many instructions as the multiply itself.  Here
//   * ONE CTA per SM stages the WHOLE x[M, K] (and its per-group sums) in shared memory ONCE;
//   * its warps form TEAMS of kTeam warps; a team owns 16-row output slabs
//     team, team + T, team + 2T, ... (T = teams on the chip, numbered SM-interleaved), so at
//     any moment the chip reads a contiguous window of the weight array (HBM wants a sweep) and
//     every SM gets the same number of slabs +-1;
//   * inside a team the warps split K; a warp's weight stream is ONE continuous sequence of
//     register-direct LDG.128 with kDepth groups in flight that runs straight across slab
//     boundaries -- the next slab's bytes are already on their way while the team closes the
//     current slab (one named barrier + a 128-thread sum through shared memory, double
//     buffered, deterministic warp order);
//   * the first requests are issued before griddepcontrol.wait (weights never depend on the
//     previous kernel);
//   * fp16: subnormal tensor-core operands + group-sum correction (gemv_common.cuh); bf16: exact
//     per-weight path.
#include "common.cuh"
#include "gemv_common.cuh"
#include "host_common.h"

namespace awq {

__host__ __device__ constexpr int per_sx_off(int KT, int M) { return (KT * M * kXBlock + 127) & ~127; }
__host__ __device__ constexpr int per_red_off(int KT, int M, int nt8) { return per_sx_off(KT, M) + ((KT * nt8 * 32 + 127) & ~127); }
__host__ __device__ constexpr int per_smem_bytes(int KT, int M, int nt8, int warps) {
  return per_red_off(KT, M, nt8) + 2 * warps * nt8 * 512;
}

// kWarps warps per CTA (1 CTA per SM), teams of `team` warps (8, 16 or 32; runtime), kDepth
// groups in flight per warp.  kEven: every warp has the same number of groups per slab and it is
// a multiple of kDepth (host-checked) -- the stream loop then carries no per-group bookkeeping.
template <typename T, int M, int kWarps, int kDepth, bool kEven, int kMinB, bool kPush>
__global__ void __launch_bounds__(32 * kWarps, kMinB)
gemv_persist_kernel(const T* __restrict__ x, const uint8_t* __restrict__ tiles, const T* __restrict__ bias,
                    T* __restrict__ y, int K, int N, int KT, int lg_team, int nslabs, long long pf_bytes, int silu, const ArPush push) {
  constexpr bool kFast = (Elem<T>::kDtype == 0);
  constexpr int NT8 = (M + 7) / 8;
  extern __shared__ __align__(128) uint8_t smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, c = lane & 3;
  if (pf_bytes > 0 && lane == 0) l2_prefetch_sweep(tiles, pf_bytes, int(blockIdx.x), int(gridDim.x), warp, kWarps);
  const int team_warps = 1 << lg_team;
  const int teams_per_cta = kWarps >> lg_team;
  const int team_in_cta = warp >> lg_team, wt = warp & (team_warps - 1);
  const int T_total = int(gridDim.x) * teams_per_cta;
  const int team = team_in_cta * int(gridDim.x) + int(blockIdx.x);  // SM-interleaved numbering
  // this warp's groups [kb, ke) of every slab of its team
  const int kb = (wt * KT) >> lg_team, ke = ((wt + 1) * KT) >> lg_team;
  const int ng = ke - kb;
  const int my_slabs = team < nslabs ? (nslabs - 1 - team) / T_total + 1 : 0;
  const int total = my_slabs * ng;  // groups in this warp's stream

  // prefetch stream state: slab ps, group pq of it, pointers to that group
  int pq = 0, pleft = total, ps = team;
  const int lane_w = g * 64 + 16 * (c ^ ((g >> 1) & 3)), lane_m = kTileMetaOff + 4 * g;
  const int dz = 32 - 2 * g;
  const uint8_t *pw, *pm;
  auto set_slab_ptrs = [&]() {
    const uint8_t* rec = tiles + (int64_t(ps >> 3) * KT + kb) * kTileBytes;
    pw = rec + ((ps & 7) * 1024 + lane_w);
    pm = rec + ((ps & 7) * kTileMetaBlock + lane_m);
  };
  set_slab_ptrs();

  uint4 wa[kDepth], wb[kDepth];
  uint32_t s2[kDepth], z2[kDepth];
  auto prefetch = [&](int s) {
    if (pleft > 0) {
      wa[s] = ldg128_stream(pw);
      wb[s] = ldg128_stream(pw + 512);
      s2[s] = ldg32_stream(pm);
      z2[s] = ldg16_stream(pm + dz);
      pw += kTileBytes;
      pm += kTileBytes;
      --pleft;
      if (++pq == ng) {
        pq = 0;
        ps += T_total;
        set_slab_ptrs();
      }
    }
  };
  if constexpr (kEven) {
    if (my_slabs > 0) {
#pragma unroll
      for (int s = 0; s < kDepth; ++s) {
        wa[s] = ldg128_stream(pw + s * kTileBytes);
        wb[s] = ldg128_stream(pw + s * kTileBytes + 512);
        s2[s] = ldg32_stream(pm + s * kTileBytes);
        z2[s] = ldg16_stream(pm + s * kTileBytes + dz);
      }
    }
  } else {
#pragma unroll
    for (int s = 0; s < kDepth; ++s) prefetch(s);
  }

  pdl_launch_dependents();
  pdl_wait();  // x and y belong to the previous kernel until here

  // ---- stage all of x (+ group sums) once per SM ----------------------------------------------
  const uint32_t xs = smem_u32(smem);
  const uint32_t sxs = xs + uint32_t(per_sx_off(KT, M));
  stage_x<T, M, NT8 * 8, kFast>(x, K, 0, KT, xs, sxs, warp, kWarps, lane);
  __syncthreads();

  // ---- stream ---------------------------------------------------------------------------------
  uint32_t xq[NT8];
#pragma unroll
  for (int t = 0; t < NT8; ++t) {
    const int tok = 8 * t + g;
    xq[t] = xs + uint32_t(((kb * M) + (tok < M ? tok : M - 1)) * kXBlock + 16 * c);
  }
  uint32_t sxq = sxs + uint32_t(kb * NT8 * 32) + 8u * uint32_t(c);
  float* red = reinterpret_cast<float*>(smem + per_red_off(KT, M, NT8)) + team_in_cta * (2 * team_warps * NT8 * 128);
  const int bar_id = 1 + team_in_cta, bar_threads = 32 * team_warps;
  const int tt = tid & (bar_threads - 1);  // thread index inside the team

  float acc[NT8][4];
#pragma unroll
  for (int t = 0; t < NT8; ++t)
#pragma unroll
    for (int i = 0; i < 4; ++i) acc[t][i] = 0.f;

  int q = 0, slab = team, parity = 0;
  // closes the team's current slab: partials -> smem, team barrier, 128 * NT8 threads sum and store
  auto close_slab = [&]() {
    float* mine = red + parity * (team_warps * NT8 * 128);
#pragma unroll
    for (int t = 0; t < NT8; ++t) {
      *reinterpret_cast<float4*>(&mine[(wt * NT8 + t) * 128 + 4 * lane]) = make_float4(acc[t][0], acc[t][1], acc[t][2], acc[t][3]);
      acc[t][0] = acc[t][1] = acc[t][2] = acc[t][3] = 0.f;
    }
    named_bar_sync(bar_id, bar_threads);
    if (silu) {
      // gate_up tiles (awq_b200_repack_gate_up): rows 0..7 of the slab are gate rows, rows 8..15 the up rows of
      // the same 8 output columns -> y[tok, 8 slab + row] = silu(gate) * up, y is [M, N / 2]
      // (SiluAndMul after gate_up_proj: python/sglang/srt/models/llama.py:94-106, activation.py)
      const int NH = N >> 1;
      for (int e = tt; e < NT8 * 64; e += bar_threads) {
        const int t = e >> 6, L = (e & 63) >> 1, i = e & 1;
        const int row = L >> 2, tok = 8 * t + 2 * (L & 3) + i;
        const int col = slab * 8 + row;
        if (tok < M && col < NH) {
          float sg = 0.f, su = 0.f;
          for (int w = 0; w < team_warps; ++w) {  // warp order: deterministic
            sg += mine[(w * NT8 + t) * 128 + 4 * L + i];
            su += mine[(w * NT8 + t) * 128 + 4 * L + i + 2];
          }
          // the unfused sequence rounds gate_up's output to T before the activation: keep its bits
          const float gq = Elem<T>::to_float(Elem<T>::from_float(sg)), uq = Elem<T>::to_float(Elem<T>::from_float(su));
          y[int64_t(tok) * NH + col] = Elem<T>::from_float(gq / (1.f + __expf(-gq)) * uq);
        }
      }
    } else
    for (int e = tt; e < NT8 * 128; e += bar_threads) {
      const int t = e >> 7, idx = e & 127;
      // element i of lane L: i = 0: (row g, tok 2c)  1: (row g, tok 2c+1)  2: (row g+8, tok 2c)  3: (row g+8, tok 2c+1)
      const int L = idx >> 2, i = idx & 3;
      const int row = (L >> 2) + 8 * (i >> 1), tok = 8 * t + 2 * (L & 3) + (i & 1);
      const int n = slab * 16 + row;
      if constexpr (kPush) {
        // row-parallel layers (awq_b200_linear_allreduce): the all-reduce's packets leave from here (gemv_common.cuh)
        const bool valid = tok < M && n < N;
        float sum = 0.f;
        if (valid) {
          for (int w = 0; w < team_warps; ++w) sum += mine[(w * NT8 + t) * 128 + idx];
          if (bias != nullptr) sum += Elem<T>::to_float(bias[n]);
          y[int64_t(tok) * N + n] = Elem<T>::from_float(sum);
        }
        ar_push_pair<T>(push, tok, n, N, valid, sum);
      } else if (tok < M && n < N) {
        float sum = 0.f;
        for (int w = 0; w < team_warps; ++w) sum += mine[(w * NT8 + t) * 128 + idx];  // warp order: deterministic
        if (bias != nullptr) sum += Elem<T>::to_float(bias[n]);
        y[int64_t(tok) * N + n] = Elem<T>::from_float(sum);
      }
    }
    parity ^= 1;
    slab += T_total;
  };

  auto compute = [&](int s, uint32_t xoff, uint32_t sxoff) {
    uint4 xf[NT8][4];
#pragma unroll
    for (int t = 0; t < NT8; ++t)
#pragma unroll
      for (int sp = 0; sp < 4; ++sp) xf[t][sp] = lds128(xq[t] + xoff + uint32_t(64 * sp));
    if constexpr (kFast) {
      float2 nsx[NT8];
#pragma unroll
      for (int t = 0; t < NT8; ++t) {
        const uint2 v = lds64(sxq + sxoff + uint32_t(32 * t));
        nsx[t] = make_float2(__uint_as_float(v.x), __uint_as_float(v.y));
      }
      group_mma_subnormal<NT8>(wa[s], wb[s], s2[s], z2[s], xf, nsx, acc);
    } else {
      gemv_group_mma<T, NT8>(wa[s], wb[s], s2[s], z2[s], xf, acc);
    }
  };

  if constexpr (kEven) {
    // pw / pm address group 0 of the current slab; the requests of its first kDepth groups are out
    for (int j = 0; j < my_slabs; ++j) {
      const bool last = j + 1 == my_slabs;
      const uint8_t *cw = pw, *cm = pm;
      ps += T_total;
      set_slab_ptrs();  // pw / pm: next slab
      for (int i = 0; i < ng; i += kDepth) {
        const bool wrap = i + kDepth >= ng;
        const uint8_t* fw = wrap ? pw : cw + kDepth * kTileBytes;
        const uint8_t* fm = wrap ? pm : cm + kDepth * kTileBytes;
        const bool more = !(wrap && last);
#pragma unroll
        for (int s = 0; s < kDepth; ++s) {
          compute(s, uint32_t(s * M * kXBlock), uint32_t(s * NT8 * 32));
          if (more) {
            wa[s] = ldg128_stream(fw + s * kTileBytes);
            wb[s] = ldg128_stream(fw + s * kTileBytes + 512);
            s2[s] = ldg32_stream(fm + s * kTileBytes);
            z2[s] = ldg16_stream(fm + s * kTileBytes + dz);
          }
        }
        cw += kDepth * kTileBytes;
        cm += kDepth * kTileBytes;
#pragma unroll
        for (int t = 0; t < NT8; ++t) xq[t] += kDepth * M * kXBlock;
        sxq += kDepth * NT8 * 32;
      }
#pragma unroll
      for (int t = 0; t < NT8; ++t) xq[t] -= uint32_t(ng * M * kXBlock);
      sxq -= uint32_t(ng * NT8 * 32);
      close_slab();
    }
  } else {
    for (int it = 0; it < total; it += kDepth) {
#pragma unroll
      for (int s = 0; s < kDepth; ++s) {
        if (it + s < total) {
          compute(s, 0u, 0u);
          prefetch(s);
#pragma unroll
          for (int t = 0; t < NT8; ++t) xq[t] += M * kXBlock;
          sxq += NT8 * 32;
          if (++q == ng) {
            q = 0;
#pragma unroll
            for (int t = 0; t < NT8; ++t) xq[t] -= uint32_t(ng * M * kXBlock);
            sxq -= uint32_t(ng * NT8 * 32);
            close_slab();
          }
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
static int g_per_team = 0;     // 0 = heuristic, else forced team size (8, 16, 32)
static int g_per_variant = -1;  // -1 = default instantiation, else forced (debug)
void gemv_persist_debug(int team, int variant) {
  g_per_team = team;
  g_per_variant = variant;
}

// team size: 8 warps per slab unless N is so small that wider teams still give every slab its
// own team (then more warps pull on each slab and everything is in flight at once)
static int per_pick_team(int KT, int nslabs, int warps) {
  int team = 8;
  if (g_per_team == 8 || g_per_team == 16 || g_per_team == 32) {
    team = g_per_team;
  } else {
    const int sms = sm_count();
    while (team * 2 <= warps && team * 2 <= KT && nslabs <= sms * (warps / (team * 2))) team *= 2;
  }
  while (team > 8 && (team > warps || team > KT || warps % team != 0)) team >>= 1;
  return team;
}

template <typename T, int M, int kWarps, int kDepth, bool kEven, int kMinB, bool kPush>
static int gemv_persist_launch_e(const void* x, const void* tiles, const void* bias, void* y, int64_t K, int64_t N,
                               cudaStream_t st, int silu, const ArPush& push) {
  constexpr int NT8 = (M + 7) / 8;
  const int KT = int(K / kTileK);
  const int nslabs = int((N + 15) / 16);
  const int sms = sm_count();
  const int team = per_pick_team(KT, nslabs, kWarps);
  AWQ_CHECK(team <= KT && kWarps % team == 0, "awq_b200_gemv: K too small for the persistent decode kernel");
  const int lg_team = team == 8 ? 3 : (team == 16 ? 4 : 5);
  const int tpc = kWarps / team;
  const int grid = nslabs < sms * tpc ? (nslabs < sms ? nslabs : sms) : sms;
  const int smem = per_smem_bytes(KT, M, NT8, kWarps);
  AWQ_CHECK(smem <= 227 * 1024, "awq_b200_gemv: K * M too large for the persistent decode kernel (%d B)", smem);
  static int configured[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (smem > configured[dev]) {
    AWQ_CUDA_CHECK(cudaFuncSetAttribute(gemv_persist_kernel<T, M, kWarps, kDepth, kEven, kMinB, kPush>,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, smem),
                   "awq_b200_gemv: smem opt-in (persistent)");
    configured[dev] = smem;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(unsigned(grid));
  cfg.blockDim = dim3(32 * kWarps);
  cfg.dynamicSmemBytes = size_t(smem);
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  AWQ_CUDA_CHECK(cudaLaunchKernelEx(&cfg, gemv_persist_kernel<T, M, kWarps, kDepth, kEven, kMinB, kPush>, (const T*)x,
                                    (const uint8_t*)tiles, (const T*)bias, (T*)y, int(K), int(N), KT, lg_team, nslabs,
                                    (long long)l2_prefetch_bytes(int64_t((N + kTileN - 1) / kTileN) * KT * kTileBytes), silu, push),
                 "awq_b200_gemv: launch (persistent)");
  return 0;
}

template <typename T, int M, int kWarps, int kDepth, int kMinB = 1>
static int gemv_persist_launch(const void* x, const void* tiles, const void* bias, void* y, int64_t K, int64_t N,
                               cudaStream_t st, int silu, const ArPush& push) {
  const int KT = int(K / kTileK);
  const int team = per_pick_team(KT, int((N + 15) / 16), kWarps);
  const bool even = KT % (team * kDepth) == 0;
  if (push.peers != nullptr) {  // packet-push instantiations: the plain ones are untouched
    if (even) return gemv_persist_launch_e<T, M, kWarps, kDepth, true, kMinB, true>(x, tiles, bias, y, K, N, st, silu, push);
    return gemv_persist_launch_e<T, M, kWarps, kDepth, false, kMinB, true>(x, tiles, bias, y, K, N, st, silu, push);
  }
  if (even) return gemv_persist_launch_e<T, M, kWarps, kDepth, true, kMinB, false>(x, tiles, bias, y, K, N, st, silu, push);
  return gemv_persist_launch_e<T, M, kWarps, kDepth, false, kMinB, false>(x, tiles, bias, y, K, N, st, silu, push);
}

bool gemv_persist_fits(int64_t M, int64_t K) {
  if (M < 1 || M > 16 || K < 8 * kTileK) return false;
  return per_smem_bytes(int(K / kTileK), int(M), M <= 8 ? 1 : 2, M <= 8 ? 32 : 16) <= 200 * 1024;
}

template <typename T, int M>
static int gemv_persist_m(const void* x, const void* tiles, const void* bias, void* y, int64_t K, int64_t N,
                          cudaStream_t st, int silu, const ArPush& push) {
  if constexpr (M <= 8) {
    if constexpr (Elem<T>::kDtype == 1) {
      return gemv_persist_launch<T, M, 16, 4>(x, tiles, bias, y, K, N, st, silu, push);
    } else {
      switch (g_per_variant < 0 ? 2 : g_per_variant) {
        case 0: return gemv_persist_launch<T, M, 32, 2>(x, tiles, bias, y, K, N, st, silu, push);
        case 3: return gemv_persist_launch<T, M, 24, 3>(x, tiles, bias, y, K, N, st, silu, push);
        default: return gemv_persist_launch<T, M, 16, 4>(x, tiles, bias, y, K, N, st, silu, push);
      }
    }
  } else {
    if constexpr (Elem<T>::kDtype == 1) {
      return gemv_persist_launch<T, M, 16, 2>(x, tiles, bias, y, K, N, st, silu, push);
    } else {
      switch (g_per_variant < 0 ? 1 : g_per_variant) {
        case 0: return gemv_persist_launch<T, M, 16, 3>(x, tiles, bias, y, K, N, st, silu, push);
        default: return gemv_persist_launch<T, M, 16, 4>(x, tiles, bias, y, K, N, st, silu, push);
      }
    }
  }
}

template <typename T>
static int gemv_persist_t(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K,
                          int64_t N, cudaStream_t st, int silu, const ArPush& push) {
  switch (int(M)) {
#define AWQ_PER_CASE(MM) \
  case MM: return gemv_persist_m<T, MM>(x, tiles, bias, y, K, N, st, silu, push)
    AWQ_PER_CASE(1); AWQ_PER_CASE(2); AWQ_PER_CASE(3); AWQ_PER_CASE(4);
    AWQ_PER_CASE(5); AWQ_PER_CASE(6); AWQ_PER_CASE(7); AWQ_PER_CASE(8);
    AWQ_PER_CASE(9); AWQ_PER_CASE(10); AWQ_PER_CASE(11); AWQ_PER_CASE(12);
    AWQ_PER_CASE(13); AWQ_PER_CASE(14); AWQ_PER_CASE(15); AWQ_PER_CASE(16);
#undef AWQ_PER_CASE
    default: return set_error(1, "awq_b200_gemv: persistent kernel supports 1 <= M <= 16");
  }
}

int gemv_persist(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K, int64_t N,
                 int dtype, cudaStream_t st, int silu, const ArPush* push_or_null) {
  const ArPush none = {nullptr, 0, 0, 0, 0, 0, 0};
  const ArPush& push = push_or_null != nullptr ? *push_or_null : none;
  if (dtype == AWQ_B200_F16) return gemv_persist_t<__half>(x, tiles, bias, y, M, K, N, st, silu, push);
  return gemv_persist_t<__nv_bfloat16>(x, tiles, bias, y, M, K, N, st, silu, push);
}

}  // namespace awq
