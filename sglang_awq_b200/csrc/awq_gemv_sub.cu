// awq_gemv_sub.cu -- kernel 2e: persistent decode kernel for AWQ group sizes 64 and 32 (fp16),
// y[M,N] = x[M,K] @ W (+ bias) on the sub-group records of awq_b200_repack_g (8192 B of weights +
// SUB = 128 / G metadata sections of 384 B).
//
// Replaces the three launches of the reference's AWQLinearMethod.apply
// (python/sglang/srt/layers/quantization/awq.py:434-451; group size = K / scales.shape[0], any value the
// checkpoint's quant config names: awq_dequantize, sgl-kernel/csrc/gemm/awq_kernel.cu:186-221) for decode batches of
// group-64 / group-32 checkpoints, which round 2 served through the tcgen05 GEMM kernel at 0.30-0.53 of HBM.
//
// It is the persistent team kernel (awq_gemv_persist.cu: one CTA per SM, x staged once, teams of warps own 16-row
// slabs, one continuous register-direct LDG.128 stream per warp) with one change of the MATH.  The fp16 fast path
// multiplies the masked nibbles as subnormal fp16 operands and applies scale and zero point to the fp32 sum of a
// whole scale group -- but an m16n8k16 MMA sums over the four lanes of a quad, and a lane's 16-byte granule is 32
// consecutive k: every MMA of a record spans all 128 k, i.e. 2 or 4 scale groups.  Instead of transposing the
// weights (shuffles) or issuing one MMA per sub-group (2-4x the tensor-pipe work), the sub-group goes into the
// TOKEN dimension, which decode leaves mostly empty: token column j of the MMA's 8 (or 16) columns carries
// token j % MP of sub-group j / MP (MP = columns / SUB), and the B operand (x) of column j is ZERO for the k
// outside its sub-group -- the lanes of a quad whose granule lies outside read a block of zeros in shared memory
// (an address select, no extra instruction in the stream).  Column j's fp32 sum is then the partial dot product of
// ONE scale group: scale, zero point (-z * sum of x over the sub-group, staged per column) are applied per column
// exactly as in the group-128 kernel, and the SUB columns of a token are added when the team closes the slab.
// Same instruction count per record as the group-128 kernel plus one metadata load per extra 8-column block.
//   G = 64: M <= 4,   G = 32: M <= 2   (8 token columns = SUB sub-groups x MP tokens).
// 16 columns (the kernel is written for them: NT8 = 2) double the MMAs and the shared-memory reads per record: 74 us
// against 50 us on Llama-3-70B gate_up at G = 32, M = 4, and 87 us at G = 64, M = 8 (taking the lane's x once and
// masking it per block with AND: 77 us, the loop is issue-bound).  The per-weight variant below does those batches in
// 68-69 us, so it takes every batch up to 8 that the token columns do not hold, and bf16; larger batches stay on the
// tcgen05 GEMM kernel (awq_gemm_tcgen05.cu, 80 us).
#include "common.cuh"
#include "gemv_common.cuh"
#include "host_common.h"

namespace awq {

constexpr int kSubZeroBlock = 256;  // bytes of zeros the masked B lanes read

__host__ __device__ constexpr int sub_zero_off(int KT, int M) { return (KT * M * kXBlock + 127) & ~127; }
__host__ __device__ constexpr int sub_sx_off(int KT, int M) { return sub_zero_off(KT, M) + kSubZeroBlock; }
__host__ __device__ constexpr int sub_red_off(int KT, int M, int nt8) { return sub_sx_off(KT, M) + ((KT * nt8 * 32 + 127) & ~127); }
__host__ __device__ constexpr int sub_smem_bytes(int KT, int M, int nt8, int warps) {
  return sub_red_off(KT, M, nt8) + 2 * warps * nt8 * 512;
}
__host__ __device__ constexpr int sub_nt8(int M, int SUB) { return SUB * M <= 8 ? 1 : 2; }

// group_mma_subnormal (gemv_common.cuh) with scale / zero point per 8-column block: block t's two columns of this
// lane belong to one sub-group, whose {scale g, scale g+8} / {zero g, zero g+8} are s2[t] / z2[t]
template <int NT8>
__device__ __forceinline__ void group_mma_subnormal_cols(const uint4& wa4, const uint4& wb4, const uint32_t* s2, const uint32_t* z2,
                                                         const uint4 (*xf)[4], const float2* nsx, float (*acc)[4]) {
  const uint32_t wa[4] = {wa4.x, wa4.y, wa4.z, wa4.w};
  const uint32_t wb[4] = {wb4.x, wb4.y, wb4.z, wb4.w};
  float ge[NT8][4], go[NT8][4];
#pragma unroll
  for (int t = 0; t < NT8; ++t)
#pragma unroll
    for (int i = 0; i < 4; ++i) ge[t][i] = go[t][i] = 0.f;
#pragma unroll
  for (int sp = 0; sp < 4; ++sp) {
    const uint32_t ta = wa[sp] >> 8, tb = wb[sp] >> 8;
    uint32_t ae[4], ao[4];
    ae[0] = wa[sp] & 0x000f000fu;
    ae[1] = wb[sp] & 0x000f000fu;
    ae[2] = ta & 0x000f000fu;
    ae[3] = tb & 0x000f000fu;
    ao[0] = wa[sp] & 0x00f000f0u;
    ao[1] = wb[sp] & 0x00f000f0u;
    ao[2] = ta & 0x00f000f0u;
    ao[3] = tb & 0x00f000f0u;
#pragma unroll
    for (int t = 0; t < NT8; ++t) {
      mma_16816<__half>(ge[t], ae, xf[t][sp].x, xf[t][sp].y);
      mma_16816<__half>(go[t], ao, xf[t][sp].z, xf[t][sp].w);
    }
  }
#pragma unroll
  for (int t = 0; t < NT8; ++t) {
    const float sa = bits16_to_float<__half>(s2[t] & 0xFFFFu) * 16777216.f, sb = bits16_to_float<__half>(s2[t] >> 16) * 16777216.f;
    const float za = float(z2[t] & 0xFFu), zb = float(z2[t] >> 8);
    float u0 = fmaf(go[t][0], 0.0625f, ge[t][0]);
    float u1 = fmaf(go[t][1], 0.0625f, ge[t][1]);
    float u2 = fmaf(go[t][2], 0.0625f, ge[t][2]);
    float u3 = fmaf(go[t][3], 0.0625f, ge[t][3]);
    u0 = fmaf(za, nsx[t].x, u0);
    u1 = fmaf(za, nsx[t].y, u1);
    u2 = fmaf(zb, nsx[t].x, u2);
    u3 = fmaf(zb, nsx[t].y, u3);
    acc[t][0] = fmaf(sa, u0, acc[t][0]);
    acc[t][1] = fmaf(sa, u1, acc[t][1]);
    acc[t][2] = fmaf(sb, u2, acc[t][2]);
    acc[t][3] = fmaf(sb, u3, acc[t][3]);
  }
}

// kWarps warps per CTA (1 CTA per SM), teams of `team` warps (8, 16 or 32; run time), kDepth records in flight per
// warp.  kEven: every warp has the same number of records per slab and it is a multiple of kDepth (host-checked).
template <int M, int SUB, int kWarps, int kDepth, bool kEven>
__global__ void __launch_bounds__(32 * kWarps, 1)
gemv_sub_kernel(const __half* __restrict__ x, const uint8_t* __restrict__ tiles, const __half* __restrict__ bias,
                __half* __restrict__ y, int K, int N, int KT, int lg_team, int nslabs) {
  using T = __half;
  constexpr int NT8 = sub_nt8(M, SUB);
  constexpr int MP = 8 * NT8 / SUB;  // token slots per sub-group
  constexpr int kRec = tile_bytes_sub(SUB);
  static_assert(M <= MP, "batch does not fit the token columns");
  extern __shared__ __align__(128) uint8_t smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, c = lane & 3;
  const int team_warps = 1 << lg_team;
  const int teams_per_cta = kWarps >> lg_team;
  const int team_in_cta = warp >> lg_team, wt = warp & (team_warps - 1);
  const int T_total = int(gridDim.x) * teams_per_cta;
  const int team = team_in_cta * int(gridDim.x) + int(blockIdx.x);  // SM-interleaved numbering
  const int kb = (wt * KT) >> lg_team, ke = ((wt + 1) * KT) >> lg_team;
  const int ng = ke - kb;
  const int my_slabs = team < nslabs ? (nslabs - 1 - team) / T_total + 1 : 0;
  const int total = my_slabs * ng;  // records in this warp's stream

  // ---- weight / metadata stream (addresses as in gemv_persist_kernel; records are kRec bytes) ----
  int pq = 0, pleft = total, ps = team;
  const int lane_w = g * 64 + 16 * (c ^ ((g >> 1) & 3)), lane_m = kTileMetaOff + 4 * g;
  const int dz = 32 - 2 * g;
  // the sub-group of this lane's two C columns in block t: columns 8t + 2c, 8t + 2c + 1
  int sec[NT8];
#pragma unroll
  for (int t = 0; t < NT8; ++t) sec[t] = ((8 * t + 2 * c) / MP) * kTileMetaSection;
  const uint8_t *pw, *pm;
  auto set_slab_ptrs = [&]() {
    const uint8_t* rec = tiles + (int64_t(ps >> 3) * KT + kb) * kRec;
    pw = rec + ((ps & 7) * 1024 + lane_w);
    pm = rec + ((ps & 7) * kTileMetaBlock + lane_m);
  };
  set_slab_ptrs();

  uint4 wa[kDepth], wb[kDepth];
  uint32_t s2[kDepth][NT8], z2[kDepth][NT8];
  auto load = [&](int s, const uint8_t* w, const uint8_t* m) {
    wa[s] = ldg128_stream(w);
    wb[s] = ldg128_stream(w + 512);
#pragma unroll
    for (int t = 0; t < NT8; ++t) {
      s2[s][t] = ldg32_stream(m + sec[t]);
      z2[s][t] = ldg16_stream(m + sec[t] + dz);
    }
  };
  auto prefetch = [&](int s) {
    if (pleft > 0) {
      load(s, pw, pm);
      pw += kRec;
      pm += kRec;
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
      for (int s = 0; s < kDepth; ++s) load(s, pw + s * kRec, pm + s * kRec);
    }
  } else {
#pragma unroll
    for (int s = 0; s < kDepth; ++s) prefetch(s);
  }

  pdl_launch_dependents();
  pdl_wait();  // x and y belong to the previous kernel until here

  // ---- stage all of x, the block of zeros and the per-column sub-group sums once per SM ----
  const uint32_t xs = smem_u32(smem);
  const uint32_t zs = xs + uint32_t(sub_zero_off(KT, M));
  const uint32_t sxs = xs + uint32_t(sub_sx_off(KT, M));
  if (tid < kSubZeroBlock / 4) sts32(zs + 4u * uint32_t(tid), 0u);
  for (int i = tid; i < KT * NT8 * 8; i += 32 * kWarps) sts32(sxs + 4u * uint32_t(i), 0u);  // columns of absent tokens
  __syncthreads();
  stage_x<T, M, NT8 * 8, true, 4, SUB, MP>(x, K, 0, KT, xs, sxs, warp, kWarps, lane);
  __syncthreads();

  // ---- B operand addresses: lane (g, c) feeds column 8t + g = (sub-group, token) with the x of ITS granule
  // (k 32c .. 32c + 31) when that lies in the column's sub-group, else with zeros ----
  constexpr int kGranulesPerSub = 4 / SUB;  // 2 (G = 64) or 1 (G = 32)
  uint32_t xq[NT8], xmask[NT8];
#pragma unroll
  for (int t = 0; t < NT8; ++t) {
    const int col = 8 * t + g;
    const int csub = col / MP, tok = col % MP;
    const bool live = (c / kGranulesPerSub) == csub;
    xmask[t] = live ? 0xffffffffu : 0u;
    xq[t] = live ? xs + uint32_t(((kb * M) + (tok < M ? tok : M - 1)) * kXBlock + 16 * c) : zs + uint32_t(16 * c);
  }
  uint32_t sxq = sxs + uint32_t(kb * NT8 * 32) + 8u * uint32_t(c);
  float* red = reinterpret_cast<float*>(smem + sub_red_off(KT, M, NT8)) + team_in_cta * (2 * team_warps * NT8 * 128);
  const int bar_id = 1 + team_in_cta, bar_threads = 32 * team_warps;
  const int tt = tid & (bar_threads - 1);  // thread index inside the team

  float acc[NT8][4];
#pragma unroll
  for (int t = 0; t < NT8; ++t)
#pragma unroll
    for (int i = 0; i < 4; ++i) acc[t][i] = 0.f;

  int q = 0, slab = team, parity = 0;
  // closes the team's current slab: per-column partials -> smem, team barrier, 16 * M threads add the SUB columns of
  // their (row, token) over the team's warps (fixed order: deterministic) and store
  auto close_slab = [&]() {
    float* mine = red + parity * (team_warps * NT8 * 128);
#pragma unroll
    for (int t = 0; t < NT8; ++t) {
      *reinterpret_cast<float4*>(&mine[(wt * NT8 + t) * 128 + 4 * lane]) = make_float4(acc[t][0], acc[t][1], acc[t][2], acc[t][3]);
      acc[t][0] = acc[t][1] = acc[t][2] = acc[t][3] = 0.f;
    }
    named_bar_sync(bar_id, bar_threads);
    for (int e = tt; e < 16 * M; e += bar_threads) {
      const int tok = e >> 4, row = e & 15;
      const int n = slab * 16 + row;
      if (n < N) {
        float sum = 0.f;
#pragma unroll
        for (int sb = 0; sb < SUB; ++sb) {
          // element i of lane L: i = 0: (row g, col 2c)  1: (row g, col 2c+1)  2: (row g+8, col 2c)  3: (row g+8, col 2c+1)
          const int col = sb * MP + tok;
          const int t = col >> 3, L = (row & 7) * 4 + ((col & 7) >> 1), i = (col & 1) + 2 * (row >> 3);
          for (int w = 0; w < team_warps; ++w) sum += mine[(w * NT8 + t) * 128 + 4 * L + i];
        }
        if (bias != nullptr) sum += __half2float(bias[n]);
        y[int64_t(tok) * N + n] = __float2half_rn(sum);
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
      for (int sp = 0; sp < 4; ++sp) xf[t][sp] = lds128(xq[t] + (xoff & xmask[t]) + uint32_t(64 * sp));
    float2 nsx[NT8];
#pragma unroll
    for (int t = 0; t < NT8; ++t) {
      const uint2 v = lds64(sxq + sxoff + uint32_t(32 * t));
      nsx[t] = make_float2(__uint_as_float(v.x), __uint_as_float(v.y));
    }
    group_mma_subnormal_cols<NT8>(wa[s], wb[s], s2[s], z2[s], xf, nsx, acc);
  };
  auto advance_x = [&](uint32_t blocks) {  // the zero block of a masked lane stays where it is
#pragma unroll
    for (int t = 0; t < NT8; ++t) xq[t] += (blocks * uint32_t(M * kXBlock)) & xmask[t];
    sxq += blocks * uint32_t(NT8 * 32);
  };
  auto rewind_x = [&]() {
#pragma unroll
    for (int t = 0; t < NT8; ++t) xq[t] -= uint32_t(ng * M * kXBlock) & xmask[t];
    sxq -= uint32_t(ng * NT8 * 32);
  };

  if constexpr (kEven) {
    for (int j = 0; j < my_slabs; ++j) {
      const bool last = j + 1 == my_slabs;
      const uint8_t *cw = pw, *cm = pm;
      ps += T_total;
      set_slab_ptrs();  // pw / pm: next slab
      for (int i = 0; i < ng; i += kDepth) {
        const bool wrap = i + kDepth >= ng;
        const uint8_t* fw = wrap ? pw : cw + kDepth * kRec;
        const uint8_t* fm = wrap ? pm : cm + kDepth * kRec;
        const bool more = !(wrap && last);
#pragma unroll
        for (int s = 0; s < kDepth; ++s) {
          compute(s, uint32_t(s * M * kXBlock), uint32_t(s * NT8 * 32));
          if (more) load(s, fw + s * kRec, fm + s * kRec);
        }
        cw += kDepth * kRec;
        cm += kDepth * kRec;
        advance_x(kDepth);
      }
      rewind_x();
      close_slab();
    }
  } else {
    for (int it = 0; it < total; it += kDepth) {
#pragma unroll
      for (int s = 0; s < kDepth; ++s) {
        if (it + s < total) {
          compute(s, 0u, 0u);
          prefetch(s);
          advance_x(1);
          if (++q == ng) {
            q = 0;
            rewind_x();
            close_slab();
          }
        }
      }
    }
  }
}

// ---- per-weight variant: bf16 at any decode batch up to 8, and fp16 for the batches the token-column trick does
// not take (M = 5 .. 8).  Every weight is dequantised as the reference does it, (w - z) * s rounded once to T
// (gemv_group_mma with the scale per weight), so the scale group is simply the lane's own: lane c of a quad holds
// granule c = k 32c .. 32c + 31 and reads the (scale, zero) pair of section c / (4 / SUB).  Nothing else differs from
// the bf16 path of gemv_persist_kernel; about 2x the ALU work per record of the subnormal fast path. ----
__host__ __device__ constexpr int subpw_red_off(int KT, int M) { return (KT * M * kXBlock + 127) & ~127; }
__host__ __device__ constexpr int subpw_smem_bytes(int KT, int M, int warps) { return subpw_red_off(KT, M) + 2 * warps * 512; }

template <typename T, int M, int SUB, int kWarps, int kDepth>
__global__ void __launch_bounds__(32 * kWarps, 1)
gemv_sub_pw_kernel(const T* __restrict__ x, const uint8_t* __restrict__ tiles, const T* __restrict__ bias, T* __restrict__ y,
                   int K, int N, int KT, int lg_team, int nslabs) {
  static_assert(M >= 1 && M <= 8, "one block of 8 token columns");
  constexpr int kRec = tile_bytes_sub(SUB);
  constexpr int kGranulesPerSub = 4 / SUB;
  extern __shared__ __align__(128) uint8_t smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, c = lane & 3;
  const int team_warps = 1 << lg_team;
  const int teams_per_cta = kWarps >> lg_team;
  const int team_in_cta = warp >> lg_team, wt = warp & (team_warps - 1);
  const int T_total = int(gridDim.x) * teams_per_cta;
  const int team = team_in_cta * int(gridDim.x) + int(blockIdx.x);
  const int kb = (wt * KT) >> lg_team, ke = ((wt + 1) * KT) >> lg_team;
  const int ng = ke - kb;
  const int my_slabs = team < nslabs ? (nslabs - 1 - team) / T_total + 1 : 0;
  const int total = my_slabs * ng;

  int pq = 0, pleft = total, ps = team;
  const int lane_w = g * 64 + 16 * (c ^ ((g >> 1) & 3));
  const int lane_m = kTileMetaOff + (c / kGranulesPerSub) * kTileMetaSection + 4 * g;  // the section of this lane's granule
  const int dz = 32 - 2 * g;
  const uint8_t *pw, *pm;
  auto set_slab_ptrs = [&]() {
    const uint8_t* rec = tiles + (int64_t(ps >> 3) * KT + kb) * kRec;
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
      pw += kRec;
      pm += kRec;
      --pleft;
      if (++pq == ng) {
        pq = 0;
        ps += T_total;
        set_slab_ptrs();
      }
    }
  };
#pragma unroll
  for (int s = 0; s < kDepth; ++s) prefetch(s);

  pdl_launch_dependents();
  pdl_wait();  // x and y belong to the previous kernel until here

  const uint32_t xs = smem_u32(smem);
  stage_x<T, M, 8, false>(x, K, 0, KT, xs, 0u, warp, kWarps, lane);
  __syncthreads();

  uint32_t xq = xs + uint32_t(((kb * M) + (g < M ? g : M - 1)) * kXBlock + 16 * c);
  float* red = reinterpret_cast<float*>(smem + subpw_red_off(KT, M)) + team_in_cta * (2 * team_warps * 128);
  const int bar_id = 1 + team_in_cta, bar_threads = 32 * team_warps;
  const int tt = tid & (bar_threads - 1);
  float acc[1][4] = {{0.f, 0.f, 0.f, 0.f}};
  int q = 0, slab = team, parity = 0;

  for (int it = 0; it < total; it += kDepth) {
#pragma unroll
    for (int s = 0; s < kDepth; ++s) {
      if (it + s < total) {
        uint4 xf[1][4];
#pragma unroll
        for (int sp = 0; sp < 4; ++sp) xf[0][sp] = lds128(xq + uint32_t(64 * sp));
        gemv_group_mma<T, 1, true>(wa[s], wb[s], s2[s], z2[s], xf, acc);
        prefetch(s);
        xq += M * kXBlock;
        if (++q == ng) {  // close the team's slab: partials -> smem, team barrier, 128 threads sum in warp order and store
          q = 0;
          xq -= uint32_t(ng * M * kXBlock);
          float* mine = red + parity * (team_warps * 128);
          *reinterpret_cast<float4*>(&mine[wt * 128 + 4 * lane]) = make_float4(acc[0][0], acc[0][1], acc[0][2], acc[0][3]);
          acc[0][0] = acc[0][1] = acc[0][2] = acc[0][3] = 0.f;
          named_bar_sync(bar_id, bar_threads);
          if (tt < 128) {
            const int L = tt >> 2, i = tt & 3;
            const int row = (L >> 2) + 8 * (i >> 1), tok = 2 * (L & 3) + (i & 1);
            const int n = slab * 16 + row;
            if (tok < M && n < N) {
              float sum = 0.f;
              for (int w = 0; w < team_warps; ++w) sum += mine[w * 128 + tt];
              if (bias != nullptr) sum += Elem<T>::to_float(bias[n]);
              y[int64_t(tok) * N + n] = Elem<T>::from_float(sum);
            }
          }
          parity ^= 1;
          slab += T_total;
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
static int sub_pick_team(int KT, int nslabs, int warps) {
  int team = 8;
  const int sms = sm_count();
  while (team * 2 <= warps && team * 2 <= KT && nslabs <= sms * (warps / (team * 2))) team *= 2;
  while (team > 8 && (team > warps || team > KT || warps % team != 0)) team >>= 1;
  return team;
}

constexpr int kSubWarps = 16, kSubDepth = 4;

template <int M, int SUB, bool kEven>
static int gemv_sub_launch_e(const void* x, const void* tiles, const void* bias, void* y, int64_t K, int64_t N, int team,
                             cudaStream_t st) {
  constexpr int NT8 = sub_nt8(M, SUB);
  const int KT = int(K / kTileK);
  const int nslabs = int((N + 15) / 16);
  const int sms = sm_count();
  const int lg_team = team == 8 ? 3 : (team == 16 ? 4 : 5);
  const int tpc = kSubWarps / team;
  const int grid = nslabs < sms * tpc ? (nslabs < sms ? nslabs : sms) : sms;
  const int smem = sub_smem_bytes(KT, M, NT8, kSubWarps);
  AWQ_CHECK(smem <= 227 * 1024, "awq_b200_linear_g: K * M too large for the sub-group decode kernel (%d B)", smem);
  auto kern = gemv_sub_kernel<M, SUB, kSubWarps, kSubDepth, kEven>;
  static int configured[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (smem > configured[dev]) {
    AWQ_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem),
                   "awq_b200_linear_g: smem opt-in (sub-group decode)");
    configured[dev] = smem;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(unsigned(grid));
  cfg.blockDim = dim3(32 * kSubWarps);
  cfg.dynamicSmemBytes = size_t(smem);
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  AWQ_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, (const __half*)x, (const uint8_t*)tiles, (const __half*)bias, (__half*)y,
                                    int(K), int(N), KT, lg_team, nslabs),
                 "awq_b200_linear_g: launch (sub-group decode)");
  return 0;
}

template <int M, int SUB>
static int gemv_sub_launch(const void* x, const void* tiles, const void* bias, void* y, int64_t K, int64_t N, cudaStream_t st) {
  const int KT = int(K / kTileK);
  const int team = sub_pick_team(KT, int((N + 15) / 16), kSubWarps);
  AWQ_CHECK(team <= KT && kSubWarps % team == 0, "awq_b200_linear_g: K too small for the sub-group decode kernel");
  if (KT % (team * kSubDepth) == 0) return gemv_sub_launch_e<M, SUB, true>(x, tiles, bias, y, K, N, team, st);
  return gemv_sub_launch_e<M, SUB, false>(x, tiles, bias, y, K, N, team, st);
}

template <typename T, int M, int SUB>
static int gemv_sub_pw_launch(const void* x, const void* tiles, const void* bias, void* y, int64_t K, int64_t N, cudaStream_t st) {
  const int KT = int(K / kTileK);
  const int nslabs = int((N + 15) / 16);
  const int sms = sm_count();
  const int team = sub_pick_team(KT, nslabs, kSubWarps);
  AWQ_CHECK(team <= KT && kSubWarps % team == 0, "awq_b200_linear_g: K too small for the sub-group decode kernel");
  const int lg_team = team == 8 ? 3 : (team == 16 ? 4 : 5);
  const int tpc = kSubWarps / team;
  const int grid = nslabs < sms * tpc ? (nslabs < sms ? nslabs : sms) : sms;
  const int smem = subpw_smem_bytes(KT, M, kSubWarps);
  AWQ_CHECK(smem <= 227 * 1024, "awq_b200_linear_g: K * M too large for the sub-group decode kernel (%d B)", smem);
  constexpr int kDepth = 4;
  auto kern = gemv_sub_pw_kernel<T, M, SUB, kSubWarps, kDepth>;
  static int configured[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (smem > configured[dev]) {
    AWQ_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem),
                   "awq_b200_linear_g: smem opt-in (sub-group decode, per weight)");
    configured[dev] = smem;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(unsigned(grid));
  cfg.blockDim = dim3(32 * kSubWarps);
  cfg.dynamicSmemBytes = size_t(smem);
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  AWQ_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, (const T*)x, (const uint8_t*)tiles, (const T*)bias, (T*)y, int(K), int(N), KT,
                                    lg_team, nslabs),
                 "awq_b200_linear_g: launch (sub-group decode, per weight)");
  return 0;
}

template <typename T, int SUB>
static int gemv_sub_pw_m(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K, int64_t N, cudaStream_t st) {
  switch (int(M)) {
#define AWQ_SUBPW_CASE(MM) \
  case MM: return gemv_sub_pw_launch<T, MM, SUB>(x, tiles, bias, y, K, N, st)
    AWQ_SUBPW_CASE(1); AWQ_SUBPW_CASE(2); AWQ_SUBPW_CASE(3); AWQ_SUBPW_CASE(4);
    AWQ_SUBPW_CASE(5); AWQ_SUBPW_CASE(6); AWQ_SUBPW_CASE(7); AWQ_SUBPW_CASE(8);
#undef AWQ_SUBPW_CASE
    default: return set_error(1, "awq_b200_linear_g: batch %lld does not fit the sub-group decode kernel", (long long)M);
  }
}

static int g_sub_force_pw = 0;  // debug: 1 sends every problem the per-weight kernel can take to it
void gemv_sub_debug(int force_pw) { g_sub_force_pw = force_pw; }

// token-column kernel: fp16, SUB * M <= 8 columns
static bool sub_cols_fits(int64_t M, int64_t K, int SUB, int dtype) {
  if (dtype != AWQ_B200_F16 || SUB * M > 8 || g_sub_force_pw) return false;
  return sub_smem_bytes(int(K / kTileK), int(M), sub_nt8(int(M), SUB), kSubWarps) <= 200 * 1024;
}
// per-weight kernel: fp16 / bf16, M <= 8
static bool sub_pw_fits(int64_t M, int64_t K) { return M <= 8 && subpw_smem_bytes(int(K / kTileK), int(M), kSubWarps) <= 200 * 1024; }

// group 64 or 32, K >= 1024, activations within one SM's shared memory: fp16 with SUB * M <= 8 (token columns), else M <= 8 (per weight)
bool gemv_sub_fits(int64_t M, int64_t K, int64_t group, int dtype) {
  if ((dtype != AWQ_B200_F16 && dtype != AWQ_B200_BF16) || (group != 64 && group != 32)) return false;
  const int SUB = int(kTileK / group);
  if (M < 1 || K < 8 * kTileK || K % kTileK != 0) return false;
  return sub_cols_fits(M, K, SUB, dtype) || sub_pw_fits(M, K);
}

int gemv_sub(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K, int64_t N, int64_t group,
             int dtype, cudaStream_t st) {
  AWQ_CHECK(gemv_sub_fits(M, K, group, dtype), "awq_b200_linear_g: the sub-group decode kernel does not take this problem");
  if (!sub_cols_fits(M, K, int(kTileK / group), dtype)) {
    if (dtype == AWQ_B200_F16)
      return group == 64 ? gemv_sub_pw_m<__half, 2>(x, tiles, bias, y, M, K, N, st) : gemv_sub_pw_m<__half, 4>(x, tiles, bias, y, M, K, N, st);
    return group == 64 ? gemv_sub_pw_m<__nv_bfloat16, 2>(x, tiles, bias, y, M, K, N, st)
                       : gemv_sub_pw_m<__nv_bfloat16, 4>(x, tiles, bias, y, M, K, N, st);
  }
#define AWQ_SUB_CASE(MM, SS) \
  case MM: return gemv_sub_launch<MM, SS>(x, tiles, bias, y, K, N, st)
  if (group == 64) {
    switch (int(M)) {
      AWQ_SUB_CASE(1, 2); AWQ_SUB_CASE(2, 2); AWQ_SUB_CASE(3, 2); AWQ_SUB_CASE(4, 2);
      default: break;
    }
  } else {
    switch (int(M)) {
      AWQ_SUB_CASE(1, 4); AWQ_SUB_CASE(2, 4);
      default: break;
    }
  }
#undef AWQ_SUB_CASE
  return set_error(1, "awq_b200_linear_g: batch %lld does not fit the sub-group decode kernel", (long long)M);
}

}  // namespace awq
