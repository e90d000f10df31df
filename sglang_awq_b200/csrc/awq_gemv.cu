// awq_gemv.cu -- kernel 2: decode path  y[M,N] = x[M,K] @ W (+ bias),  1 <= M <= 64.
//
// Replaces, for small M, the three launches of the reference's AWQLinearMethod.apply
// (python/sglang/srt/layers/quantization/awq.py:434-451: awq_dequantize -> torch.matmul ->
// add_), which move 4.5 B per weight through HBM.  This kernel moves the algorithmic 0.53 B.
//
// The op is HBM-bound; the design is about keeping HBM saturated (measured on B200, see
// ~14 MB must be in flight chip-wide, and 1-D bulk-TMA copies cost ~27 cycles +
// bytes/22 B/clk per request per SM, so a ring fed by several small bulk copies per stage
// tops out near 3-4 TB/s while plain LSU traffic reaches 7 TB/s):
//   * warp-autonomous streaming: each of the 8 warps of a CTA owns 16 output rows and runs
//     its OWN cp.async (LDGSTS, LSU path) ring of 8 x 1072 B stages straight out of the tile
//     records -- no producer warp, no mbarriers, no block-level barrier anywhere in the
//     kernel.  3 CTAs/SM x 8 warps x 7 KB = 168 KB in flight per SM;
//   * stream-K work split: the (n-tile, k-tile) sequence IS the memory order of the tile
//     records, so CTA i simply streams the contiguous byte range of its Ttot/G tiles; every
//     shape is balanced to +-1 tile over all resident CTAs (no wave quantisation, no tail);
//   * activations come straight from L1/L2 as 128-bit loads: the tile layout packs 8
//     consecutive k per word, so a lane's B fragments for two MMA k-steps are one aligned
//     16-byte piece of an x row -- no shared-memory staging, no ldmatrix;
//   * int4 -> exact (w - z) fp16/bf16 register fragments with 9 ALU ops per 8 weights, reduced
//     over k with mma.sync.m16n8k16 (fp32 accumulate, two independent chains).  CUDA-core FMAs
//     cannot do this at HBM rate (>= 25 issue slots per 8 weights at M = 1 against a budget of
//     ~22); the group scale is applied in fp32 once per (row, group);
//   * programmatic dependent launch: the ring prologue (weights only) is issued before
//     griddepcontrol.wait, so layer i+1's HBM stream starts while layer i drains;
//   * n-tiles that straddle CTAs are closed through an L2-resident fp32 workspace: each warp
//     publishes its 16-row partial (coalesced float4), the last arriver per (n-tile, warp)
//     semaphore sums the partials in RANK order -> deterministic, no float atomics.
#include "common.cuh"
#include "host_common.h"
#include "gemv_route_table.h"

#include <cmath>

namespace awq {

constexpr int kGemvWarps = 8;
constexpr int kGemvThreads = 32 * kGemvWarps;
// cp.async stages per warp (measured: 6 stages / 3 CTAs per SM is no faster than 8 / 2 for the shared-x variant)
__host__ __device__ constexpr int gemv_ring(int nt8, bool shared_x) { return 8; }
constexpr int kSliceW = 1024;          // 16 rows x 64 B of packed weights
constexpr int kSliceBytes = kSliceW + kTileMetaBlock;  // + 48 B metadata block = 1072 = 67 x 16
constexpr int kSemInts = 8192;         // semaphores: n_tile * 8 + warp

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

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
// predicated variant: no branch around the copy
__device__ __forceinline__ void cp_async16_if(bool pred, uint32_t dst, const void* src) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %0, 0;\n\t@p cp.async.cg.shared.global [%1], [%2], 16;\n\t}" ::"r"(
          uint32_t(pred)),
      "r"(dst), "l"(src)
      : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ uint4 ldg128(const void* p) {
  uint4 v;
  asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}

template <typename T>
__device__ __forceinline__ float bits16_to_float(uint32_t b) {
  const uint16_t h = uint16_t(b);
  return Elem<T>::to_float(*reinterpret_cast<const T*>(&h));
}

// stream-K geometry shared by host and device: CTA i owns tiles [tile_begin(i), tile_begin(i+1))
// (32-bit math: the host guarantees ttot * grid < 2^31; 64-bit division is ~100 instructions)
__host__ __device__ __forceinline__ int sk_begin(int i, int ttot, int grid) {
  return int(unsigned(i) * unsigned(ttot) / unsigned(grid));
}

constexpr int kXPitch = 288;  // shared-x row pitch: 256 B of k + 32 B, conflict-free with the chunk swizzle
__host__ __device__ constexpr int gemv_smem_bytes(int nt8, bool shared_x) {
  return gemv_ring(nt8, shared_x) * (kGemvWarps * kSliceBytes + (shared_x ? 8 * nt8 * kXPitch : 0));
}
__host__ __device__ constexpr int gemv_min_ctas(int nt8, bool shared_x) {
  return (nt8 == 1 && !shared_x) ? 3 : (nt8 <= 4 ? 2 : 1);
}

// kSharedX = false (M <= 2): every warp reads its activation fragments straight from L1/L2.
// kSharedX = true  (M  > 2): the CTA's 8 warps share one swizzled smem copy of each x k-tile
//   (one cooperative cp.async per thread per tile, one CTA barrier per tile); without it x is
//   re-read 8x per CTA through a ~20 KB L1 and the kernel becomes L2-bound from M = 4 up.
template <typename T, int NT8, bool kSharedX>
__global__ void __launch_bounds__(kGemvThreads, gemv_min_ctas(NT8, kSharedX))
gemv_kernel(const T* __restrict__ x, const uint8_t* __restrict__ tiles, const T* __restrict__ bias,
            T* __restrict__ y, float* __restrict__ ws_partial, int* __restrict__ ws_sem, int M, int K, int N,
            int KT, int ttot, int maxc) {
  constexpr int kRing = gemv_ring(NT8, kSharedX);
  extern __shared__ __align__(128) uint8_t smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, c = lane & 3;
  const int grid = gridDim.x;
  const int t_begin = sk_begin(blockIdx.x, ttot, grid);
  const int t_end = sk_begin(blockIdx.x + 1, ttot, grid);
  const int nk = t_end - t_begin;

  // running cp.async state: source pointers advance one record per tile, the destination
  // walks the warp's ring (all 32-bit adds; no per-tile multiplies)
  const uint32_t ring = smem_u32(smem) + uint32_t(warp) * (kRing * kSliceBytes);
  const uint32_t ring_end = ring + kRing * kSliceBytes;
  const uint8_t* src_w = tiles + int64_t(t_begin) * kTileBytes + warp * kSliceW + 16 * lane;
  const uint8_t* src_m = tiles + int64_t(t_begin) * kTileBytes + kTileMetaOff + warp * kTileMetaBlock + 16 * (lane % 3);
  uint32_t st_issue = ring;
  int n_issue = nk;  // tiles still to request
  const bool meta_lane = lane < 3;

  // shared-x ring (kSharedX): [kRing][8*NT8 rows][288 B], 16-byte chunk j of a row stored at j ^ (j >> 3)
  constexpr int kXRows = 8 * NT8;
  constexpr int kXStage = kXRows * kXPitch;
  constexpr int kXPerThread = (kXRows * 16 + kGemvThreads - 1) / kGemvThreads;  // 16-byte chunks per thread per tile
  const uint32_t xring = smem_u32(smem) + kGemvWarps * kRing * kSliceBytes;
  const uint32_t xring_end = xring + kRing * kXStage;
  uint32_t xst_issue = xring;
  const T* xsrc[kXPerThread];   // this thread's source chunk(s) of the next x tile to request
  uint32_t xdst[kXPerThread];
  bool xok[kXPerThread];
  int kx_issue = t_begin % KT;  // k-tile index xsrc[] currently addresses
  if constexpr (kSharedX) {
#pragma unroll
    for (int i = 0; i < kXPerThread; ++i) {
      const int ch = tid + i * kGemvThreads;
      const int row = ch >> 4, j = ch & 15;
      xok[i] = row < kXRows && row < M;
      xsrc[i] = x + int64_t(xok[i] ? row : 0) * K + kx_issue * kTileK + 8 * j;
      xdst[i] = uint32_t(row * kXPitch + 16 * (j ^ (j >> 3)));
    }
  }
  auto advance_x = [&]() {
    if constexpr (kSharedX) {
#pragma unroll
      for (int i = 0; i < kXPerThread; ++i) xsrc[i] += kTileK;
      if (++kx_issue == KT) {  // next n-tile restarts at k = 0
        kx_issue = 0;
#pragma unroll
        for (int i = 0; i < kXPerThread; ++i) xsrc[i] -= K;
      }
      xst_issue += kXStage;
      if (xst_issue == xring_end) xst_issue = xring;
    }
  };

  // cp.async this warp's slice of the next tile into its ring (+ this thread's share of the x tile)
  auto issue = [&](bool with_x) {
    if (n_issue > 0) {
      cp_async16(st_issue + 16 * lane, src_w);
      cp_async16(st_issue + 512 + 16 * lane, src_w + 512);
      cp_async16_if(meta_lane, st_issue + kSliceW + 16 * (lane % 3), src_m);
      src_w += kTileBytes;
      src_m += kTileBytes;
      st_issue += kSliceBytes;
      if (st_issue == ring_end) st_issue = ring;
      --n_issue;
      if constexpr (kSharedX) {
        if (with_x) {
#pragma unroll
          for (int i = 0; i < kXPerThread; ++i) cp_async16_if(xok[i], xst_issue + xdst[i], xsrc[i]);
          advance_x();
        }
      }
    }
    cp_async_commit();  // always commit: group accounting stays uniform
  };

  // ring prologue: weights never depend on the previous kernel, so they start before the PDL wait
#pragma unroll
  for (int i = 0; i < kRing - 1; ++i) issue(false);
  pdl_launch_dependents();
  pdl_wait();  // x, y and the workspace belong to the previous kernel until here
  if (nk <= 0) return;
  if constexpr (kSharedX) {
    // x of the prologue tiles: plain loads (they may not join the weight groups issued before
    // the PDL wait); made visible by the CTA barrier of the first fetch
    const int pre = nk < kRing - 1 ? nk : kRing - 1;
    for (int i = 0; i < pre; ++i) {
#pragma unroll
      for (int q = 0; q < kXPerThread; ++q)
        if (xok[q]) {
          const uint4 v = ldg128(xsrc[q]);
          asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(xst_issue + xdst[q]), "r"(v.x), "r"(v.y), "r"(v.z),
                       "r"(v.w)
                       : "memory");
        }
      advance_x();
    }
  }

  const uint32_t w_off_a = uint32_t(g) * 64u + 16u * uint32_t(c ^ ((g >> 1) & 3));
  const uint32_t w_off_b = w_off_a + 512u;
  const uint32_t s_off = kSliceW + 4u * uint32_t(g);        // scales of rows (g, g+8): one 32-bit word
  const uint32_t z_off = kSliceW + 32u + 2u * uint32_t(g);  // zeros of rows (g, g+8): two bytes
  // fp16: the group scale is applied in fp32 when a group's partial sum is folded in.
  // bf16: scale every weight in bf16 exactly as the reference's dequant does (its 8-bit
  //       significand makes the two orders differ by ~1e-3, the whole error budget).
  constexpr bool kScalePerWeight = (Elem<T>::kDtype == 1);

  float acc[NT8][4];
#pragma unroll
  for (int t = 0; t < NT8; ++t)
#pragma unroll
    for (int i = 0; i < 4; ++i) acc[t][i] = 0.f;

  // raw words of the NEXT tile are pulled into registers before the current tile is
  // multiplied, so cp.async completion + LDS latency hide behind the MMA work.  Two explicit
  // register buffers (A/B) and a x2-unrolled loop keep that free of register copies.
  struct Raw {
    uint4 wa, wb;
    uint32_t s2, z2;
  };
  uint32_t st_fetch = ring;
  auto fetch = [&](Raw& r) {
    if constexpr (kSharedX) {
      cp_async_wait<kRing - 3>();  // issue() runs AFTER the barrier in this mode: one group fewer in flight
      __syncthreads();             // x tile complete for every warp; everyone is done with the stage issue() reuses
    } else {
      cp_async_wait<kRing - 2>();
      __syncwarp();
    }
    r.wa = lds128(st_fetch + w_off_a);
    r.wb = lds128(st_fetch + w_off_b);
    r.s2 = lds32(st_fetch + s_off);
    r.z2 = lds16(st_fetch + z_off);
    st_fetch += kSliceBytes;
    if (st_fetch == ring_end) st_fetch = ring;
  };

  int nt = t_begin / KT, kt = t_begin - nt * KT;
  int kx = kt;  // k-tile index the activation pointers xp[] currently address
  // activations: lane (g, c) reads x[token 8t+g][128kt + 32c + 8sp .. +8).  Token indices are
  // clamped to M-1 (their columns are never stored), so the loads need no predicate.
  const T* xp[NT8];
#pragma unroll
  for (int t = 0; t < NT8; ++t) {
    int tok = 8 * t + g;
    tok = tok < M ? tok : M - 1;
    xp[t] = x + int64_t(tok) * K + kt * kTileK + 32 * c;
  }

  int left = nk;  // tiles of this CTA still to multiply
  uint32_t xst_cur = xring;  // shared-x stage of the tile being multiplied
  // lane (g, c) reads chunk j = 4c + sp of row 8t + g; physical chunk = j ^ (j >> 3) = 4c + (sp ^ (c >> 1))
  const uint32_t xfrag_off = uint32_t(g * kXPitch + 64 * c);
  const uint32_t xflip = uint32_t(c >> 1);
  // one k-tile: cur -> MMA, nxt <- ring
  auto step = [&](const Raw& cur, Raw& nxt) {
    uint4 xf[NT8][4];
    if constexpr (kSharedX) {
      if (left > 1) fetch(nxt);  // CTA barrier inside; nk is uniform over the CTA
      issue(true);
#pragma unroll
      for (int t = 0; t < NT8; ++t)
#pragma unroll
        for (int sp = 0; sp < 4; ++sp)
          xf[t][sp] = lds128(xst_cur + uint32_t(t * 8 * kXPitch) + xfrag_off + 16u * (uint32_t(sp) ^ xflip));
      xst_cur += kXStage;
      if (xst_cur == xring_end) xst_cur = xring;
    } else {
      // activations of this k-tile straight from L1/L2 (the line was prefetched during the previous tile)
#pragma unroll
      for (int t = 0; t < NT8; ++t) {
#pragma unroll
        for (int sp = 0; sp < 4; ++sp) xf[t][sp] = ldg128(xp[t] + 8 * sp);
        xp[t] += kTileK;
      }
      if (++kx == KT) {  // next n-tile restarts at k = 0
        kx = 0;
#pragma unroll
        for (int t = 0; t < NT8; ++t) xp[t] -= K;
      }
#pragma unroll
      for (int t = 0; t < NT8; ++t) asm volatile("prefetch.global.L1 [%0];" ::"l"(xp[t]));
      issue(false);
      if (left > 1) fetch(nxt);
    }

    const uint32_t wa[4] = {cur.wa.x, cur.wa.y, cur.wa.z, cur.wa.w};
    const uint32_t wb[4] = {cur.wb.x, cur.wb.y, cur.wb.z, cur.wb.w};
    uint32_t za_e, za_o, zb_e, zb_o;
    Unpack<T>::zero_consts_scalar(cur.z2 & 0xFFu, za_e, za_o);
    Unpack<T>::zero_consts_scalar(cur.z2 >> 8, zb_e, zb_o);
    const uint32_t sa2 = __byte_perm(cur.s2, 0, 0x1010), sb2 = __byte_perm(cur.s2, 0, 0x3232);

    // two independent accumulation chains (even / odd k-steps) halve the HMMA latency chain
    float ge[NT8][4], go[NT8][4];
#pragma unroll
    for (int t = 0; t < NT8; ++t)
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        ge[t][i] = kScalePerWeight ? acc[t][i] : 0.f;
        go[t][i] = 0.f;
      }

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
      if constexpr (kScalePerWeight) {
        a0[0] = Unpack<T>::mul(a0[0], sa2);
        a0[1] = Unpack<T>::mul(a0[1], sb2);
        a0[2] = Unpack<T>::mul(a0[2], sa2);
        a0[3] = Unpack<T>::mul(a0[3], sb2);
        a1[0] = Unpack<T>::mul(a1[0], sa2);
        a1[1] = Unpack<T>::mul(a1[1], sb2);
        a1[2] = Unpack<T>::mul(a1[2], sa2);
        a1[3] = Unpack<T>::mul(a1[3], sb2);
      }
#pragma unroll
      for (int t = 0; t < NT8; ++t) {
        mma_16816<T>(ge[t], a0, xf[t][sp].x, xf[t][sp].y);
        mma_16816<T>(go[t], a1, xf[t][sp].z, xf[t][sp].w);
      }
    }

    if constexpr (kScalePerWeight) {
#pragma unroll
      for (int t = 0; t < NT8; ++t)
#pragma unroll
        for (int i = 0; i < 4; ++i) acc[t][i] = ge[t][i] + go[t][i];
    } else {
      const float sa = bits16_to_float<T>(cur.s2 & 0xFFFFu), sb = bits16_to_float<T>(cur.s2 >> 16);
#pragma unroll
      for (int t = 0; t < NT8; ++t) {
        acc[t][0] = fmaf(sa, ge[t][0] + go[t][0], acc[t][0]);
        acc[t][1] = fmaf(sa, ge[t][1] + go[t][1], acc[t][1]);
        acc[t][2] = fmaf(sb, ge[t][2] + go[t][2], acc[t][2]);
        acc[t][3] = fmaf(sb, ge[t][3] + go[t][3], acc[t][3]);
      }
    }
    --left;
  };

  // close this CTA's segment of n-tile nt (per warp; no block barrier anywhere)
  auto close_segment = [&]() {
    // contributors of n-tile nt: CTAs whose tile range intersects [nt*KT, (nt+1)*KT)
    const int lo = nt * KT, hi = lo + KT;
    int first = int(unsigned(lo) * unsigned(grid) / unsigned(ttot));
    while (first > 0 && sk_begin(first, ttot, grid) > lo) --first;
    while (sk_begin(first + 1, ttot, grid) <= lo) ++first;
    int last = int(unsigned(hi - 1) * unsigned(grid) / unsigned(ttot));
    if (last >= grid) last = grid - 1;
    while (last + 1 < grid && sk_begin(last + 1, ttot, grid) < hi) ++last;
    while (sk_begin(last, ttot, grid) >= hi) --last;
    const int ncontrib = last - first + 1;
    const int rank = int(blockIdx.x) - first;

    bool do_store = true;
    if (ncontrib > 1) {
      float4* mine = reinterpret_cast<float4*>(ws_partial) +
                     ((int64_t(nt) * maxc + rank) * kGemvWarps + warp) * (32 * NT8) + lane;
#pragma unroll
      for (int t = 0; t < NT8; ++t) __stcg(mine + 32 * t, make_float4(acc[t][0], acc[t][1], acc[t][2], acc[t][3]));
      __syncwarp();  // the lane-0 release below publishes the whole warp's partial (cumulativity)
      unsigned ticket = 0;
      if (lane == 0)
        asm volatile("atom.add.acq_rel.gpu.global.u32 %0, [%1], 1;"
                     : "=r"(ticket)
                     : "l"(ws_sem + nt * kGemvWarps + warp)
                     : "memory");
      ticket = __shfl_sync(0xffffffffu, ticket, 0);
      do_store = (int(ticket) == ncontrib - 1);
      if (do_store) {
        const float4* base = reinterpret_cast<const float4*>(ws_partial) +
                             (int64_t(nt) * maxc * kGemvWarps + warp) * (32 * NT8) + lane;
#pragma unroll
        for (int t = 0; t < NT8; ++t) acc[t][0] = acc[t][1] = acc[t][2] = acc[t][3] = 0.f;
#pragma unroll 4
        for (int r = 0; r < ncontrib; ++r) {  // fixed rank order -> deterministic
#pragma unroll
          for (int t = 0; t < NT8; ++t) {
            const float4 p = __ldcg(base + int64_t(r) * kGemvWarps * (32 * NT8) + 32 * t);
            acc[t][0] += p.x;
            acc[t][1] += p.y;
            acc[t][2] += p.z;
            acc[t][3] += p.w;
          }
        }
        __syncwarp();
        if (lane == 0) ws_sem[nt * kGemvWarps + warp] = 0;  // leave the workspace clean
      }
    }
    if (do_store) {
      // acc[t] = {(row g, tok 8t+2c), (row g, tok 8t+2c+1), (row g+8, tok 8t+2c), (row g+8, tok 8t+2c+1)}
      const int n_lo = nt * kTileN + 16 * warp + g, n_hi = n_lo + 8;
      float b_lo = 0.f, b_hi = 0.f;
      if (bias != nullptr) {
        if (n_lo < N) b_lo = Elem<T>::to_float(bias[n_lo]);
        if (n_hi < N) b_hi = Elem<T>::to_float(bias[n_hi]);
      }
#pragma unroll
      for (int t = 0; t < NT8; ++t) {
        const int tok = 8 * t + 2 * c;
        if (tok < M) {
          if (n_lo < N) y[int64_t(tok) * N + n_lo] = Elem<T>::from_float(acc[t][0] + b_lo);
          if (n_hi < N) y[int64_t(tok) * N + n_hi] = Elem<T>::from_float(acc[t][2] + b_hi);
        }
        if (tok + 1 < M) {
          if (n_lo < N) y[int64_t(tok + 1) * N + n_lo] = Elem<T>::from_float(acc[t][1] + b_lo);
          if (n_hi < N) y[int64_t(tok + 1) * N + n_hi] = Elem<T>::from_float(acc[t][3] + b_hi);
        }
      }
    }
#pragma unroll
    for (int t = 0; t < NT8; ++t) acc[t][0] = acc[t][1] = acc[t][2] = acc[t][3] = 0.f;
  };

  Raw A, B;
  fetch(A);
  while (left > 0) {
    int seg = KT - kt;  // tiles of n-tile nt handled by this CTA in this segment
    if (seg > left) seg = left;
    kt += seg;
    for (; seg >= 2; seg -= 2) {
      step(A, B);
      step(B, A);
    }
    if (seg == 1) {
      step(A, B);
      A = B;
    }
    close_segment();
    if (kt == KT) {
      kt = 0;
      ++nt;
    }
  }
  cp_async_wait<0>();
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------

struct GemvPlan {
  int nt8, grid, n_tiles, k_tiles, ttot, maxc, smem;
};

template <typename T, int NT8, bool kSharedX>
static int gemv_resident_ctas() {
  static int cached[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (cached[dev] == 0) {
    const int smem = gemv_smem_bytes(NT8, kSharedX);
    cudaFuncSetAttribute(gemv_kernel<T, NT8, kSharedX>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, gemv_kernel<T, NT8, kSharedX>, kGemvThreads, smem) != cudaSuccess ||
        per_sm < 1)
      per_sm = 1;
    cached[dev] = per_sm * sm_count();
  }
  return cached[dev];
}

static int g_gemv_ctas_per_sm = 0;    // 0 = as many as fit
static int g_gemv_shared_x_min_m = 1;  // M >= this: activations shared through smem (65 = never)
static int g_gemv_direct_max_m = 4;    // M <= this: one-slab-per-CTA register-streaming kernel (awq_gemv_direct.cu)

int gemv_direct(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K, int64_t N,
                int dtype, cudaStream_t st, const ArPush* push_or_null = nullptr);
bool gemv_direct_fits(int64_t M, int64_t K);
void gemv_direct_set_variant(int v);
int gemv_cluster(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K, int64_t N,
                 int dtype, cudaStream_t st, const ArPush* push_or_null = nullptr);
bool gemv_cluster_fits(int64_t M, int64_t K);
void gemv_cluster_debug(int split, int variant);
int gemv_persist(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K, int64_t N,
                 int dtype, cudaStream_t st, int silu, const ArPush* push_or_null = nullptr);
bool gemv_persist_fits(int64_t M, int64_t K);
void gemv_persist_debug(int team, int variant);
int gemv_tc(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K, int64_t N, int dtype,
            void* workspace, int64_t workspace_bytes, cudaStream_t st, int silu);
bool gemv_tc_fits(int64_t M, int64_t K, int64_t N, int dtype);
int64_t gemv_tc_workspace_bytes(int64_t max_M, int64_t max_N);
void gemv_tc_debug(int ctas, int pf);
static int g_gemv_impl = 0;  // 0 = auto, 1 = direct, 2 = legacy stream-K, 3 = persistent, 4 = cluster (1-4: where the kernel fits, else legacy)

static GemvPlan gemv_plan(int64_t M, int64_t K, int64_t N, int resident, bool shared_x) {
  if (g_gemv_ctas_per_sm > 0 && g_gemv_ctas_per_sm * sm_count() < resident) resident = g_gemv_ctas_per_sm * sm_count();
  GemvPlan p;
  p.nt8 = M <= 8 ? 1 : (M <= 16 ? 2 : (M <= 32 ? 4 : 8));
  p.n_tiles = int((N + kTileN - 1) / kTileN);
  p.k_tiles = int(K / kTileK);
  p.ttot = p.n_tiles * p.k_tiles;
  p.grid = resident < p.ttot ? resident : p.ttot;
  const int tpc_min = p.ttot / p.grid;  // >= 1
  p.maxc = (p.k_tiles + tpc_min - 1) / tpc_min + 1;
  if (p.maxc > p.grid) p.maxc = p.grid;
  p.smem = gemv_smem_bytes(p.nt8, shared_x);
  return p;
}

static int64_t gemv_ws_bytes(const GemvPlan& p) {
  return int64_t(kSemInts) * 4 + int64_t(p.n_tiles) * p.maxc * kGemvWarps * 32 * p.nt8 * 16;
}

template <typename T, int NT8, bool kSharedX>
static int gemv_launch(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K, int64_t N,
                       void* workspace, int64_t workspace_bytes, cudaStream_t st) {
  const GemvPlan p = gemv_plan(M, K, N, gemv_resident_ctas<T, NT8, kSharedX>(), kSharedX);
  AWQ_CHECK(p.n_tiles * kGemvWarps <= kSemInts, "awq_b200_gemv: N too large");
  AWQ_CHECK(int64_t(p.ttot) * (p.grid + 1) < (int64_t(1) << 31), "awq_b200_gemv: problem too large");
  const int64_t need = gemv_ws_bytes(p);
  AWQ_CHECK(workspace != nullptr && workspace_bytes >= need,
            "awq_b200_gemv: workspace too small (%lld < %lld bytes); size it with awq_b200_workspace_bytes",
            (long long)workspace_bytes, (long long)need);
  int* sem = reinterpret_cast<int*>(workspace);
  float* partial = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(workspace) + kSemInts * sizeof(int));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(unsigned(p.grid));
  cfg.blockDim = dim3(kGemvThreads);
  cfg.dynamicSmemBytes = size_t(p.smem);
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  AWQ_CUDA_CHECK(cudaLaunchKernelEx(&cfg, gemv_kernel<T, NT8, kSharedX>, (const T*)x, (const uint8_t*)tiles, (const T*)bias,
                                    (T*)y, partial, sem, int(M), int(K), int(N), p.k_tiles, p.ttot, p.maxc),
                 "awq_b200_gemv: launch");
  return 0;
}

static bool gemv_impl_fits(int impl, int64_t M, int64_t K, int64_t N, int dtype) {
  switch (impl) {
    case 1: return M <= g_gemv_direct_max_m && gemv_direct_fits(M, K);
    case 2: return true;
    case 3: return gemv_persist_fits(M, K);
    case 4: return gemv_cluster_fits(M, K);
    case 5: return gemv_tc_fits(M, K, N, dtype);
    default: return false;
  }
}

// The router is a measured table (gemv_route_table.h, generated from profiles/r02_kernel_matrix.json:
// every decode kernel timed on every per-rank linear shape of Llama-3-8B/70B, Qwen2.5-32B and the
// Llama-2-7B MLP at TP 1..8, batch 1..32): the nearest measured (K, N) in log space, the batch rounded
// up to its bucket, and the fastest kernel of that cell that can take the actual problem.
// bf16 has its own table (kGemvRouteTableBf16, profiles/r02_kernel_matrix_bf16.json): its kernels scale every weight,
// which moves the crossovers (and the tcgen05 decode kernel is fp16 only)
static const GemvRouteEntry* gemv_route_entry(int64_t K, int64_t N, int dtype) {
  struct Slot {
    int64_t K, N;
    const GemvRouteEntry* e;
  };
  const bool bf = dtype == AWQ_B200_BF16;
  static thread_local Slot cache[2][8] = {};
  Slot& slot = cache[bf ? 1 : 0][(uint64_t(K) * 31u + uint64_t(N)) & 7u];
  if (slot.e != nullptr && slot.K == K && slot.N == N) return slot.e;
  const GemvRouteEntry* table = bf ? kGemvRouteTableBf16 : kGemvRouteTable;
  const int entries = bf ? kGemvRouteEntriesBf16 : kGemvRouteEntries;
  const float lk = log2f(float(K)), ln = log2f(float(N));
  const GemvRouteEntry* best = &table[0];
  float best_d = 1e30f;
  for (int i = 0; i < entries; ++i) {
    const GemvRouteEntry& e = table[i];
    const float d = fabsf(lk - log2f(float(e.K))) + fabsf(ln - log2f(float(e.N)));
    if (d < best_d) {
      best_d = d;
      best = &e;
    }
  }
  slot.K = K;
  slot.N = N;
  slot.e = best;
  return best;
}

static int gemv_route(int64_t M, int64_t K, int64_t N, int dtype) {
  if (M > 32) return 2;
  const int bucket = M <= 1 ? 0 : (M <= 2 ? 1 : (M <= 4 ? 2 : (M <= 8 ? 3 : (M <= 16 ? 4 : 5))));
  const GemvRouteEntry* e = gemv_route_entry(K, N, dtype);
  for (int r = 0; r < 5; ++r) {
    const int impl = e->rank[bucket][r];
    if (impl != 0 && gemv_impl_fits(impl, M, K, N, dtype)) return impl;
  }
  static const int generic[4] = {3, 4, 5, 2};
  for (int impl : generic)
    if (gemv_impl_fits(impl, M, K, N, dtype)) return impl;
  return 2;
}

// SiluAndMul-fused decode launch on gate_up tiles: the persistent kernel for M <= 8 (and for bf16 up to 16), the
// tcgen05 decode kernel for fp16 up to M = 32 -- the two kernels the measured table routes gate_up-sized
// matrices to.  Returns -1 when neither takes the problem (the caller then uses the tcgen05 GEMM).
int gemv_silu_mul(const void* x, const void* tiles, void* y, int64_t M, int64_t K, int64_t N, int dtype, void* workspace,
                  int64_t workspace_bytes, cudaStream_t st) {
  const int forced = g_gemv_impl;
  if (forced != 5 && M <= 8 && gemv_persist_fits(M, K)) return gemv_persist(x, tiles, nullptr, y, M, K, N, dtype, st, 1);
  if (forced != 3 && gemv_tc_fits(M, K, N, dtype)) return gemv_tc(x, tiles, nullptr, y, M, K, N, dtype, workspace, workspace_bytes, st, 1);
  if (forced != 5 && gemv_persist_fits(M, K)) return gemv_persist(x, tiles, nullptr, y, M, K, N, dtype, st, 1);
  return -1;
}

// Row-parallel decode launch whose epilogue also pushes y to the TP peers as all-reduce packets (gemv_common.cuh
// ArPush): the one-slab-per-CTA, the cluster and the persistent kernel, i.e. what the measured table routes the
// per-rank o_proj / down_proj shards to at TP = 2 / 4 / 8 for M <= 8.  Returns -1 when the table routes (M, K, N) to another kernel (the caller then
// runs the plain linear and the plain collective).
int gemv_pushed(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K, int64_t N, int dtype,
                const ArPush& push, cudaStream_t st) {
  if (g_gemv_impl != 0 || M < 1 || M > 16 || N % 16 != 0) return -1;
  switch (gemv_route(M, K, N, dtype)) {
    case 1: return gemv_direct(x, tiles, bias, y, M, K, N, dtype, st, &push);
    case 3: return gemv_persist(x, tiles, bias, y, M, K, N, dtype, st, 0, &push);
    case 4: return gemv_cluster(x, tiles, bias, y, M, K, N, dtype, st, &push);
    default: return -1;
  }
}

}  // namespace awq

using namespace awq;

extern "C" int awq_b200_gemv_route(int64_t M, int64_t K, int64_t N, int dtype) {
  if (M < 1 || M > 64 || K <= 0 || K % kTileK != 0 || N <= 0) return 0;
  return gemv_route(M, K, N, dtype);
}

extern "C" int awq_b200_debug_gemv_impl_fits(int impl, int64_t M, int64_t K, int64_t N, int dtype) {
  if (M < 1 || M > 64 || K <= 0 || K % kTileK != 0 || N <= 0) return 0;
  return gemv_impl_fits(impl, M, K, N, dtype) ? 1 : 0;
}

extern "C" void awq_b200_set_gemv_ctas_per_sm(int n) { g_gemv_ctas_per_sm = n; }
extern "C" void awq_b200_set_gemv_shared_x_min_m(int m) { g_gemv_shared_x_min_m = m < 1 ? 1 : m; }
extern "C" void awq_b200_debug_set_gemv_impl(int impl, int split, int variant) {
  g_gemv_impl = impl;
  gemv_persist_debug(split, variant);
  gemv_cluster_debug(split, variant);
  gemv_direct_set_variant(variant);
  gemv_tc_debug(impl == 5 ? split : 0, impl == 5 ? variant : 0);
}
extern "C" void awq_b200_set_gemv_direct_max_m(int m) { g_gemv_direct_max_m = m < 0 ? 0 : (m > 4 ? 4 : m); }

extern "C" int64_t awq_b200_workspace_bytes(int64_t max_M, int64_t max_N) {
  if (max_M < 1) max_M = 1;
  if (max_N < 1) max_N = 1;
  const int64_t nt = (max_N + kTileN - 1) / kTileN;
  const int64_t nt8 = max_M <= 8 ? 1 : (max_M <= 16 ? 2 : (max_M <= 32 ? 4 : 8));
  // sum over n-tiles of contributors <= 2 * resident CTAs + 2 * n_tiles; 1024 CTAs covers any sm_100 part
  const int64_t slots = 2 * 1024 + 2 * nt;
  const int64_t gemv = int64_t(kSemInts) * 4 + slots * kGemvWarps * 32 * nt8 * 16;
  const int64_t gemm = gemm_workspace_bytes(max_M, max_N);
  const int64_t tc = gemv_tc_workspace_bytes(max_M > 32 ? 32 : max_M, max_N);
  const int64_t a = gemv > gemm ? gemv : gemm;
  return a > tc ? a : tc;
}

extern "C" int awq_b200_gemv(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K,
                             int64_t N, int dtype, void* workspace, int64_t workspace_bytes, void* stream) {
  if (M == 0 || N == 0) return 0;
  AWQ_CHECK(M >= 1 && M <= 64, "awq_b200_gemv: M (%lld) must be in [1, 64]", (long long)M);
  AWQ_CHECK(K > 0 && K % kTileK == 0, "awq_b200_gemv: K (%lld) must be a positive multiple of 128", (long long)K);
  AWQ_CHECK(N > 0 && N % 8 == 0, "awq_b200_gemv: N (%lld) must be a positive multiple of 8", (long long)N);
  AWQ_CHECK(dtype == AWQ_B200_F16 || dtype == AWQ_B200_BF16, "awq_b200_gemv: bad dtype %d", dtype);
  AWQ_CHECK(x && tiles && y, "awq_b200_gemv: null pointer");
  AWQ_CHECK(((uintptr_t)x % 16) == 0 && ((uintptr_t)tiles % 16) == 0 && ((uintptr_t)y % 2) == 0,
            "awq_b200_gemv: x and tiles must be 16-byte aligned");
  AWQ_CHECK(((uintptr_t)workspace % 16) == 0, "awq_b200_gemv: workspace must be 16-byte aligned");
  AWQ_CHECK(K * N / 2 < (int64_t(1) << 40), "awq_b200_gemv: problem too large");
  if (int rc = awq_b200_check_device()) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  if (g_gemv_impl == 0) {
    switch (gemv_route(M, K, N, dtype)) {
      case 1: return gemv_direct(x, tiles, bias, y, M, K, N, dtype, st);
      case 3: return gemv_persist(x, tiles, bias, y, M, K, N, dtype, st, 0);
      case 4: return gemv_cluster(x, tiles, bias, y, M, K, N, dtype, st);
      case 5: return gemv_tc(x, tiles, bias, y, M, K, N, dtype, workspace, workspace_bytes, st, 0);
      default: break;  // 2: legacy stream-K ring kernel below
    }
  }
  if (g_gemv_impl == 5 && gemv_tc_fits(M, K, N, dtype))
    return gemv_tc(x, tiles, bias, y, M, K, N, dtype, workspace, workspace_bytes, st, 0);
  if (g_gemv_impl == 3 && gemv_persist_fits(M, K)) return gemv_persist(x, tiles, bias, y, M, K, N, dtype, st, 0);
  if (g_gemv_impl == 4 && gemv_cluster_fits(M, K)) return gemv_cluster(x, tiles, bias, y, M, K, N, dtype, st);
  if (g_gemv_impl == 1 && M <= g_gemv_direct_max_m && gemv_direct_fits(M, K))
    return gemv_direct(x, tiles, bias, y, M, K, N, dtype, st);
#define AWQ_GEMV_CASE(TT, NN, SX) \
  return gemv_launch<TT, NN, SX>(x, tiles, bias, y, M, K, N, workspace, workspace_bytes, st)
  const int nt8 = M <= 8 ? 1 : (M <= 16 ? 2 : (M <= 32 ? 4 : 8));
  const bool shared_x = M > g_gemv_shared_x_min_m - 1;
  if (dtype == AWQ_B200_F16) {
    switch (nt8) {
      case 1: if (shared_x) AWQ_GEMV_CASE(__half, 1, true); else AWQ_GEMV_CASE(__half, 1, false);
      case 2: AWQ_GEMV_CASE(__half, 2, true);
      case 4: AWQ_GEMV_CASE(__half, 4, true);
      default: AWQ_GEMV_CASE(__half, 8, true);
    }
  } else {
    switch (nt8) {
      case 1: if (shared_x) AWQ_GEMV_CASE(__nv_bfloat16, 1, true); else AWQ_GEMV_CASE(__nv_bfloat16, 1, false);
      case 2: AWQ_GEMV_CASE(__nv_bfloat16, 2, true);
      case 4: AWQ_GEMV_CASE(__nv_bfloat16, 4, true);
      default: AWQ_GEMV_CASE(__nv_bfloat16, 8, true);
    }
  }
#undef AWQ_GEMV_CASE
}
