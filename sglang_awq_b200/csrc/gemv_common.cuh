// gemv_common.cuh -- pieces shared by the decode kernels: legacy-tensor-pipe MMA wrappers,
// streaming global loads, and the per-group fragment math (one 16-row x 128-k slice).
#pragma once

#include "common.cuh"

namespace awq {

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

// activations / small reusable data: read-only path, default caching
__device__ __forceinline__ uint4 ldg128(const void* p) {
  uint4 v;
  asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}
// weights: read exactly once -> do not allocate in L1, first to leave L2
__device__ __forceinline__ uint4 ldg128_stream(const void* p) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p));
  return v;
}
__device__ __forceinline__ uint32_t ldg32_stream(const void* p) {
  uint32_t v;
  asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ uint32_t ldg16_stream(const void* p) {
  uint32_t v;  // 32-bit destination, zero-extended by the load: no widening instruction behind it
  asm volatile("ld.global.nc.L1::no_allocate.u16 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}

template <typename T>
__device__ __forceinline__ float bits16_to_float(uint32_t b) {
  const uint16_t h = uint16_t(b);
  return Elem<T>::to_float(*reinterpret_cast<const T*>(&h));
}

// One 16-row x 128-k slice of a tile against NT8 blocks of 8 tokens.
//   wa / wb : the lane's 16-byte granule of row g / row g+8 (words sp = 0..3, 8 consecutive k each)
//   s2      : {scale g, scale g+8} (2 x 16 bit)      z2 : {zero g, zero g+8} (2 bytes)
//   xf[t][sp]: token 8t+g, k = 32c + 8sp .. +8  (one 16-byte piece of the x row)
//   acc[t]  : {(row g, tok 8t+2c), (row g, tok 8t+2c+1), (row g+8, tok 8t+2c), (row g+8, tok 8t+2c+1)}
// fp16: the group scale is applied in fp32 when the group's partial sum is folded in.
// bf16: every weight is scaled in bf16 exactly as the reference's dequant does (its 8-bit
//       significand makes the two orders differ by ~1e-3, the whole error budget).
// kScalePerWeight can be forced for fp16 too (sub-group records: a lane's scale belongs to ITS granule's group, so it
// cannot be applied to the sum an MMA takes over the four lanes of a quad; awq_gemv_sub.cu).
template <typename T, int NT8, bool kScalePerWeight = (Elem<T>::kDtype == 1)>
__device__ __forceinline__ void gemv_group_mma(const uint4& wa4, const uint4& wb4, uint32_t s2, uint32_t z2,
                                               const uint4 (*xf)[4], float (*acc)[4]) {
  const uint32_t wa[4] = {wa4.x, wa4.y, wa4.z, wa4.w};
  const uint32_t wb[4] = {wb4.x, wb4.y, wb4.z, wb4.w};
  uint32_t za_e, za_o, zb_e, zb_o;
  Unpack<T>::zero_consts_scalar(z2 & 0xFFu, za_e, za_o);
  Unpack<T>::zero_consts_scalar(z2 >> 8, zb_e, zb_o);
  const uint32_t sa2 = __byte_perm(s2, 0, 0x1010), sb2 = __byte_perm(s2, 0, 0x3232);

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
    const float sa = bits16_to_float<T>(s2 & 0xFFFFu), sb = bits16_to_float<T>(s2 >> 16);
#pragma unroll
    for (int t = 0; t < NT8; ++t) {
      acc[t][0] = fmaf(sa, ge[t][0] + go[t][0], acc[t][0]);
      acc[t][1] = fmaf(sa, ge[t][1] + go[t][1], acc[t][1]);
      acc[t][2] = fmaf(sb, ge[t][2] + go[t][2], acc[t][2]);
      acc[t][3] = fmaf(sb, ge[t][3] + go[t][3], acc[t][3]);
    }
  }
}


// Fused row-parallel epilogue: besides storing y, a decode kernel pushes it to the TP peers as the all-reduce's
// self-validating packets (awq_allreduce.cu, one-shot packet protocol: vector v = 8 elements -> two 16-byte words
// {d, epoch, d, epoch} at [slot][source rank][2 v]), so the NVLink transfer overlaps the rest of the launch and the
// collective that follows only polls and sums.  peers == nullptr: plain kernel.

// Call with all 32 lanes.  Lane l holds element (tok, n), lane l + 16 element (tok, n + 1), n even for l < 16 (the
// decode kernels' reduction order: consecutive rows of a slab are 16 lanes apart).  `v` = the final output value.
template <typename T>
__device__ __forceinline__ void ar_push_pair(const ArPush& p, int tok, int n, int N, bool valid, float v) {
  const T tv = Elem<T>::from_float(v);
  const uint32_t lo = uint32_t(*reinterpret_cast<const uint16_t*>(&tv));
  const uint32_t hi = __shfl_down_sync(0xffffffffu, lo, 16);
  if (p.peers == nullptr) return;  // uniform
  if (valid && (threadIdx.x & 16) == 0) {
    const long long e = (long long)tok * N + n;  // element of the [M, N] message (n even: a whole 4-byte packet)
    const long long vec = e >> 3;
    const uint8_t* mine = reinterpret_cast<const uint8_t*>(p.peers[p.rank]);
    const unsigned int epoch = *reinterpret_cast<const volatile unsigned int*>(mine + p.epoch_off + 4 * (vec >> 7)) + 1u;
    const long long off = p.ll_off + (long long)(epoch & 1u) * p.slot_stride + p.rank * p.src_stride + 32 * vec + 8 * ((e & 7) >> 1);
    const uint32_t data = lo | (hi << 16);
    for (int j = 1; j < p.world; ++j) {
      int r = p.rank + j;
      if (r >= p.world) r -= p.world;
      asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1, %2};" ::"l"(reinterpret_cast<uint8_t*>(p.peers[r]) + off), "r"(data),
                   "r"(epoch)
                   : "memory");
    }
  }
}

constexpr int kXBlock = 320;  // bytes per (group, token) block of staged x: 256 B + 64 B skew

__device__ __forceinline__ uint2 lds64(uint32_t addr) {
  uint2 v;
  asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
  return v;
}

// fp16 fast path: one 16-row x 128-k slice against NT8 blocks of 8 tokens.
//   even quads (nibbles 0,4 / 2,6 of a word) -> value w * 2^-24, k pairs (8w+0,1) and (8w+4,5)
//   odd  quads (nibbles 1,5 / 3,7)           -> value w * 2^-20, k pairs (8w+2,3) and (8w+6,7)
//   nsx[t] = -2^-24 * {Sx(tok 8t+2c), Sx(tok 8t+2c+1)}
//   xf[t][sp] = x of token 8t+g, k = 8(4c+sp) + {0,1, 4,5, 2,3, 6,7}   (permuted while staging)
template <int NT8>
__device__ __forceinline__ void group_mma_subnormal(const uint4& wa4, const uint4& wb4, uint32_t s2, uint32_t z2,
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
    ae[0] = wa[sp] & 0x000f000fu;  // row g,   k 8w+0,1
    ae[1] = wb[sp] & 0x000f000fu;  // row g+8
    ae[2] = ta & 0x000f000fu;      // row g,   k 8w+4,5
    ae[3] = tb & 0x000f000fu;
    ao[0] = wa[sp] & 0x00f000f0u;  // row g,   k 8w+2,3  (x16)
    ao[1] = wb[sp] & 0x00f000f0u;
    ao[2] = ta & 0x00f000f0u;      // row g,   k 8w+6,7  (x16)
    ao[3] = tb & 0x00f000f0u;
#pragma unroll
    for (int t = 0; t < NT8; ++t) {
      mma_16816<__half>(ge[t], ae, xf[t][sp].x, xf[t][sp].y);  // staged as (k01, k45 | k23, k67): B pairs are
      mma_16816<__half>(go[t], ao, xf[t][sp].z, xf[t][sp].w);  // adjacent registers, no moves
    }
  }
  const float sa = bits16_to_float<__half>(s2 & 0xFFFFu) * 16777216.f, sb = bits16_to_float<__half>(s2 >> 16) * 16777216.f;
  const float za = float(z2 & 0xFFu), zb = float(z2 >> 8);
#pragma unroll
  for (int t = 0; t < NT8; ++t) {
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


// Stage x[tok][k0 + 128 q + ...] for tok < M, q < ng into `xs` as (group, token) blocks of
// kXBlock bytes (chunk (c, sp) of a block at 64 sp + 16 c: conflict-free fragment reads) and,
// on the fp16 fast path, the negated, 2^-24-scaled group sums into `sxs` ([q][kSxStride] floats).
// Work items = (pair of groups, token); the caller's `nlanes` cooperating warps take items
// wid, wid + nlanes, ...; a warp issues the loads of up to kU items before it touches any (each
// round is one L2 round trip: at M = 16 a 16-warp CTA stages 256 items, 4 rounds of 4 or 2 of 8).
// kSub > 1 (group size 64 / 32, awq_gemv_sub.cu): the sums are taken per SUB-group and stored per MMA token COLUMN,
// column = sub * kMP + token (kMP token slots per sub-group), see gemv_sub_kernel.
template <typename T, int M, int kSxStride, bool kFast, int kU = 4, int kSub = 1, int kMP = 0>
__device__ __forceinline__ void stage_x(const T* __restrict__ x, int K, int k0, int ng, uint32_t xs, uint32_t sxs, int wid,
                                        int nlanes, int lane) {
  const int half = lane >> 4, j = lane & 15;
  const uint32_t dst_j = uint32_t(((j & 3) * 4 + (j >> 2)) * 16);
  const int nqp = (ng + 1) >> 1;
  const int nitems = nqp * M;
  const T* xcol = x + k0 + 8 * j;
  for (int it0 = wid; it0 < nitems; it0 += kU * nlanes) {
    uint4 v[kU];
    int qs[kU], toks[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int it = it0 + u * nlanes;
      const int qp = M == 1 ? it : it / M;  // M is a compile-time constant
      toks[u] = it - qp * M;
      qs[u] = 2 * qp + half;
      v[u] = make_uint4(0u, 0u, 0u, 0u);
      if (it < nitems && qs[u] < ng) v[u] = ldg128(xcol + int64_t(toks[u]) * K + qs[u] * kTileK);
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      if (it0 + u * nlanes < nitems) {  // warp-uniform
        const bool ok = qs[u] < ng;
        const uint32_t dst = xs + uint32_t((qs[u] * M + toks[u]) * kXBlock) + dst_j;
        if constexpr (kFast) {
          // (k01, k45 | k23, k67): the B register pairs of the even / odd MMAs become adjacent
          if (ok) sts128(dst, make_uint4(v[u].x, v[u].z, v[u].y, v[u].w));
          // group sums on the tensor pipe: A = the lane's 8 x values, B = ones
          float d[4] = {0.f, 0.f, 0.f, 0.f};
          const uint32_t a[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
          mma_16816<__half>(d, a, 0x3c003c00u, 0x3c003c00u);
          float sum = d[0] + d[2];  // the quad's 32 k
          if constexpr (kSub < 4) sum += __shfl_xor_sync(0xffffffffu, sum, 4);
          if constexpr (kSub < 2) sum += __shfl_xor_sync(0xffffffffu, sum, 8);
          if constexpr (kSub == 1) {
            if (ok && j == 0) sts32(sxs + uint32_t((qs[u] * kSxStride + toks[u]) * 4), __float_as_uint(sum * -5.9604644775390625e-08f));
          } else {
            constexpr int kLanesPerSub = 16 / kSub;  // lanes j of one sub-group (8 k each)
            if (ok && (j & (kLanesPerSub - 1)) == 0)
              sts32(sxs + uint32_t((qs[u] * kSxStride + (j / kLanesPerSub) * kMP + toks[u]) * 4),
                    __float_as_uint(sum * -5.9604644775390625e-08f));
          }
        } else {
          if (ok) sts128(dst, v[u]);
        }
      }
    }
  }
}

}  // namespace awq
