// common.cuh -- device helpers shared by the sm_100a AWQ kernels (PTX wrappers, 4-bit unpack).
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace awq {

constexpr int kTileN = 128;           // output features per tile
constexpr int kTileK = 128;           // input features per tile (== AWQ group size)
constexpr int kTileWBytes = 8192;     // 128 rows x 16 int32
constexpr int kTileWWords = 2048;
constexpr int kTileMetaOff = 8192;    // 8 blocks x 48 B: [16 scales | 16 zero bytes], both in paired row order
constexpr int kTileMetaBlock = 48;
constexpr int kTileBytes = kTileWBytes + 8 * kTileMetaBlock;  // 8576 = 67 * 128: one self-contained record
// group sizes 64 / 32: a record carries 2 / 4 metadata sections of 8 x 48 B, one per 64- / 32-k sub-group, behind the
// weights (8960 / 9728 bytes; sub-group j of a row: the offsets below + kTileMetaSection * j).  Only the tcgen05 GEMM
// and the tiled dequant read such records.
constexpr int kTileMetaSection = 8 * kTileMetaBlock;  // 384
__host__ __device__ constexpr int tile_bytes_sub(int sub) { return kTileWBytes + kTileMetaSection * sub; }
constexpr int kTileBytesMax = kTileWBytes + 4 * kTileMetaSection;  // 9728
// scale / zero of row r inside a record
__host__ __device__ constexpr int tile_scale_off(int r) {
  return kTileMetaOff + kTileMetaBlock * (r >> 4) + 2 * ((((r & 15) & 7) << 1) | ((r & 15) >> 3));
}
__host__ __device__ constexpr int tile_zero_off(int r) {
  return kTileMetaOff + kTileMetaBlock * (r >> 4) + 32 + ((((r & 15) & 7) << 1) | ((r & 15) >> 3));
}

// see gemv_common.cuh (ar_push_pair) and awq_allreduce.cu
struct ArPush {
  void* const* peers;     // device array of the world's staging-buffer bases as mapped here (entry `rank` = local)
  long long ll_off;       // packet region inside a buffer
  long long src_stride;   // one source rank's packets of one slot
  long long slot_stride;  // one slot (all sources)
  int epoch_off;          // per-block epoch words (block = 128 vectors) inside the local buffer
  int rank, world;
};

// ---------------------------------------------------------------------------------------
// element traits
// ---------------------------------------------------------------------------------------
template <typename T>
struct Elem;
template <>
struct Elem<__half> {
  using T2 = __half2;
  static constexpr int kDtype = 0;
  static __device__ __forceinline__ float to_float(__half v) { return __half2float(v); }
  static __device__ __forceinline__ __half from_float(float v) { return __float2half_rn(v); }
};
template <>
struct Elem<__nv_bfloat16> {
  using T2 = __nv_bfloat162;
  static constexpr int kDtype = 1;
  static __device__ __forceinline__ float to_float(__nv_bfloat16 v) { return __bfloat162float(v); }
  static __device__ __forceinline__ __nv_bfloat16 from_float(float v) { return __float2bfloat16_rn(v); }
};

// ---------------------------------------------------------------------------------------
// bit tricks
// ---------------------------------------------------------------------------------------
template <int kLut>
__device__ __forceinline__ uint32_t lop3(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t r;
  asm("lop3.b32 %0, %1, %2, %3, %4;" : "=r"(r) : "r"(a), "r"(b), "r"(c), "n"(kLut));
  return r;
}
// (a & b) | c
__device__ __forceinline__ uint32_t and_or(uint32_t a, uint32_t b, uint32_t c) {
  return lop3<(0xf0 & 0xcc) | 0xaa>(a, b, c);
}

__device__ __forceinline__ uint32_t h2_sub(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("sub.rn.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
__device__ __forceinline__ uint32_t h2_add(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("add.rn.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
__device__ __forceinline__ float bits16_as_half_to_float(uint32_t b) {
  const uint16_t h = uint16_t(b);
  return __half2float(*reinterpret_cast<const __half*>(&h));
}
__device__ __forceinline__ uint32_t h2_mul(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("mul.rn.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
__device__ __forceinline__ uint32_t h2_fma(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t r;
  asm("fma.rn.f16x2 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  return r;
}
__device__ __forceinline__ uint32_t bf2_sub(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("sub.rn.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
__device__ __forceinline__ uint32_t bf2_mul(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("mul.rn.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}

// ---------------------------------------------------------------------------------------
// 4-bit unpack.  A 32-bit word holds nibbles p0..p7; the "quad" q (0..3) is the pair of
// nibbles (q, q+4), returned as a packed 2x16-bit float pair {lo = nibble q, hi = nibble q+4}.
//
// fp16: "raw" forms avoid shifts:  q even -> 1024 + v    (v in the low mantissa bits)
//                                  q odd  -> 1024 + 16 v (v four bits higher)
//       unpack4_raw returns the four raw forms; sub_zero turns raw(w) and the matching
//       zero constant into the EXACT integer (w - z) as fp16.
// bf16: mantissa is 7 bits, so every quad is shifted down: raw = 128 + v for all q.
// ---------------------------------------------------------------------------------------
template <typename T>
struct Unpack;

template <>
struct Unpack<__half> {
  // raw[q] for q = 0..3
  static __device__ __forceinline__ void raw(uint32_t w, uint32_t (&r)[4]) {
    const uint32_t top = w >> 8;
    r[0] = and_or(w, 0x000f000fu, 0x64006400u);    // {1024 + p0, 1024 + p4}
    r[1] = and_or(w, 0x00f000f0u, 0x64006400u);    // {1024 + 16 p1, 1024 + 16 p5}
    r[2] = and_or(top, 0x000f000fu, 0x64006400u);  // {1024 + p2, 1024 + p6}
    r[3] = and_or(top, 0x00f000f0u, 0x64006400u);  // {1024 + 16 p3, 1024 + 16 p7}
  }
  // zero constants from the raw forms of the zero word: zc[q] such that diff() is exact
  static __device__ __forceinline__ void zero_consts(const uint32_t (&zr)[4], uint32_t (&zc)[4]) {
    zc[0] = zr[0];                              // 1024 + z
    zc[1] = h2_mul(zr[1], 0xac00ac00u);         // (1024 + 16 z) * (-1/16) = -(64 + z)
    zc[2] = zr[2];
    zc[3] = h2_mul(zr[3], 0xac00ac00u);
  }
  // exact (w - z) for quad q
  template <int q>
  static __device__ __forceinline__ uint32_t diff(uint32_t wr, uint32_t zc) {
    if constexpr ((q & 1) == 0) {
      return h2_sub(wr, zc);                    // (1024 + w) - (1024 + z)
    } else {
      return h2_fma(wr, 0x2c002c00u, zc);       // (1024 + 16 w)/16 - (64 + z)
    }
  }
  // zero constants for a scalar zero z (same z for both halves, any quad parity)
  static __device__ __forceinline__ void zero_consts_scalar(uint32_t z, uint32_t& zc_even, uint32_t& zc_odd) {
    zc_even = z * 0x00010001u + 0x64006400u;    // {1024 + z} x2  (ulp at 2^10 is 1)
    zc_odd = z * 0x00100010u + 0xd400d400u;     // {-(64 + z)} x2 (ulp at 2^6 is 1/16)
  }
  static __device__ __forceinline__ uint32_t mul(uint32_t a, uint32_t b) { return h2_mul(a, b); }
};

template <>
struct Unpack<__nv_bfloat16> {
  static __device__ __forceinline__ void raw(uint32_t w, uint32_t (&r)[4]) {
    r[0] = and_or(w, 0x000f000fu, 0x43004300u);        // {128 + p0, 128 + p4}
    r[1] = and_or(w >> 4, 0x000f000fu, 0x43004300u);   // {128 + p1, 128 + p5}
    r[2] = and_or(w >> 8, 0x000f000fu, 0x43004300u);
    r[3] = and_or(w >> 12, 0x000f000fu, 0x43004300u);
  }
  static __device__ __forceinline__ void zero_consts(const uint32_t (&zr)[4], uint32_t (&zc)[4]) {
#pragma unroll
    for (int q = 0; q < 4; ++q) zc[q] = zr[q];
  }
  template <int q>
  static __device__ __forceinline__ uint32_t diff(uint32_t wr, uint32_t zc) {
    return bf2_sub(wr, zc);                            // (128 + w) - (128 + z)
  }
  static __device__ __forceinline__ void zero_consts_scalar(uint32_t z, uint32_t& zc_even, uint32_t& zc_odd) {
    zc_even = z * 0x00010001u + 0x43004300u;    // {128 + z} x2
    zc_odd = zc_even;
  }
  static __device__ __forceinline__ uint32_t mul(uint32_t a, uint32_t b) { return bf2_mul(a, b); }
};

// ---------------------------------------------------------------------------------------
// shared memory / mbarrier / bulk-copy PTX
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// raises the barrier's pending transaction bytes without arriving (the issuing thread arrives separately)
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}
// Same, with a back-off between polls.  Polling is not free: try_wait issues on the XU pipe, and
// a CTA with ~20 warps spinning saturates it (ncu: xu pipe 100 %, profiles/), which slows every
// barrier operation of the warps that do have work.
__device__ __forceinline__ void mbar_wait_sleep(uint32_t bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) asm volatile("nanosleep.u32 128;" ::: "memory");
}

// Same, suspending in hardware: try_wait with a suspend-time hint parks the thread until the phase
// completes (or the hint expires), so a long wait costs a handful of instructions instead of a
// poll loop (ncu on the tcgen05 decode kernel: 25 % of all issued instructions were polls).
__device__ __forceinline__ void mbar_wait_susp(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}"
      ::"r"(bar), "r"(parity), "r"(0x989680u)
      : "memory");
}

// 1-D bulk async copy global -> shared, completion counted in bytes on an mbarrier (UBLKCP)
__device__ __forceinline__ void bulk_g2s(uint32_t dst_smem, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_smem),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
// same with an L2 eviction-priority hint (createpolicy value)
__device__ __forceinline__ void bulk_g2s_hint(uint32_t dst_smem, const void* src, uint32_t bytes, uint32_t bar,
                                              uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
          dst_smem),
      "l"(src), "r"(bytes), "r"(bar), "l"(policy)
      : "memory");
}
__device__ __forceinline__ uint64_t policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t policy_evict_last() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}

// L2 prefetch of a contiguous byte range (multiple of 16 B) through the bulk-copy engine: no
// shared memory, no registers, no completion to wait for
__device__ __forceinline__ void bulk_prefetch_l2(const void* src, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(bytes) : "memory");
}
constexpr int kPfPiece = 16384;
// The first `ncta` CTAs of a grid walk [base, base + bytes) front to back in kPfPiece pieces
// (piece index = round * ncta * nissuers + issuer * ncta + cta): the whole chip's prefetch front
// sweeps the array in memory order, which is also the order the weights are consumed in.
__device__ __forceinline__ void l2_prefetch_sweep(const uint8_t* base, int64_t bytes, int cta, int ncta, int issuer,
                                                  int nissuers) {
  const int64_t npieces = (bytes + kPfPiece - 1) / kPfPiece;
  for (int64_t p = int64_t(issuer) * ncta + cta; p < npieces; p += int64_t(nissuers) * ncta) {
    const int64_t off = p * kPfPiece;
    const int64_t len = bytes - off < kPfPiece ? bytes - off : kPfPiece;
    bulk_prefetch_l2(base + off, uint32_t(len));
  }
}

// 16-byte asynchronous copy global -> shared through the LSU path (LDGSTS), L2 evict-first hint
__device__ __forceinline__ void cp_async16_hint(uint32_t dst, const void* src, uint64_t policy) {
  asm volatile("cp.async.cg.shared.global.L2::cache_hint [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "l"(policy) : "memory");
}
__device__ __forceinline__ void cp_async16_plain(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
// the mbarrier receives one arrival from this thread once all its prior cp.async have landed
// (the arrival is pre-counted in the barrier's init count)
__device__ __forceinline__ void cp_async_mbar_arrive_noinc(uint32_t bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}
// orders this thread's later async-proxy operations (tcgen05.mma / TMA reads of shared memory)
// after generic-proxy writes it has observed
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// programmatic dependent launch (no-ops when the launch carries no PDL attribute)
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts128(uint32_t addr, const uint4& v) {
  asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void sts32(uint32_t addr, uint32_t v) {
  asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint32_t lds16(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint32_t lds8(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

}  // namespace awq
