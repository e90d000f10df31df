// awq_gemv_tc.cu -- kernel 2c: decode path on the 5th-gen tensor cores (fp16, 1 <= M <= 32).
//
//   y[M,N] = x[M,K] @ W (+ bias)
//
// Replaces the three launches of the reference's AWQLinearMethod.apply
// (python/sglang/srt/layers/quantization/awq.py:434-451) for decode batches.  The op is
// HBM-bound; what the mma.sync decode kernels run out of above M = 8 is the legacy tensor pipe
// (two HMMA.16816 per k-step, ncu profiles/r01b_gemv_m16_gate_up.txt) and below that the x
// fragment loads.  Here the streaming warps do NOTHING but load, mask and store:
//   * weights travel HBM -> REGISTERS (LDG.128, no shared-memory ring, no bulk-copy engine in the
//     way: its ~22 B/clk/SM is below the HBM share of an SM); a convert warp owns 32 rows of a
//     128 x 128 record and keeps kDepth records in flight;
//   * int4 -> tensor-core operand costs 5 ALU ops per 8 weights and no conversion: a nibble masked
//     in place IS the fp16 subnormal w * 2^-24 (even quads) or 16 w * 2^-24 (odd quads).  The
//     masked words go straight to tensor memory (tcgen05.st.16x256b: the lane-quad register
//     layout of the loads is exactly that shape's fragment layout, one instruction per 16-row
//     slab); the x16 of the odd quads is undone EXACTLY by the MMA itself: the four odd k-steps
//     run first, the first even k-step then accumulates with scale-input-d = 2^-4;
//   * the zero point never touches the weight operands either: (w - z) . x = w . x + (-z) . x, and the
//     second term is eight more MMAs per tile against a 16-column operand that holds the row's
//     constant -z * 2^-24 (-16 z * 2^-24 for the odd k-steps): the tensor pipe is ~90 % idle in this
//     kernel, the ALU and issue slots of the streaming warps are not;
//   * the x tile of a record is staged by the same four warps that convert the record, a quarter
//     each, into the group's 128-byte-swizzled stage with its k order permuted to the operand order
//     of the masked words (8-byte pieces): A operand and x stage share one full / empty barrier
//     pair, and no single warp carries a per-tile job of more than a few dozen instructions (one
//     warp issues an instruction every 4-6 clocks; a tile has ~460);
//   * two threads issue tcgen05.mma for alternate tiles (M = 128 rows, N = 16/32 tokens, K = 16
//     per instruction, A from TMEM); every 128-k tile (= one AWQ group) gets a fresh fp32
//     accumulator that ends up holding 2^-24 * Sum (w - z) x exactly; four drain warps read it back
//     (tcgen05.ld), apply the group scale in fp32 and keep the running sum in registers; the scale
//     comes straight from the record's metadata block (L2-prefetched 24 tiles ahead);
//   * work split = stream-K over the (n-tile, k-tile) records in memory order, one CTA per SM,
//     split n-tiles closed through the fp32 workspace in RANK order (deterministic).
#include "tcgen05.cuh"
#include "gemv_common.cuh"
#include "host_common.h"

#include <cuda.h>
#include <mutex>
#include <utility>

namespace awq {

constexpr int kTcMaxGroups = 4;       // convert groups of four warps (one warp per TMEM lane quarter); group g owns A buffer g
constexpr int kTcDrainWarps = 4;
constexpr int kTcFirstConvert = 4;    // warps 0..3: drain + x staging (TMEM lane quarter = warp); then the convert groups:
                                      // a group = 4 consecutive warps = 4 distinct TMEM lane quarters (warp % 4);
                                      // the last kTcIssuers warps issue the MMAs
__host__ __device__ constexpr int tc_first_mma(int groups) { return kTcFirstConvert + 4 * groups; }
constexpr int kTcIssuers = 4;         // issuer warps: issuer j takes tiles j, j + 4, ... (one thread issues a tcgen05.mma every ~50 clocks)
__host__ __device__ constexpr int tc_threads(int groups) { return 32 * (tc_first_mma(groups) + kTcIssuers); }
constexpr int kTcND = 8;              // most accumulator buffers in TMEM (BN columns each)
constexpr int kTcSemInts = 8192;
constexpr int kTcL2Rounds = 5;        // a group pushes whole records into L2 this many of its rounds ahead of its register loads
constexpr int kTcL2MaxTiles = 128;    // ... when the CTA's stream is shorter than this many records
constexpr int kTcMetaDepth = 4;       // tiles between a drain thread's metadata load and its use

struct TcGeom {
  int M, K, N, KT, S, ttot;
  int silu;  // gate_up tiles: y[M, N / 2] = silu(gate) * up (rows r, r + 8 of a 16-row slab)
#ifdef AWQ_TC_TRACE
  unsigned long long* trace;  // development aid (awq_b200_debug_set_tc_trace): phase time stamps, else null
  int trace_seq;              // launch number since tracing was switched on (host side)
#endif
};

__device__ __forceinline__ unsigned long long tc_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// The stamps cost registers in a kernel that has none to spare (96 per thread, the in-flight weight records live
// there): they exist only in -DAWQ_TC_TRACE builds (scripts/trace_tc.py builds one), never in the shipped library.
#ifdef AWQ_TC_TRACE
// slot s of launch `launch` (16 slots per launch)
#define TC_STAMP(s)                                                                     \
  do {                                                                                  \
    if (geo.trace != nullptr && blockIdx.x == 0) geo.trace[16 + 16 * trace_launch + (s)] = tc_now(); \
  } while (0)
// every CTA, launches 4..7 of a traced sequence: slot s of [launch - 4][cta][8] behind the CTA-0 region
#define TC_STAMP_ALL(s)                                                                                   \
  do {                                                                                                    \
    if (geo.trace != nullptr && (geo.trace_seq & ~3) == 4)                                                \
      geo.trace[16 + 64 * 16 + (((geo.trace_seq & 3) * 160 + blockIdx.x) << 3) + (s)] = tc_now();         \
  } while (0)
#else
#define TC_STAMP(s) do {} while (0)
#define TC_STAMP_ALL(s) do {} while (0)
#endif

// first tile of CTA i (host guarantees ttot * (grid + 1) < 2^31)
__host__ __device__ __forceinline__ int tc_begin(int i, int grid, int ttot) {
  return int(unsigned(i) * unsigned(ttot) / unsigned(grid));
}

__host__ __device__ constexpr int tc_x_bytes(int bn) { return 2 * bn * 128; }

// D[tmem] = A[tmem] * B[smem]^T + D * 2^-4   (kind::f16 "scale-input-d")
__device__ __forceinline__ void mma_ts_scale_d4(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, 1, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p, 4;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc)
      : "memory");
}

// 16 TMEM lanes x 32 columns: thread (g = lane >> 2, c = lane & 3) supplies, for repetition i,
// v[4i], v[4i+1] -> lane g, columns 8i + 2c, 8i + 2c + 1 and v[4i+2], v[4i+3] -> lane g + 8, same columns
__device__ __forceinline__ void st_16x256b_x4(uint32_t taddr, const uint32_t* v) {
  asm volatile(
      "tcgen05.st.sync.aligned.16x256b.x4.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
      ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]),
      "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
      : "memory");
}

// f(integral_constant<int, 0>) ... f(integral_constant<int, kN - 1>): a loop whose index is usable as a template argument
template <class F, int... kI>
__device__ __forceinline__ void tc_static_for_impl(F&& f, std::integer_sequence<int, kI...>) {
  (f(std::integral_constant<int, kI>{}), ...);
}
template <int kN, class F>
__device__ __forceinline__ void tc_static_for(F&& f) {
  tc_static_for_impl(static_cast<F&&>(f), std::make_integer_sequence<int, kN>{});
}

template <int kOff>
__device__ __forceinline__ void sts32(uint32_t addr, uint32_t a) {
  asm volatile("st.shared.u32 [%0+%2], %1;" ::"r"(addr), "r"(a), "n"(kOff) : "memory");
}
// st.shared.v2 at register + compile-time offset (the offset folds into the instruction)
template <int kOff>
__device__ __forceinline__ void sts64(uint32_t addr, uint32_t a, uint32_t b) {
  asm volatile("st.shared.v2.u32 [%0+%3], {%1, %2};" ::"r"(addr), "r"(a), "r"(b), "n"(kOff) : "memory");
}

// One 16-row slab (rows g, g + 8 of it: granules wa, wb) -> operand registers of two st_16x256b_x4.
//   k-step sp     (even quads, value w * 2^-24):    columns 2c, 2c+1 = k 8w + {0,1}, 8w + {4,5},  w = 4c + sp
//   k-step 4 + sp (odd quads,  value 16 w * 2^-24): columns 2c, 2c+1 = k 8w + {2,3}, 8w + {6,7}
template <uint32_t kMask>
__device__ __forceinline__ void tc_mask_slab(const uint4& wa4, const uint4& wb4, uint32_t (&v)[16]) {
  const uint32_t wa[4] = {wa4.x, wa4.y, wa4.z, wa4.w};
  const uint32_t wb[4] = {wb4.x, wb4.y, wb4.z, wb4.w};
  // (w >> 8 as __umulhi(w, 1 << 24) would move the shift to the FMA pipe, but IMAD.HI is slower: measured)
#pragma unroll
  for (int sp = 0; sp < 4; ++sp) {
    v[4 * sp + 0] = wa[sp] & kMask;
    v[4 * sp + 1] = (wa[sp] >> 8) & kMask;
    v[4 * sp + 2] = wb[sp] & kMask;
    v[4 * sp + 3] = (wb[sp] >> 8) & kMask;
  }
}
// Predicated streaming loads INTO the given registers ("+r": the destination is the in-flight
// buffer itself).  A conditional assignment of a load's result makes the compiler load into a
// temporary and copy, and the copy waits for the data: the full memory latency per record.
__device__ __forceinline__ void ldg128_stream_if(bool pred, uint4& v, const void* p) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "@p ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%5];\n\t}"
      : "+r"(v.x), "+r"(v.y), "+r"(v.z), "+r"(v.w)
      : "r"(uint32_t(pred)), "l"(p));
}
__device__ __forceinline__ void ldg16_stream_if(bool pred, uint32_t& v, const void* p) {
  // a 32-bit destination (zero-extended by the load itself): a separate widening instruction
  // would sit right behind the load and wait out the memory latency
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %1, 0;\n\t"
      "@p ld.global.nc.L1::no_allocate.u16 %0, [%2];\n\t}"
      : "+r"(v)
      : "r"(uint32_t(pred)), "l"(p));
}
__device__ __forceinline__ void st_16x256b_x1(uint32_t taddr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("tcgen05.st.sync.aligned.16x256b.x1.b32 [%0], {%1,%2,%3,%4};" ::"r"(taddr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
// masks, stores and (optionally) refills one slab: TMEM columns [0, 32) = even k-steps, [32, 64) = odd.
// z2 = the zero points of rows g, g + 8 (two bytes); zaddr = the slab's lanes of the 8-column zero
// operand: every column of a row holds -z * 2^-24 (0x8000 | z read as fp16).
__device__ __forceinline__ void tc_store_slab(uint32_t taddr, uint32_t zaddr, uint4& wa, uint4& wb, uint32_t& z2, bool more,
                                              const uint8_t* refill, const uint8_t* zrefill) {
  uint32_t v[16];
  tc_mask_slab<0x000f000fu>(wa, wb, v);
  st_16x256b_x4(taddr, v);
  tc_mask_slab<0x00f000f0u>(wa, wb, v);
  ldg128_stream_if(more, wa, refill);
  ldg128_stream_if(more, wb, refill + 512);
  st_16x256b_x4(taddr + 32, v);
  const uint32_t za = __byte_perm(z2, 0x80u, 0x4040), zb = __byte_perm(z2, 0x80u, 0x4141);  // one byte permute per value
  ldg16_stream_if(more, z2, zrefill);
  st_16x256b_x1(zaddr, za, za, zb, zb);
}

// BN = UMMA N = token rows of the B operand = token columns drained (16 or 32); kTcGroups convert
// groups (each owns one A operand in tensor memory and one activation stage in shared memory);
// kDepth = records in flight per convert warp; kBuf = operand buffers per group (2: a group
// converts its next record while the MMAs of its previous one are still in flight).  The register
// file is the in-flight buffer: five warps per scheduler (17..20 warps) leave 96 registers per
// thread, four (<= 16 warps) 128.
template <int BN, int kTcGroups, int kDepth, int kBuf>
__global__ void __launch_bounds__(tc_threads(kTcGroups), 1)
gemv_tc_kernel(const __half* __restrict__ xg, const uint8_t* __restrict__ tiles, const __half* __restrict__ bias,
               __half* __restrict__ y, float* __restrict__ ws_partial, int* __restrict__ ws_sem, const TcGeom geo) {
  constexpr int kXBytes = tc_x_bytes(BN);
  constexpr int kXHalf = BN * 128;
  constexpr uint32_t kIdesc = tc::instr_desc_f16(0, 128, BN);
  constexpr int kTmemCols = 512;
  constexpr int kNB = kBuf * kTcGroups;                      // A operands / x stages: tile `it` uses it % kNB
  constexpr int kNF = kNB % 4 == 0 ? kNB : (kNB % 2 == 0 ? 2 * kNB : 4 * kNB);  // "full" barriers: tile `it` uses it % kNF (a multiple of the issuer count)
  constexpr int kND = (8 * BN + kNB * (64 + 8) <= kTmemCols) ? 8 : 4;  // accumulator buffers (BN columns each)
  static_assert(kND * BN + kNB * (64 + 8) <= kTmemCols, "tensor memory budget");
  static_assert(kNF <= 4 * kTcMaxGroups && kNB <= 2 * kTcMaxGroups, "barrier space");
  static_assert(kNF % kTcIssuers == 0 && kND % kTcIssuers == 0, "every barrier meets one issuer");
  // Every mbarrier is waited on phase by phase by ONE waiter (a waiter two phases ahead would see
  // a stale parity): accumulator d belongs to issuer d % 2 (kND is even); the A buffer and x stage
  // of tile `it` are it % kNB and always belong to group it % G; their "full" barrier is it % kNF
  // (an even period), so that barrier too always meets the same issuer.
  static_assert(kND % 2 == 0, "ring sizes");
  static_assert(kTcGroups >= 2 && kTcGroups <= kTcMaxGroups, "convert groups");
  constexpr int kFirstMma = tc_first_mma(kTcGroups);

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t x_base = smem_base;                                     // kNB stages of kXBytes (1024-aligned: SWIZZLE_128B)
  const uint32_t bar_afull = x_base + kNB * kXBytes;
  const uint32_t bar_aempty = bar_afull + 8 * kNF;
  const uint32_t bar_dfull = bar_aempty + 8 * kNB;
  const uint32_t bar_dempty = bar_dfull + 8 * kND;
  const uint32_t tmem_holder = bar_dempty + 8 * kND;

  // the shuffle tells the compiler that the warp index (and every role / loop decision made from it)
  // is warp-uniform: branches, barrier addresses and descriptors stay in the uniform datapath
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
#ifdef AWQ_TC_TRACE
  const unsigned int trace_launch = (unsigned int)geo.trace_seq & 63u;
  if (tid == 0) {
    TC_STAMP_ALL(0);
    TC_STAMP(0);
  }
#endif
  const int grid = gridDim.x;
  const int t_begin = tc_begin(blockIdx.x, grid, geo.ttot);
  const int t_end = tc_begin(blockIdx.x + 1, grid, geo.ttot);
  const int nk = t_end - t_begin;
  const int KT = geo.KT;
  const uint8_t* rec0 = tiles + int64_t(t_begin) * kTileBytes;

  // ---- convert warps: the first weight records are on their way into L2 before anything else
  // happens (weights never depend on the previous kernel) ----
  const bool is_convert = warp >= kTcFirstConvert && warp < kFirstMma;
  const int grp = (warp - kTcFirstConvert) >> 2;
  const int quarter = warp & 3;
  const int g8 = lane >> 2, c4 = lane & 3;
  const uint8_t* wptr = rec0 + int64_t(grp) * kTileBytes + quarter * 2048 + g8 * 64 + 16 * (c4 ^ ((g8 >> 1) & 3));
  if (is_convert && quarter == 0 && lane < kDepth + kTcL2Rounds && grp + kTcGroups * lane < nk)
    bulk_prefetch_l2(rec0 + int64_t(grp + kTcGroups * lane) * kTileBytes, kTileBytes);

  if (tid == 0) {
    for (int a = 0; a < kNF; ++a) mbar_init(bar_afull + 8 * a, 4);   // the 4 convert warps of the group
    for (int a = 0; a < kNB; ++a) mbar_init(bar_aempty + 8 * a, 1);  // tcgen05.commit
    for (int d = 0; d < kND; ++d) {
      mbar_init(bar_dfull + 8 * d, 1);
      mbar_init(bar_dempty + 8 * d, kTcDrainWarps);
    }
    mbar_fence_init();
  }
  if (warp == kFirstMma) {
    tc::tmem_alloc(tmem_holder, kTmemCols);
    tc::tmem_relinquish();
  }
  tc::fence_before_thread_sync();
  __syncthreads();
  tc::fence_after_thread_sync();
  if (tid == 0) TC_STAMP(1);  // setup done (barriers, TMEM)
  const uint32_t tmem_base = lds32(tmem_holder);
  const uint32_t tmem_d = tmem_base;                               // columns [0, ND * BN)
  const uint32_t tmem_a = tmem_base + uint32_t(kND * BN);        // kNB weight operands of 64 columns
  const uint32_t tmem_z = tmem_a + uint32_t(kNB * 64);           // kNB zero-point operands of 8 columns
  pdl_launch_dependents();

  if (warp >= kFirstMma) {
    // =============================== MMA issuers ===============================
    // The issuers take tiles round-robin: their accumulators, A buffers and x stages are disjoint, so
    // no ordering between them is needed, and the barrier round trips of one overlap the issue of
    // the other.  "A full" covers the group's x stage too (same producers, same arrive).
    const uint64_t desc0 = tc::smem_desc_k_sw128(x_base);
    for (int it = warp - kFirstMma; it < nk; it += kTcIssuers) {
      const uint32_t a = uint32_t(it) % kNB, af = uint32_t(it) % kNF, d = uint32_t(it) & (kND - 1);
      mbar_wait_susp(bar_dempty + 8 * d, ((uint32_t(it) / kND) & 1) ^ 1);
      mbar_wait_susp(bar_afull + 8 * af, (uint32_t(it) / kNF) & 1);
      fence_proxy_async_smem();
      tc::fence_after_thread_sync();
      if (tc::elect_one()) {
        const uint64_t bd = desc0 + uint64_t((a * kXBytes) >> 4);
        const uint32_t ta = tmem_a + a * 64, tz = tmem_z + a * 8, td = tmem_d + d * BN;
        // odd quads first (their operands carry 16 w); the first even k-step folds them in at 2^-4;
        // the zero point last: one 8-column operand (-z) against all eight k-blocks of x
#pragma unroll
        for (int i = 4; i < 8; ++i) tc::mma_ts(td, ta + 8 * i, bd + uint64_t((kXHalf + (i & 3) * 32) >> 4), kIdesc, i != 4);
        mma_ts_scale_d4(td, ta, bd, kIdesc);
#pragma unroll
        for (int i = 1; i < 4; ++i)
          tc::mma_ts(td, ta + 8 * i, bd + uint64_t((i * 32) >> 4), kIdesc, true);
#pragma unroll
        for (int i = 0; i < 8; ++i)
          tc::mma_ts(td, tz, bd + uint64_t(((i >> 2) * kXHalf + (i & 3) * 32) >> 4), kIdesc, true);
        tc::commit(bar_aempty + 8 * a);
        tc::commit(bar_dfull + 8 * d);
        if (it == 0) TC_STAMP(8);        // first tile's MMAs issued
        if (it == nk - 1) TC_STAMP(9);   // last tile's MMAs issued
      }
      __syncwarp();
    }
  } else if (is_convert) {
    // =============================== convert ===============================
    // group grp takes tiles grp, grp + G, ...; a warp owns slabs 2 * quarter, 2 * quarter + 1 of each
    // (TMEM lanes 32 * quarter .. + 31).  Registers of a consumed slab are refilled at once with the
    // same slab of the tile kDepth rounds ahead.
    const uint32_t ta0 = tmem_a + (uint32_t(32 * quarter) << 16);  // + 64 * buffer
    const uint32_t tz0 = tmem_z + (uint32_t(32 * quarter) << 16);  // + 8 * buffer
    // zero points of rows g, g + 8 of slab 2 * quarter (+ kTileMetaBlock: slab 2 * quarter + 1)
    const uint8_t* zptr = rec0 + int64_t(grp) * kTileBytes + kTileMetaOff + (2 * quarter) * kTileMetaBlock + 32 + 2 * g8;
    uint4 w[kDepth][4];
    uint32_t z[kDepth][2];  // zero points of rows (g, g + 8) of slab a / slab b, one byte each
#pragma unroll
    for (int u = 0; u < kDepth; ++u) {
      const bool have = grp + kTcGroups * u < nk;
      const uint8_t* p = wptr + int64_t(kTcGroups * u) * kTileBytes;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        w[u][q] = make_uint4(0u, 0u, 0u, 0u);
        ldg128_stream_if(have, w[u][q], p + 512 * q);
      }
      z[u][0] = z[u][1] = 0u;
      ldg16_stream_if(have, z[u][0], zptr + int64_t(kTcGroups * u) * kTileBytes);
      ldg16_stream_if(have, z[u][1], zptr + int64_t(kTcGroups * u) * kTileBytes + kTileMetaBlock);
    }
    // refill pointers: the record kDepth rounds ahead of the one being converted
    const uint8_t* pn = wptr + int64_t(kTcGroups * kDepth) * kTileBytes;
    const int zdelta = int(zptr - wptr);  // zero points of the same record, relative to the thread's weights
    // ---- activations: the group's 128 threads also stage the x tile of each of its records into
    // the group's stage.  Thread t stages the 16-byte piece j = t % 16 (8 consecutive k = one packed
    // word's worth) of tokens t / 16 + 8 rr: the 8 bytes the even-quad k-step wants and the 8 bytes
    // the odd-quad k-step wants go to the two halves of the stage.  Token rows >= M are CLAMPED to
    // row M - 1 rather than predicated (same cache lines; the copies land in operand rows whose
    // accumulator columns nobody reads).  No registers are held for x: each thread prefetches the
    // line of its pieces of the group's NEXT record into L1 (the weights bypass L1, nothing evicts
    // it) and loads this record's pieces at the top of the step, an L1 hit that completes under the
    // masking of the weights.
    constexpr int kCh = BN / 8;
    const int tg = 32 * quarter + lane, jx = tg & 15, cx = jx >> 2, spx = jx & 3, tk = tg >> 4;
    const uint32_t xdst0 = x_base + uint32_t(tk * 128) + 8u * uint32_t(cx & 1) +
                          16u * (uint32_t(2 * spx + (cx >> 1)) ^ uint32_t(tk));
    // one pointer (the thread's piece of token tk in the next record's x tile) + byte offsets of the other tokens
    const int row0 = min(tk, geo.M - 1);
    int rdelta[kCh];
#pragma unroll
    for (int rr = 0; rr < kCh; ++rr) rdelta[rr] = (min(tk + 8 * rr, geo.M - 1) - row0) * geo.K * 2;
    const uint32_t kwrap = uint32_t(geo.K), kadv = uint32_t(kTcGroups % KT) * kTileK;
    uint32_t ko = uint32_t((t_begin + grp) % KT) * kTileK;  // k offset of the next record's x tile
    const uint8_t* xp = reinterpret_cast<const uint8_t*>(xg + int64_t(row0) * geo.K + 8 * jx + ko);
    auto prefetch_x = [&]() {
#pragma unroll
      for (int rr = 0; rr < kCh; ++rr) asm volatile("prefetch.global.L1 [%0];" ::"l"(xp + rdelta[rr]));
    };
    pdl_wait();  // x belongs to the previous kernel until here (the weight loads above do not wait for it)
    if (grp == 0 && quarter == 0 && lane == 0) {
      TC_STAMP(2);  // previous kernel complete
      TC_STAMP_ALL(1);
    }
    if (grp < nk) prefetch_x();
    // the record kDepth + kTcL2Rounds rounds ahead -> L2 (weights for this group's loads, metadata
    // for the drain warps' scale loads)
    // long streams run faster without the steady-state prefetch (measured: 70B gate_up 58 vs 61 us);
    // short ones need it to get going
    const int kL2Ahead = nk < kTcL2MaxTiles ? kTcGroups * (kDepth + kTcL2Rounds) : (1 << 30);
    int it = grp;
    while (it < nk) {
#pragma unroll
      for (int u = 0; u < kDepth; ++u) {
        if (it < nk) {  // warp-uniform
          const bool more = it + kTcGroups * kDepth < nk;
          uint4 xr[kCh];
          {
#pragma unroll
            for (int rr = 0; rr < kCh; ++rr) xr[rr] = ldg128(xp + rdelta[rr]);
            ko += kadv;
            xp += 2 * kadv;
            if (ko >= kwrap) {
              ko -= kwrap;
              xp -= 2 * kwrap;
            }
          }
          const uint32_t b = uint32_t(it) % kNB, af = uint32_t(it) % kNF;
          const uint32_t ta = ta0 + 64 * b, tz = tz0 + 8 * b, xdst = xdst0 + b * kXBytes;
          // the MMAs of the buffer's previous tile have retired (a fresh barrier passes parity 1 at once)
          mbar_wait_susp(bar_aempty + 8 * b, ((uint32_t(it) / kNB) & 1) ^ 1);
          tc::fence_after_thread_sync();
          tc_store_slab(ta, tz, w[u][0], w[u][1], z[u][0], more, pn, pn + zdelta);
          tc_store_slab(ta + (16u << 16), tz + (16u << 16), w[u][2], w[u][3], z[u][1], more, pn + 1024, pn + zdelta + kTileMetaBlock);
          tc_static_for<kCh>([&](auto rc) {
            constexpr int rr = decltype(rc)::value;
            // 32-bit stores: a 64-bit one needs an aligned register pair, i.e. two moves per store
            sts32<rr * 1024>(xdst, xr[rr].x);                // k-step sp:     k 8w + {0,1}
            sts32<rr * 1024 + 4>(xdst, xr[rr].z);            //                k 8w + {4,5}
            sts32<rr * 1024 + kXHalf>(xdst, xr[rr].y);       // k-step 4 + sp: k 8w + {2,3}
            sts32<rr * 1024 + kXHalf + 4>(xdst, xr[rr].w);   //                k 8w + {6,7}
          });
          if (it + kTcGroups < nk) prefetch_x();
          if (quarter == 0 && it + kL2Ahead < nk) {  // warp-uniform condition and address
            if (tc::elect_one()) bulk_prefetch_l2(rec0 + int64_t(it + kL2Ahead) * kTileBytes, kTileBytes);
          }
          fence_proxy_async_smem();  // the stage is read by tcgen05.mma (async proxy)
          tc::wait_st();
          tc::fence_before_thread_sync();
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_afull + 8 * af);
          if (quarter == 0 && lane == 0) {
            if (it == 0) TC_STAMP(3);                     // first operand in tensor memory
            if (it + kTcGroups >= nk) TC_STAMP(4 + grp);  // a group's last operand (slots 4..7)
          }
          it += kTcGroups;
          pn += kTcGroups * kTileBytes;
        }
      }
    }
  } else {
    // =============================== drain + close ===============================
    const int r = 32 * quarter + lane;
    const uint32_t lane_base = uint32_t(32 * quarter) << 16;
    const uint8_t* smeta = rec0 + tile_scale_off(r);
    uint32_t sb[kTcMetaDepth];
#pragma unroll
    for (int u = 0; u < kTcMetaDepth; ++u) {
      sb[u] = 0u;
      if (u < nk) sb[u] = ldg16_stream(smeta + int64_t(u) * kTileBytes);
    }
    float acc[BN];
#pragma unroll
    for (int jj = 0; jj < BN; ++jj) acc[jj] = 0.f;
    int seg = t_begin / KT, kt = t_begin - seg * KT;
    bool waited = false;

    // closes this CTA's part of n-tile `seg` when tile `it` was its last one
    auto step = [&](int it) {
      ++kt;
      if (quarter == 0 && lane == 0) {
        if (it == 0) TC_STAMP(10);       // first accumulator drained
        if (it == nk - 1) TC_STAMP(11);  // last accumulator drained
      }
      if (kt == KT || it + 1 == nk) {
        if (!waited) {
          pdl_wait();  // y / workspace belong to the previous kernel until here
          waited = true;
        }
        const int lo = seg * KT, hi = lo + KT;
        int first = int(unsigned(lo) * unsigned(grid) / unsigned(geo.ttot));
        while (first > 0 && tc_begin(first, grid, geo.ttot) > lo) --first;
        while (tc_begin(first + 1, grid, geo.ttot) <= lo) ++first;
        int last = int(unsigned(hi - 1) * unsigned(grid) / unsigned(geo.ttot));
        if (last >= grid) last = grid - 1;
        while (last + 1 < grid && tc_begin(last + 1, grid, geo.ttot) < hi) ++last;
        while (tc_begin(last, grid, geo.ttot) >= hi) --last;
        const int ncontrib = last - first + 1;
        const int rank = int(blockIdx.x) - first;
        const int n = seg * kTileN + r;
        if (quarter == 0 && lane == 0 && it + 1 == nk) TC_STAMP_ALL(4);  // final closure begins
        const float bias_v = (bias != nullptr && n < geo.N) ? __half2float(bias[n]) : 0.f;
        // SiluAndMul epilogue: the up row of gate row r is TMEM lane r + 8 = lane + 8 of the same warp
        auto store_silu = [&](const float (&val)[BN]) {
          const int NH = geo.N >> 1, col = seg * (kTileN / 2) + (r >> 4) * 8 + (r & 7);
#pragma unroll
          for (int jj = 0; jj < BN; ++jj) {
            const float gq = __half2float(__float2half_rn(val[jj]));  // the unfused sequence rounds gate_up's output first
            const float uq = __shfl_xor_sync(0xffffffffu, gq, 8);
            if ((lane & 8) == 0 && jj < geo.M && col < NH)
              y[int64_t(jj) * NH + col] = __float2half_rn(gq / (1.f + __expf(-gq)) * uq);
          }
        };
        if (ncontrib == 1) {
          if (geo.silu) {
            store_silu(acc);
          } else if (n < geo.N) {
#pragma unroll
            for (int jj = 0; jj < BN; ++jj)
              if (jj < geo.M) y[int64_t(jj) * geo.N + n] = __float2half_rn(acc[jj] + bias_v);
          }
        } else {
          // Split n-tile: closed by its OWNER, the contributor with the first k-range (rank 0).  That CTA reaches the
          // n-tile at the very end of its own stream (its range ENDS inside it), while the others start with it or
          // hold nothing else, so the owner is the last to arrive anyway -- and as owner it needs no partial store, no
          // release fence and no ticket round trip of its own: it waits for the arrival count (normally already
          // there), pulls the other partials and adds them to its registers in rank order.  The others publish
          // [row][BN floats] at slot first + seg + rank and signal with a release-add; nobody waits for a return value.
          int* sem = ws_sem + seg * kTcDrainWarps + quarter;
          if (rank != 0) {
            float4* part = reinterpret_cast<float4*>(ws_partial) + (int64_t(first + seg + rank) * kTileN + r) * (BN / 4);
#pragma unroll
            for (int q = 0; q < BN / 4; ++q) __stcg(part + q, make_float4(acc[4 * q], acc[4 * q + 1], acc[4 * q + 2], acc[4 * q + 3]));
            __syncwarp();
            if (lane == 0) asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(sem) : "memory");
            if (quarter == 0 && lane == 0 && it + 1 == nk) TC_STAMP_ALL(5);  // published
          } else {
            if (lane == 0) {
              unsigned have = 0, spins = 0;
              do {
                asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(have) : "l"(sem) : "memory");
                if (++spins > (1u << 28)) __trap();  // a contributor never arrived: fail the launch, never store a partial sum
              } while (int(have) < ncontrib - 1);
            }
            __syncwarp();
            if (quarter == 0 && lane == 0 && it + 1 == nk) TC_STAMP_ALL(5);  // everybody has arrived
            const float4* base = reinterpret_cast<const float4*>(ws_partial) + (int64_t(first + seg) * kTileN + r) * (BN / 4);
            // kU contributors' loads are in flight together; added in rank order (rank 0 = the registers)
            constexpr int kU = 1;
#pragma unroll 1
            for (int rr0 = 1; rr0 < ncontrib; rr0 += kU) {
              float4 f[kU][BN / 4];
#pragma unroll
              for (int u = 0; u < kU; ++u) {
                const float4* pp = base + int64_t(rr0 + u < ncontrib ? rr0 + u : rr0) * kTileN * (BN / 4);
#pragma unroll
                for (int q = 0; q < BN / 4; ++q) f[u][q] = __ldcg(pp + q);
              }
#pragma unroll
              for (int u = 0; u < kU; ++u) {
                if (rr0 + u < ncontrib) {
#pragma unroll
                  for (int q = 0; q < BN / 4; ++q) {
                    acc[4 * q] += f[u][q].x;
                    acc[4 * q + 1] += f[u][q].y;
                    acc[4 * q + 2] += f[u][q].z;
                    acc[4 * q + 3] += f[u][q].w;
                  }
                }
              }
            }
            if (geo.silu) {
              store_silu(acc);
            } else if (n < geo.N) {
#pragma unroll
              for (int jj = 0; jj < BN; ++jj)
                if (jj < geo.M) y[int64_t(jj) * geo.N + n] = __float2half_rn(acc[jj] + bias_v);
            }
            __syncwarp();
            if (quarter == 0 && lane == 0 && it + 1 == nk) TC_STAMP_ALL(6);  // owner: summed and stored
            if (lane == 0) *sem = 0;
          }
        }
#pragma unroll
        for (int jj = 0; jj < BN; ++jj) acc[jj] = 0.f;
        if (kt == KT) {
          kt = 0;
          ++seg;
        }
      }
    };

    auto load_meta = [&](int u, int it) { sb[u] = ldg16_stream(smeta + int64_t(it) * kTileBytes); };
    // acc[16 ch ..] += sf * v for one 16-column chunk of one tile (v = 2^-24 * Sum (w - z) x, exact)
    auto fold = [&](int ch, const uint32_t (&v)[16], float sf) {
#pragma unroll
      for (int q = 0; q < 16; ++q) acc[16 * ch + q] = fmaf(sf, __uint_as_float(v[q]), acc[16 * ch + q]);
    };
    for (int it0 = 0; it0 < nk; it0 += kTcMetaDepth) {
      if constexpr (BN == 16) {
        // two tiles per round: one barrier / tcgen05.ld / arrive round trip serves both
#pragma unroll
        for (int u = 0; u < kTcMetaDepth; u += 2) {
          const int it = it0 + u;  // even
          if (it < nk) {           // warp-uniform
            const bool two = it + 1 < nk;
            const float sf0 = bits16_as_half_to_float(sb[u]) * 16777216.f;  // the operands carry 2^-24
            const float sf1 = bits16_as_half_to_float(sb[u + 1]) * 16777216.f;
            if (it + kTcMetaDepth < nk) load_meta(u, it + kTcMetaDepth);
            if (it + 1 + kTcMetaDepth < nk) load_meta(u + 1, it + 1 + kTcMetaDepth);
            const uint32_t d0 = uint32_t(it) & (kND - 1), par = (uint32_t(it) / kND) & 1;  // it + 1 shares the parity
            mbar_wait_susp(bar_dfull + 8 * d0, par);
            if (two) mbar_wait_susp(bar_dfull + 8 * (d0 + 1), par);
            tc::fence_after_thread_sync();
            uint32_t v0[16], v1[16];
            tc::ld_32x32b_x16(tmem_d + lane_base + d0 * BN, v0);
            if (two) tc::ld_32x32b_x16(tmem_d + lane_base + (d0 + 1) * BN, v1);
            tc::wait_ld();
            tc::fence_before_thread_sync();
            __syncwarp();
            if (lane == 0) {
              mbar_arrive(bar_dempty + 8 * d0);
              if (two) mbar_arrive(bar_dempty + 8 * (d0 + 1));
            }
            fold(0, v0, sf0);
            step(it);
            if (two) {
              fold(0, v1, sf1);
              step(it + 1);
            }
          }
        }
      } else {
#pragma unroll
        for (int u = 0; u < kTcMetaDepth; ++u) {
          const int it = it0 + u;
          if (it < nk) {  // warp-uniform
            const float sf = bits16_as_half_to_float(sb[u]) * 16777216.f;
            if (it + kTcMetaDepth < nk) load_meta(u, it + kTcMetaDepth);
            const uint32_t d = uint32_t(it) & (kND - 1);
            mbar_wait_susp(bar_dfull + 8 * d, (uint32_t(it) / kND) & 1);
            tc::fence_after_thread_sync();
#pragma unroll
            for (int ch = 0; ch < BN / 16; ++ch) {
              uint32_t v[16];
              tc::ld_32x32b_x16(tmem_d + lane_base + uint32_t(d * BN + 16 * ch), v);
              tc::wait_ld();
              fold(ch, v, sf);
            }
            tc::fence_before_thread_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_dempty + 8 * d);
            step(it);
          }
        }
      }
    }
  }

  // ---- teardown ----
  if (tid == 0) {
    TC_STAMP(12);  // drain warp 0 done (outputs / partials written)
    TC_STAMP_ALL(2);
  }
  tc::fence_before_thread_sync();
  __syncthreads();
  if (tid == 0) TC_STAMP(13);  // all roles done
  if (warp == kFirstMma) {
    tc::fence_after_thread_sync();
    tc::tmem_dealloc(tmem_base, kTmemCols);
    if (lane == 0) {
      TC_STAMP(14);
      TC_STAMP_ALL(3);
    }
  }
}

}  // namespace awq

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
namespace awq {

__host__ __device__ constexpr int tc_smem_bytes(int bn, int stages) {
  return stages * tc_x_bytes(bn) + 1024 /*align slack*/ + 8 * (6 * kTcMaxGroups + 2 * kTcND) + 16;
}

static unsigned long long* g_tc_trace = nullptr;
static int g_tc_trace_seq = 0;
static int g_tc_ctas = 0;     // debug: grid cap (0 = one CTA per SM)
static int g_tc_variant = 0;  // debug: (groups, records in flight per warp, buffers per group), see gemv_tc below
void gemv_tc_debug(int ctas, int variant) {
  g_tc_ctas = ctas;
  g_tc_variant = variant < 0 ? 0 : variant;
}

bool gemv_tc_fits(int64_t M, int64_t K, int64_t N, int dtype) {
  if (dtype != AWQ_B200_F16 || M < 1 || M > 32) return false;
  const int64_t S = (N + kTileN - 1) / kTileN;
  return S * kTcDrainWarps <= kTcSemInts && S * (K / kTileK) < (int64_t(1) << 30);
}

// scratch for problems up to (max_M, max_N): semaphores + one partial slot per (n-tile, contributor)
int64_t gemv_tc_workspace_bytes(int64_t max_M, int64_t max_N) {
  const int64_t S = (max_N + kTileN - 1) / kTileN;
  const int64_t bn = max_M <= 16 ? 16 : 32;
  return int64_t(kTcSemInts) * 4 + (1024 + S + 1) * kTileN * bn * 4;
}

template <int BN, int G, int kDepth, int kBuf>
static int gemv_tc_launch(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K, int64_t N,
                          void* workspace, int64_t workspace_bytes, cudaStream_t st, int silu) {
  auto kern = gemv_tc_kernel<BN, G, kDepth, kBuf>;
  constexpr int smem = tc_smem_bytes(BN, G * kBuf);
  static_assert(smem <= 227 * 1024, "shared memory budget");
  static bool attr_done[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && !attr_done[dev]) {
    AWQ_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem),
                   "awq_b200_gemv: cudaFuncSetAttribute (tc)");
    attr_done[dev] = true;
  }
  TcGeom geo;
  geo.M = int(M);
  geo.K = int(K);
  geo.N = int(N);
  geo.KT = int(K / kTileK);
  geo.S = int((N + kTileN - 1) / kTileN);
  geo.ttot = geo.S * geo.KT;
  geo.silu = silu;
#ifdef AWQ_TC_TRACE
  geo.trace = g_tc_trace;
  geo.trace_seq = g_tc_trace != nullptr ? g_tc_trace_seq++ : 0;
#endif
  int grid = sm_count();
  if (g_tc_ctas > 0) grid = g_tc_ctas;
  if (grid > geo.ttot) grid = geo.ttot;
  AWQ_CHECK(int64_t(geo.ttot) * (grid + 1) < (int64_t(1) << 31), "awq_b200_gemv: problem too large");
  const int64_t need = int64_t(kTcSemInts) * 4 + int64_t(grid + geo.S + 1) * kTileN * BN * 4;
  AWQ_CHECK(workspace != nullptr && workspace_bytes >= need,
            "awq_b200_gemv: workspace too small (%lld < %lld bytes); size it with awq_b200_workspace_bytes",
            (long long)workspace_bytes, (long long)need);
  int* sem = reinterpret_cast<int*>(workspace);
  float* partial = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(workspace) + kTcSemInts * 4);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(unsigned(grid));
  cfg.blockDim = dim3(tc_threads(G));
  cfg.dynamicSmemBytes = size_t(smem);
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  AWQ_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, (const __half*)x, (const uint8_t*)tiles, (const __half*)bias, (__half*)y,
                                    partial, sem, geo),
                 "awq_b200_gemv: launch (tc)");
  return 0;
}

int gemv_tc(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K, int64_t N, int dtype,
            void* workspace, int64_t workspace_bytes, cudaStream_t st, int silu) {
  AWQ_CHECK(gemv_tc_fits(M, K, N, dtype), "awq_b200_gemv: tensor-core decode kernel needs fp16 and 1 <= M <= 32");
  if (M <= 16) {
    if (g_tc_variant == 1) return gemv_tc_launch<16, 3, 2, 1>(x, tiles, bias, y, M, K, N, workspace, workspace_bytes, st, silu);
    if (g_tc_variant == 2) return gemv_tc_launch<16, 3, 3, 2>(x, tiles, bias, y, M, K, N, workspace, workspace_bytes, st, silu);
    if (g_tc_variant == 3) return gemv_tc_launch<16, 4, 2, 1>(x, tiles, bias, y, M, K, N, workspace, workspace_bytes, st, silu);
    return gemv_tc_launch<16, 3, 2, 2>(x, tiles, bias, y, M, K, N, workspace, workspace_bytes, st, silu);
  }
  if (g_tc_variant == 3) return gemv_tc_launch<32, 4, 2, 1>(x, tiles, bias, y, M, K, N, workspace, workspace_bytes, st, silu);
  if (g_tc_variant == 2) return gemv_tc_launch<32, 3, 3, 1>(x, tiles, bias, y, M, K, N, workspace, workspace_bytes, st, silu);
  return gemv_tc_launch<32, 3, 2, 1>(x, tiles, bias, y, M, K, N, workspace, workspace_bytes, st, silu);
}

}  // namespace awq

/* development aid: device buffer of >= (16 + 64 * 16 + 4 * 160 * 8) u64 (zeroed by the caller) that CTA 0 of every tcgen05 decode
 * launch fills with %globaltimer stamps of its phases (scripts/trace_tc.py); NULL switches it off */
extern "C" void awq_b200_debug_set_tc_trace(void* buf) {
  awq::g_tc_trace = reinterpret_cast<unsigned long long*>(buf);
  awq::g_tc_trace_seq = 0;
}

/* ---- allocation-free decode entry point (no counterpart in the reference; added
 * for callers that cannot call cudaMalloc: code under CUDA-graph capture, tight serving loops, harnesses
 * that forbid it).  The tcgen05 decode kernel above with its split-n-tile scratch (arrival counters + fp32 partial
 * slots) in a file-scope device array instead of a caller workspace.  Every GPU that loads the library holds the
 * array (AWQ_B200_STATIC_SCRATCH_BYTES) until the process exits; every call shares it, so two calls that overlap
 * on the GPU (different streams, or graph branches) race.  The name of the array is a host-side stand-in in host
 * code: its device address is fetched once per device with cudaGetSymbolAddress. ---- */
namespace awq {
constexpr int64_t kTcStaticScratchBytes = AWQ_B200_STATIC_SCRATCH_BYTES;
static_assert(kTcStaticScratchBytes > int64_t(kTcSemInts) * 4, "static scratch must hold the arrival counters");
static __device__ __align__(256) unsigned char g_tc_static_scratch[kTcStaticScratchBytes];

// device address of g_tc_static_scratch on the CURRENT device (set-up: once per device)
static int tc_static_scratch_ptr(void** out) {
  static std::mutex mu;
  static void* ptr[64] = {nullptr};
  int dev = 0;
  AWQ_CUDA_CHECK(cudaGetDevice(&dev), "awq_b200_gemv_static: cudaGetDevice");
  AWQ_CHECK(dev >= 0 && dev < 64, "awq_b200_gemv_static: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (ptr[dev] == nullptr) {
    void* p = nullptr;
    AWQ_CUDA_CHECK(cudaGetSymbolAddress(&p, g_tc_static_scratch), "awq_b200_gemv_static: cudaGetSymbolAddress");
    AWQ_CHECK(p != nullptr && ((uintptr_t)p % 256) == 0, "awq_b200_gemv_static: bad scratch address");
    ptr[dev] = p;
  }
  *out = ptr[dev];
  return 0;
}
}  // namespace awq

extern "C" int awq_b200_gemv_static(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K,
                                    int64_t N, int dtype, void* stream) {
  using namespace awq;
  if (M == 0 || N == 0) return 0;
  AWQ_CHECK(dtype == AWQ_B200_F16, "awq_b200_gemv_static: fp16 only (dtype %d)", dtype);
  AWQ_CHECK(M >= 1 && M <= 32, "awq_b200_gemv_static: M (%lld) must be in [1, 32]", (long long)M);
  AWQ_CHECK(K > 0 && K % kTileK == 0, "awq_b200_gemv_static: K (%lld) must be a positive multiple of 128", (long long)K);
  AWQ_CHECK(N > 0 && N % 8 == 0, "awq_b200_gemv_static: N (%lld) must be a positive multiple of 8", (long long)N);
  AWQ_CHECK(x && tiles && y, "awq_b200_gemv_static: null pointer");
  AWQ_CHECK(((uintptr_t)x % 16) == 0 && ((uintptr_t)tiles % 16) == 0 && ((uintptr_t)y % 2) == 0,
            "awq_b200_gemv_static: x and tiles must be 16-byte aligned");
  AWQ_CHECK(K * N / 2 < (int64_t(1) << 40), "awq_b200_gemv_static: problem too large");
  if (int rc = awq_b200_check_device()) return rc;
  // the scratch the kernel needs for this problem on this device (gemv_tc_launch): counters + one fp32 slot of
  // 128 x BN per CTA and per n-tile
  const int64_t S = (N + kTileN - 1) / kTileN;
  const int64_t bn = M <= 16 ? 16 : 32;
  int64_t grid = sm_count();
  if (grid > S * (K / kTileK)) grid = S * (K / kTileK);
  const int64_t need = int64_t(kTcSemInts) * 4 + (grid + S + 1) * kTileN * bn * 4;
  AWQ_CHECK(need <= kTcStaticScratchBytes,
            "awq_b200_gemv_static: M = %lld, N = %lld needs %lld bytes of scratch, beyond the static maximum of %lld "
            "(AWQ_B200_STATIC_SCRATCH_BYTES); use awq_b200_gemv with a caller workspace",
            (long long)M, (long long)N, (long long)need, (long long)kTcStaticScratchBytes);
  AWQ_CHECK(gemv_tc_fits(M, K, N, dtype), "awq_b200_gemv_static: N (%lld) too large for the tcgen05 decode kernel", (long long)N);
  void* scratch = nullptr;
  if (int rc = tc_static_scratch_ptr(&scratch)) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  // the arrival counters must start at zero; the array keeps whatever the last call (possibly an aborted one) left
  AWQ_CUDA_CHECK(cudaMemsetAsync(scratch, 0, size_t(kTcSemInts) * 4, st), "awq_b200_gemv_static: clearing the counters");
  return gemv_tc(x, tiles, bias, y, M, K, N, dtype, scratch, kTcStaticScratchBytes, st, 0);
}
