// awq_gemv_tc.cu -- kernel 2c: decode path on the 5th-gen tensor cores (fp16, 1 <= M <= 32).
//
//   y[M,N] = x[M,K] @ W (+ bias)
//
// Replaces the three launches of the reference's AWQLinearMethod.apply
// (python/sglang/srt/layers/quantization/awq.py:434-451) for decode batches.  The op is
// HBM-bound; what the mma.sync decode kernels run out of is ISSUE SLOTS (ncu, profiles/): every
// HMMA, every activation-fragment load and every in-flight weight byte costs the streaming
// warps an instruction or a register.  Here
//   * weights travel HBM -> shared memory as whole 8576-byte tile records by bulk copy (one
//     instruction per record, issued by one thread; bytes in flight live in shared memory);
//   * the convert warps turn nibbles into tensor-core operands with 9 ALU ops per 8 weights:
//     a nibble in the low mantissa bits of an fp16 SUBNORMAL is the exact value w * 2^-24; the
//     zero point is subtracted in the same representation ((w - z) * 2^-24, exact), so the A
//     operand needs no scale and no correction term; it is written straight into tensor memory
//     (tcgen05.st), thread = output row;
//   * one thread issues tcgen05.mma (M = 128 rows, N = 16/32 tokens, K = 16 per instruction);
//     activations come from shared memory (TMA, 128-byte swizzle) -- no fragment loads;
//   * every 128-k tile (= one AWQ group) gets a fresh fp32 accumulator in tensor memory; the
//     drain warps read it back (tcgen05.ld), apply the group's scale in fp32 and keep the
//     running sum in registers;
//   * work split = stream-K over the (n-tile, k-tile) records in memory order, one CTA per SM,
//     split n-tiles closed through the fp32 workspace in RANK order (deterministic);
//   * the CTA uses < half of an SM (shared memory, 256 TMEM columns, registers) so that under
//     programmatic dependent launch the NEXT kernel's CTAs become resident and start their
//     weight stream while this one drains (scripts/micro/chain_lat.cu).
#include "tcgen05.cuh"
#include "host_common.h"

#include <cuda.h>

namespace awq {

constexpr int kTcGroups = 3;          // convert groups of four warps (one warp per TMEM lane quarter); group g owns A buffer g
constexpr int kTcConvertWarps = 4 * kTcGroups;
constexpr int kTcDrainWarps = 4;
constexpr int kTcProducers = 3;       // producer warp p loads tiles p, p + 3, ... (stage s is always filled by warp s % 3)
constexpr int kTcFirstMma = kTcProducers;          // two MMA issuer warps (even / odd tiles)
constexpr int kTcFirstConvert = kTcFirstMma + 2;   // a group = 4 consecutive warps = 4 distinct TMEM lane quarters (warp % 4)
constexpr int kTcFirstDrain = kTcFirstConvert + kTcConvertWarps;
constexpr int kTcThreads = 32 * (kTcFirstDrain + kTcDrainWarps);
constexpr int kTcNAMax = 2 * kTcGroups;  // A buffers in TMEM (64 columns each): one or two per convert group
constexpr int kTcNDMax = 8;          // accumulator buffers in TMEM (BN columns each): 4 or 8
constexpr int kTcSemInts = 8192;
constexpr int kTcScaleSlots = 8;     // per-tile scale side array (floats, [slot][row]); reuse distance >= 4 tiles
constexpr int kTcNAF = 2 * kTcGroups;  // afull barriers: index it % 6 -> always the same issuer (it % 2) waits on it

#ifdef AWQ_TRACE
// development build only (scripts/build_trace.sh): per-tile SM-clock stamps of the first CTAs
__device__ long long* g_tc_trace = nullptr;
#define TCT(role, it)                                                                          \
  do {                                                                                         \
    if (g_tc_trace != nullptr && blockIdx.x < 4 && (it) < 256)                                 \
      g_tc_trace[(size_t(blockIdx.x) * 16 + (role)) * 256 + (it)] = clock64();                 \
  } while (0)
#else
#define TCT(role, it) do {} while (0)
#endif

struct TcGeom {
  int M, K, N, KT, S, ttot;
};

// first tile of CTA i (host guarantees ttot * (grid + 1) < 2^31)
__host__ __device__ __forceinline__ int tc_begin(int i, int grid, int ttot) {
  return int(unsigned(i) * unsigned(ttot) / unsigned(grid));
}

__host__ __device__ constexpr int tc_x_bytes(int xr) { return 2 * xr * 128; }
__host__ __device__ constexpr int tc_stage_bytes(int xr) { return ((tc_x_bytes(xr) + kTileBytes + 1023) / 1024) * 1024; }

// XR = token rows per TMA box (8, 16, 32); BN = UMMA N (16 or 32); MV = token columns drained;
// kMinB = CTAs per SM the register budget is sized for (2: a second kernel can be co-resident).
// kMinB == 1 (whole-SM footprint): every convert group owns TWO A buffers, so it converts tile
// it + 3 while the MMAs of tile it run (with one buffer the group's convert -> tcgen05.st ->
// mbarrier -> mma -> commit chain is serial: ~2150 clocks per tile per group, measured with
// scripts/trace_tc.py), and there are 8 accumulators (BN = 16) in flight.
template <int XR, int BN, int MV, int kStages, int kMinB>
__global__ void __launch_bounds__(kTcThreads, kMinB)
gemv_tc_kernel(const void* __restrict__ xsrc, const uint8_t* __restrict__ tiles,
               const __half* __restrict__ bias, __half* __restrict__ y, float* __restrict__ ws_partial,
               int* __restrict__ ws_sem, const TcGeom geo) {
  constexpr int kStageBytes = tc_stage_bytes(XR);
  constexpr int kXBytes = tc_x_bytes(XR);
  constexpr int kXHalf = XR * 128;
  constexpr uint32_t kIdesc = tc::instr_desc_f16(0, 128, BN);
  constexpr bool kBulk = kMinB == 1;
  constexpr int kNA = kMinB == 1 ? 2 * kTcGroups : kTcGroups;
  constexpr int kND = (kMinB == 1 && 8 * BN + kNA * 64 <= 512) ? 8 : 4;
  constexpr int kTmemCols = (kND * BN + kNA * 64 <= 256) ? 256 : 512;
  static_assert(kND * BN + kNA * 64 <= 512 && kND <= kTcNDMax && kND <= kTcScaleSlots, "tensor memory budget");
  // Every mbarrier must be waited on phase by phase by ONE waiter (a waiter two phases ahead
  // would see a stale parity): stage s is always converted by group s % 3, afull barrier
  // it % 6 is always waited on by issuer it % 2.
  static_assert(kStages >= 6 && kStages % kTcGroups == 0 && kStages % kTcProducers == 0, "the ring must be a multiple of the convert groups");

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sc_base = smem_base + kStages * kStageBytes;            // [slots][128] floats
  const uint32_t bar_base = sc_base + kTcScaleSlots * 128 * 4;
  const uint32_t bar_full = bar_base;
  const uint32_t bar_empty = bar_full + 8 * kStages;
  const uint32_t bar_afull = bar_empty + 8 * kStages;
  const uint32_t bar_aempty = bar_afull + 8 * kTcNAF;
  const uint32_t bar_dfull = bar_aempty + 8 * kTcNAMax;
  const uint32_t bar_dempty = bar_dfull + 8 * kTcNDMax;
  const uint32_t tmem_holder = bar_dempty + 8 * kTcNDMax;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int grid = gridDim.x;
  const int t_begin = tc_begin(blockIdx.x, grid, geo.ttot);
  const int t_end = tc_begin(blockIdx.x + 1, grid, geo.ttot);
  const int nk = t_end - t_begin;
  const int KT = geo.KT;

  if (tid == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(bar_full + 8 * s, 32);      // the 32 producer lanes (cp.async arrive-on-completion)
      mbar_init(bar_empty + 8 * s, 1 + 4);  // tcgen05.commit (x consumed) + the 4 convert warps of the tile
    }
    for (int a = 0; a < kTcNAF; ++a) mbar_init(bar_afull + 8 * a, 4);
    for (int a = 0; a < kNA; ++a) mbar_init(bar_aempty + 8 * a, 1);
    for (int d = 0; d < kND; ++d) {
      mbar_init(bar_dfull + 8 * d, 1);
      mbar_init(bar_dempty + 8 * d, kTcDrainWarps);
    }
    mbar_fence_init();
  }
  if (warp == kTcFirstMma) {
    tc::tmem_alloc(tmem_holder, kTmemCols);
    tc::tmem_relinquish();
  }
  tc::fence_before_thread_sync();
  __syncthreads();
  tc::fence_after_thread_sync();
  const uint32_t tmem_base = lds32(tmem_holder);
  const uint32_t tmem_d = tmem_base;                          // columns [0, ND * BN)
  const uint32_t tmem_a = tmem_base + uint32_t(kND * BN);     // columns [ND * BN, ND * BN + NA * 64)
  pdl_launch_dependents();

  if (warp < kTcProducers) {
    // =============================== producers ===============================
    // The whole warp copies with 16-byte cp.async (LDGSTS, LSU path).  Measured on B200
    // (profiles/): the per-SM TMA engine sustains ~24 B/clk plus ~40 clocks per request, which
    // a record + two activation boxes per tile exceeds; the LSU path has no such limit and
    // only the M real activation rows of a tile are fetched (rows >= M of the B operand may
    // hold anything: they only reach accumulator columns that are never read).
    const uint64_t pol_w = policy_evict_first();
    const uint8_t* wsrc = tiles + int64_t(t_begin) * kTileBytes + 16 * lane;
    const int M = geo.M;
    const __half* xg = reinterpret_cast<const __half*>(xsrc);
    auto issue_w = [&](int it, uint32_t stage) {
      const uint32_t st = smem_base + stage * kStageBytes;
      if constexpr (kBulk) {
        // whole-SM ring: ONE bulk copy per record.  The LDGSTS stream below tops out near 4.3 TB/s
        // chip-wide (trace: a record lands ~9400 clocks after its 17 x 32 requests were issued); one
        // 8576-byte request per tile keeps the bulk engine under its ~22 B/clk/SM limit
        // (profiles/r01_bulk_copy_microbench.txt) and x stays on the LSU path beside it.
        if (lane == 0) {
          mbar_expect_tx(bar_full + 8 * stage, kTileBytes);
          bulk_g2s_hint(st + kXBytes, tiles + (int64_t(t_begin) + it) * kTileBytes, kTileBytes, bar_full + 8 * stage, pol_w);
        }
      } else {
        const uint8_t* src = wsrc + int64_t(it) * kTileBytes;
        const uint32_t dst = st + kXBytes + 16 * lane;
#pragma unroll
        for (int i = 0; i < 16; ++i) cp_async16_hint(dst + 512 * i, src + 512 * i, pol_w);
        if (lane < (kTileBytes - 16 * 512) / 16) cp_async16_hint(dst + 512 * 16, src + 512 * 16, pol_w);
      }
    };
    auto issue_x = [&](int kt, uint32_t st) {
      // item = (row j, 16-byte chunk cc of the row's 256 bytes): half h = cc >> 3, chunk c = cc & 7
      for (int item = lane; item < 16 * M; item += 32) {
        const int j = item >> 4, cc = item & 15, h = cc >> 3, c = cc & 7;
        cp_async16_plain(st + uint32_t(h * kXHalf + j * 128 + 16 * (c ^ (j & 7))), xg + int64_t(j) * geo.K + kt * kTileK + 8 * cc);
      }
    };
    const int pre = nk < kStages ? nk : kStages;
    for (int it = warp; it < pre; it += kTcProducers) issue_w(it, uint32_t(it));  // weights do not depend on the previous kernel
    pdl_wait();
    int kt = (t_begin + warp) % KT;
    const int kstep = kTcProducers % KT;
    int it = warp;
    for (; it < pre; it += kTcProducers) {
      issue_x(kt, smem_base + it * kStageBytes);
      cp_async_mbar_arrive_noinc(bar_full + 8 * it);
      kt += kstep;
      if (kt >= KT) kt -= KT;
    }
    uint32_t s = uint32_t(it) - kStages, ph = 1;  // stage of tile `it` (it >= kStages here), parity of its previous use
    for (; it < nk; it += kTcProducers) {
      mbar_wait_sleep(bar_empty + 8 * s, ph ^ 1);
      const uint32_t st = smem_base + s * kStageBytes;
      issue_w(it, s);
      issue_x(kt, st);
      cp_async_mbar_arrive_noinc(bar_full + 8 * s);
      if (lane == 0) TCT(0, it);
      kt += kstep;
      if (kt >= KT) kt -= KT;
      s += kTcProducers;
      if (s >= uint32_t(kStages)) {
        s -= kStages;
        ph ^= 1;
      }
    }
  } else if (warp == kTcFirstMma || warp == kTcFirstMma + 1) {
    // =============================== MMA issuers ===============================
    // Two issuers take alternate tiles: their accumulators and A buffers are disjoint, so no
    // ordering between them is needed, and the per-tile barrier round trip of one overlaps the
    // issue of the other.  ONE wait per tile: afull[a] is only signalled by convert warps that
    // have seen full[s] (weights AND x of the stage landed) and dempty[d] (accumulator d drained).
    const uint64_t desc0 = tc::smem_desc_k_sw128(smem_base);
    int it = warp - kTcFirstMma;
    uint32_t s = uint32_t(it), a = uint32_t(it), q = 0;  // s = it % kStages, a = it % 3, q = it / 3
    for (; it < nk; it += 2) {
      const uint32_t af = a + kTcGroups * (q & 1);          // = it % 6
      const uint32_t ab = kNA == kTcGroups ? a : af;        // A buffer of tile it
      mbar_wait_sleep(bar_afull + 8 * af, (q >> 1) & 1);
      fence_proxy_async_smem();  // the activations were written through the generic proxy (cp.async)
      tc::fence_after_thread_sync();
      if (tc::elect_one()) {
        TCT(4, it);
        const uint32_t d = uint32_t(it) & (kND - 1);
        const uint64_t bd = desc0 + uint64_t((s * kStageBytes) >> 4);
        const uint32_t ta = tmem_a + ab * 64, td = tmem_d + d * BN;
#pragma unroll
        for (int j = 0; j < 8; ++j)
          tc::mma_ts(td, ta + 8 * j, bd + uint64_t(((j >> 2) * kXHalf + (j & 3) * 32) >> 4), kIdesc, j != 0);
        tc::commit(bar_empty + 8 * s);
        tc::commit(bar_aempty + 8 * ab);
        tc::commit(bar_dfull + 8 * d);
        TCT(5, it);
      }
      __syncwarp();
      s += 2;
      if (s >= uint32_t(kStages)) s -= kStages;
      a += 2;
      if (a >= uint32_t(kTcGroups)) {
        a -= kTcGroups;
        ++q;
      }
    }
  } else if (warp >= kTcFirstConvert && warp < kTcFirstDrain) {
    // =============================== convert ===============================
    const int grp = (warp - kTcFirstConvert) >> 2;  // converts tiles grp, grp + 3, ... into A buffer grp
    const int quarter = warp & 3;
    const int r = 32 * quarter + lane;
    const uint32_t lane_base = uint32_t(32 * quarter) << 16;
    const uint32_t row_off = uint32_t(r) * 64u;
    const uint32_t swz = uint32_t((r >> 1) & 3);
    const uint32_t s_off = uint32_t(tile_scale_off(r));
    const uint32_t z_off = uint32_t(tile_zero_off(r));
    const uint32_t a_addr0 = tmem_a + lane_base + uint32_t(grp * 64);
    uint32_t s = uint32_t(grp), ph_s = 0, j = 0;
    if (grp < nk) mbar_wait_sleep(bar_full + 8 * s, 0);
    for (int it = grp; it < nk; it += kTcGroups, ++j) {
      // the record of tile `it` is in shared memory
      const uint32_t wbase = smem_base + s * kStageBytes + kXBytes;
      if (quarter == 0 && lane == 0) TCT(1, it);
      uint4 gq[4];
#pragma unroll
      for (int c = 0; c < 4; ++c) gq[c] = lds128(wbase + row_off + 16u * (uint32_t(c) ^ swz));
      const uint32_t sbits = lds16(wbase + s_off);
      const uint32_t z = lds8(wbase + z_off);
      // A buffer free (the MMAs of tile it - 3 retired) and, on behalf of the MMA issuer,
      // accumulator it % 4 drained (tile it - 4)
      const uint32_t ab = kNA == kTcGroups ? uint32_t(grp) : uint32_t(grp) + kTcGroups * (j & 1);  // = it % kNA
      const uint32_t use = kNA == kTcGroups ? j : (j >> 1);                                          // = it / kNA
      const uint32_t a_addr = a_addr0 + (ab - uint32_t(grp)) * 64;
      mbar_wait_sleep(bar_aempty + 8 * ab, (use & 1) ^ 1);
      mbar_wait_sleep(bar_dempty + 8 * (it & (kND - 1)), ((it / kND) & 1) ^ 1);
      tc::fence_after_thread_sync();
      if (quarter == 0 && lane == 0) TCT(2, it);
      const uint32_t nz2 = 0x80008000u | (z * 0x00010001u);  // {-z * 2^-24} x2
#pragma unroll
      for (int c = 0; c < 4; ++c) {  // granule c = words 4c..4c+3 = k 32c..32c+31 = TMEM columns 16c..16c+15
        const uint32_t words[4] = {gq[c].x, gq[c].y, gq[c].z, gq[c].w};
        uint32_t av[16];
#pragma unroll
        for (int w = 0; w < 4; ++w) {
          const uint32_t lo = words[w], hi = words[w] >> 8;
          av[4 * w + 0] = h2_add(lo & 0x000f000fu, nz2);                 // k 8w+0,1   (w   - z) * 2^-24
          av[4 * w + 1] = h2_fma(lo & 0x00f000f0u, 0x2c002c00u, nz2);    // k 8w+2,3   (16w / 16 - z) * 2^-24
          av[4 * w + 2] = h2_add(hi & 0x000f000fu, nz2);                 // k 8w+4,5
          av[4 * w + 3] = h2_fma(hi & 0x00f000f0u, 0x2c002c00u, nz2);    // k 8w+6,7
        }
        tc::st_32x32b_x16(a_addr + 16 * c, av);
      }
      __syncwarp();
      if (quarter == 0 && lane == 0) TCT(9, it);
      if (lane == 0) mbar_arrive(bar_empty + 8 * s);  // this warp no longer needs the record
      // group scale (x 2^24: the operands carry 2^-24) for the drain warp of this row
      sts32(sc_base + uint32_t(((it & (kTcScaleSlots - 1)) * 128 + r) * 4),
            __float_as_uint(bits16_as_half_to_float(sbits) * 16777216.f));
      tc::wait_st();
      if (quarter == 0 && lane == 0) TCT(10, it);
      tc::fence_before_thread_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_afull + 8 * (grp + kTcGroups * (j & 1)));
      if (quarter == 0 && lane == 0) TCT(3, it);
      s += kTcGroups;
      if (s >= uint32_t(kStages)) {
        s -= kStages;
        ph_s ^= 1;
      }
      if (it + kTcGroups < nk) mbar_wait_sleep(bar_full + 8 * s, ph_s);
    }
  } else if (warp >= kTcFirstDrain) {
    // =============================== drain + close ===============================
    // Two tiles per round: one barrier / tcgen05.ld / arrive round trip serves both.
    const int quarter = warp & 3;
    const int r = 32 * quarter + lane;
    const uint32_t lane_base = uint32_t(32 * quarter) << 16;
    float acc[MV];
#pragma unroll
    for (int j = 0; j < MV; ++j) acc[j] = 0.f;
    int seg = t_begin / KT, kt = t_begin - seg * KT;
    bool waited = false;

    // closes this CTA's part of n-tile `seg` when tile `it` was its last one
    auto step = [&](int it) {
      ++kt;
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
        const float bias_v = (bias != nullptr && n < geo.N) ? __half2float(bias[n]) : 0.f;
        if (ncontrib == 1) {
          if (n < geo.N) {
#pragma unroll
            for (int j = 0; j < MV; ++j)
              if (j < geo.M) y[int64_t(j) * geo.N + n] = __float2half_rn(acc[j] + bias_v);
          }
        } else {
          // partial slot: compact index first + seg + rank (unique per (n-tile, contributor)); [row][MV floats]
          float4* part = reinterpret_cast<float4*>(ws_partial) + (int64_t(first + seg + rank) * kTileN + r) * (MV / 4);
#pragma unroll
          for (int q = 0; q < MV / 4; ++q) __stcg(part + q, make_float4(acc[4 * q], acc[4 * q + 1], acc[4 * q + 2], acc[4 * q + 3]));
          __syncwarp();
          unsigned ticket = 0;
          int* sem = ws_sem + seg * kTcDrainWarps + quarter;
          if (lane == 0)
            asm volatile("atom.add.acq_rel.gpu.global.u32 %0, [%1], 1;" : "=r"(ticket) : "l"(sem) : "memory");
          ticket = __shfl_sync(0xffffffffu, ticket, 0);
          if (int(ticket) == ncontrib - 1) {
            // last arriver for this (n-tile, quarter): sum the partials in rank order
            float sum[MV];
#pragma unroll
            for (int j = 0; j < MV; ++j) sum[j] = 0.f;
            const float4* base = reinterpret_cast<const float4*>(ws_partial) + (int64_t(first + seg) * kTileN + r) * (MV / 4);
#pragma unroll 1
            for (int rr = 0; rr < ncontrib; ++rr) {
              const float4* pp = base + int64_t(rr) * kTileN * (MV / 4);
#pragma unroll
              for (int q = 0; q < MV / 4; ++q) {
                const float4 f = __ldcg(pp + q);
                sum[4 * q] += f.x;
                sum[4 * q + 1] += f.y;
                sum[4 * q + 2] += f.z;
                sum[4 * q + 3] += f.w;
              }
            }
            if (n < geo.N) {
#pragma unroll
              for (int j = 0; j < MV; ++j)
                if (j < geo.M) y[int64_t(j) * geo.N + n] = __float2half_rn(sum[j] + bias_v);
            }
            __syncwarp();
            if (lane == 0) *sem = 0;
          }
        }
#pragma unroll
        for (int j = 0; j < MV; ++j) acc[j] = 0.f;
        if (kt == KT) {
          kt = 0;
          ++seg;
        }
      }
    };

    for (int it = 0; it < nk; it += 2) {
      const bool two = it + 1 < nk;               // warp-uniform
      const uint32_t d0 = uint32_t(it) & (kND - 1), par = uint32_t(it / kND) & 1;  // it is even: it + 1 shares the parity
      mbar_wait_sleep(bar_dfull + 8 * d0, par);
      if (two) mbar_wait_sleep(bar_dfull + 8 * (d0 + 1), par);
      tc::fence_after_thread_sync();
      if (quarter == 0 && lane == 0) TCT(6, it);
      const float s0 = __uint_as_float(lds32(sc_base + uint32_t(((it & (kTcScaleSlots - 1)) * 128 + r) * 4)));
      const float s1 = __uint_as_float(lds32(sc_base + uint32_t((((it + 1) & (kTcScaleSlots - 1)) * 128 + r) * 4)));
      if constexpr (MV <= 8) {
        // both accumulators in flight at once
        uint32_t v0[8], v1[8];
        tc::ld_32x32b_x8(tmem_d + lane_base + uint32_t(d0 * BN), v0);
        if (two) tc::ld_32x32b_x8(tmem_d + lane_base + uint32_t((d0 + 1) * BN), v1);
        tc::wait_ld();
        tc::fence_before_thread_sync();
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(bar_dempty + 8 * d0);
          if (two) mbar_arrive(bar_dempty + 8 * (d0 + 1));
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = fmaf(s0, __uint_as_float(v0[j]), acc[j]);
        step(it);
        if (two) {
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[j] = fmaf(s1, __uint_as_float(v1[j]), acc[j]);
          step(it + 1);
        }
      } else {
        // register budget: eight columns at a time, one barrier round for the pair of tiles
#pragma unroll
        for (int ch = 0; ch < MV / 8; ++ch) {
          uint32_t v[8];
          tc::ld_32x32b_x8(tmem_d + lane_base + uint32_t(d0 * BN + 8 * ch), v);
          tc::wait_ld();
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[8 * ch + j] = fmaf(s0, __uint_as_float(v[j]), acc[8 * ch + j]);
        }
        step(it);
        if (two) {
#pragma unroll
          for (int ch = 0; ch < MV / 8; ++ch) {
            uint32_t v[8];
            tc::ld_32x32b_x8(tmem_d + lane_base + uint32_t((d0 + 1) * BN + 8 * ch), v);
            tc::wait_ld();
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[8 * ch + j] = fmaf(s1, __uint_as_float(v[j]), acc[8 * ch + j]);
          }
        }
        tc::fence_before_thread_sync();
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(bar_dempty + 8 * d0);
          if (two) mbar_arrive(bar_dempty + 8 * (d0 + 1));
        }
        if (two) step(it + 1);
      }
      if (quarter == 0 && lane == 0) TCT(7, it);
    }
  }

  // ---- teardown ----
  tc::fence_before_thread_sync();
  __syncthreads();
  if (warp == kTcFirstMma) {
    tc::fence_after_thread_sync();
    tc::tmem_dealloc(tmem_base, kTmemCols);
  }
}

}  // namespace awq

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
namespace awq {

__host__ __device__ constexpr int tc_smem_bytes(int xr, int stages) {
  return stages * tc_stage_bytes(xr) + 1024 /*align slack*/ + kTcScaleSlots * 128 * 4 +
         8 * (2 * stages + kTcNAF + kTcNAMax + 2 * kTcNDMax) + 16;
}

#ifdef AWQ_TRACE
extern "C" void awq_b200_debug_set_tc_trace(void* buf) {
  long long* p = (long long*)buf;
  cudaMemcpyToSymbol(g_tc_trace, &p, sizeof(p));
}
#endif

static int g_tc_ctas = 0;     // debug: grid cap (0 = one CTA per SM)
static int g_tc_variant = 0;  // 0 = by tiles per CTA, 1 = half-SM footprint (co-resident under PDL), 2 = whole-SM ring
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
  const int64_t mv = max_M <= 8 ? 8 : (max_M <= 16 ? 16 : 32);
  return int64_t(kTcSemInts) * 4 + (1024 + S + 1) * kTileN * mv * 4;
}

template <int XR, int BN, int MV, int kStages, int kMinB>
static int gemv_tc_launch(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K, int64_t N,
                          void* workspace, int64_t workspace_bytes, cudaStream_t st) {
  auto kern = gemv_tc_kernel<XR, BN, MV, kStages, kMinB>;
  constexpr int smem = tc_smem_bytes(XR, kStages);
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
  int grid = sm_count();
  if (g_tc_ctas > 0) grid = g_tc_ctas;
  if (grid > geo.ttot) grid = geo.ttot;
  AWQ_CHECK(int64_t(geo.ttot) * (grid + 1) < (int64_t(1) << 31), "awq_b200_gemv: problem too large");
  const int64_t need = int64_t(kTcSemInts) * 4 + int64_t(grid + geo.S + 1) * kTileN * MV * 4;
  AWQ_CHECK(workspace != nullptr && workspace_bytes >= need,
            "awq_b200_gemv: workspace too small (%lld < %lld bytes); size it with awq_b200_workspace_bytes",
            (long long)workspace_bytes, (long long)need);
  int* sem = reinterpret_cast<int*>(workspace);
  float* partial = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(workspace) + kTcSemInts * 4);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(unsigned(grid));
  cfg.blockDim = dim3(kTcThreads);
  cfg.dynamicSmemBytes = size_t(smem);
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  AWQ_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, x, (const uint8_t*)tiles, (const __half*)bias, (__half*)y, partial, sem, geo),
                 "awq_b200_gemv: launch (tc)");
  return 0;
}

int gemv_tc(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K, int64_t N, int dtype,
            void* workspace, int64_t workspace_bytes, cudaStream_t st) {
  AWQ_CHECK(gemv_tc_fits(M, K, N, dtype), "awq_b200_gemv: tensor-core decode kernel needs fp16 and 1 <= M <= 32");
  // Two footprints.  Half an SM (few stages, <= 46 registers): the next kernel's CTAs become
  // resident under programmatic dependent launch and start their weight stream while this one
  // drains -- right for launches of a few tiles per CTA.  Whole SM (ring of ~200 KB): the bytes
  // in flight per SM that a long stream needs to hold the HBM rate (Little: ~2 us x 43 GB/s/SM).
  const int64_t tiles_per_cta = ((N + kTileN - 1) / kTileN) * (K / kTileK) / sm_count();
  const bool whole = g_tc_variant == 2 || (g_tc_variant == 0 && tiles_per_cta >= 24);
  if (M <= 8) {
    if (whole) return gemv_tc_launch<8, 16, 8, 18, 1>(x, tiles, bias, y, M, K, N, workspace, workspace_bytes, st);
    return gemv_tc_launch<8, 16, 8, 9, 2>(x, tiles, bias, y, M, K, N, workspace, workspace_bytes, st);
  }
  if (M <= 16) {
    if (whole) return gemv_tc_launch<16, 16, 16, 15, 1>(x, tiles, bias, y, M, K, N, workspace, workspace_bytes, st);
    return gemv_tc_launch<16, 16, 16, 6, 2>(x, tiles, bias, y, M, K, N, workspace, workspace_bytes, st);
  }
  if (whole) return gemv_tc_launch<32, 32, 32, 12, 1>(x, tiles, bias, y, M, K, N, workspace, workspace_bytes, st);
  return gemv_tc_launch<32, 32, 32, 9, 1>(x, tiles, bias, y, M, K, N, workspace, workspace_bytes, st);
}

}  // namespace awq
