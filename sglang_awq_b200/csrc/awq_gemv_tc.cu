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
//     instruction per record, issued by one thread; ~200 KB in flight per SM live in shared
//     memory, released as soon as the convert warps hold the record in registers);
//   * the convert warps turn nibbles into tensor-core operands with 9 ALU ops per 8 weights:
//     a nibble in the low mantissa bits of an fp16 SUBNORMAL is the exact value w * 2^-24; the
//     zero point is subtracted in the same representation ((w - z) * 2^-24, exact), so the A
//     operand needs no scale and no correction term; it is written straight into tensor memory
//     (tcgen05.st), thread = output row, two A buffers per convert group;
//   * two threads issue tcgen05.mma for alternate tiles (M = 128 rows, N = 16/32 tokens, K = 16
//     per instruction); activations come from a second, shallow shared-memory ring (cp.async,
//     128-byte swizzle) -- no fragment loads;
//   * every 128-k tile (= one AWQ group) gets a fresh fp32 accumulator in tensor memory; the
//     drain warps read it back (tcgen05.ld), apply the group's scale in fp32 and keep the
//     running sum in registers;
//   * work split = stream-K over the (n-tile, k-tile) records in memory order, one CTA per SM,
//     split n-tiles closed through the fp32 workspace in RANK order (deterministic).
// Status (profiles/r01d_tc_m16_70b.txt, scripts/trace_tc.py): ~600 clocks per tile where the HBM
// rate allows 386; the limit is instruction issue in the convert warps (alu pipe: 85 LOP3/SHF per
// warp-tile at 2 clocks each), not memory, barriers or tensor memory.
#include "tcgen05.cuh"
#include "host_common.h"

#include <cuda.h>

namespace awq {

constexpr int kTcGroups = 3;          // convert groups of four warps (one warp per TMEM lane quarter); group g owns A buffer g
constexpr int kTcConvertWarps = 4 * kTcGroups;
constexpr int kTcDrainWarps = 4;
constexpr int kTcProducers = 3;       // warp 0: weight records (one bulk copy per tile); warps 1, 2: activations of even / odd tiles
constexpr int kTcFirstMma = kTcProducers;          // two MMA issuer warps (even / odd tiles)
constexpr int kTcFirstConvert = kTcFirstMma + 2;   // a group = 4 consecutive warps = 4 distinct TMEM lane quarters (warp % 4)
constexpr int kTcFirstDrain = kTcFirstConvert + kTcConvertWarps;
constexpr int kTcThreads = 32 * (kTcFirstDrain + kTcDrainWarps);
constexpr int kTcNAMax = 2 * kTcGroups;  // A buffers in TMEM (64 columns each): one or two per convert group
constexpr int kTcNDMax = 8;          // accumulator buffers in TMEM (BN columns each): 4 or 8
constexpr int kTcSemInts = 8192;
constexpr int kTcScaleSlots = 16;     // per-tile scale side array (floats, [slot][row]); a slot is rewritten 16 tiles later
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

// XR = token rows of the B operand (8, 16, 32); BN = UMMA N (16 or 32); MV = token columns drained.
// Two rings, because their time scales differ by 5x (measured with scripts/trace_tc.py and ncu):
//   * kW weight stages of exactly one record.  A record is in flight for the LOADED DRAM latency
//     (2 - 3 us: the chip needs > 100 KB per SM in flight to hold the HBM rate,
//     profiles/r01_bulk_copy_microbench.txt) and is released as soon as the convert warps have
//     read it into registers -- not after the MMAs, which never touch it.
//   * kX activation stages (L2 hits, ~1 us), filled a few tiles ahead of the MMA issuers and
//     released by tcgen05.commit.
// Every convert group owns TWO A buffers in tensor memory, so it converts tile it + 3 while the
// MMAs of tile it run, and there are 8 accumulators (BN = 16) in flight.
template <int XR, int BN, int MV, int kW, int kX>
__global__ void __launch_bounds__(kTcThreads, 1)
gemv_tc_kernel(const void* __restrict__ xsrc, const uint8_t* __restrict__ tiles,
               const __half* __restrict__ bias, __half* __restrict__ y, float* __restrict__ ws_partial,
               int* __restrict__ ws_sem, const TcGeom geo) {
  constexpr int kXBytes = tc_x_bytes(XR);
  constexpr int kXHalf = XR * 128;
  constexpr uint32_t kIdesc = tc::instr_desc_f16(0, 128, BN);
  constexpr int kNA = 2 * kTcGroups;
  constexpr int kND = (8 * BN + kNA * 64 <= 512) ? 8 : 4;
  constexpr int kTmemCols = (kND * BN + kNA * 64 <= 256) ? 256 : 512;
  static_assert(kND * BN + kNA * 64 <= 512 && kND <= kTcNDMax && kND <= kTcScaleSlots, "tensor memory budget");
  // Every mbarrier must be waited on phase by phase by ONE waiter (a waiter two phases ahead
  // would see a stale parity): weight stage s is always converted by group s % 3, activation
  // stage s is always filled by producer 1 + s % 2 and consumed by issuer s % 2, afull barrier
  // it % 6 is always waited on by issuer it % 2.
  static_assert(kW >= 6 && kW % kTcGroups == 0 && kX >= 4 && kX % 2 == 0, "ring sizes");

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t x_base = smem_base;                                     // kX stages of kXBytes (1024-aligned: SWIZZLE_128B)
  const uint32_t w_base = x_base + kX * kXBytes;                         // kW stages of one record
  const uint32_t sc_base = w_base + kW * kTileBytes;                     // [slots][128] floats
  const uint32_t bar_base = sc_base + kTcScaleSlots * 128 * 4;
  const uint32_t bar_wfull = bar_base;
  const uint32_t bar_wempty = bar_wfull + 8 * kW;
  const uint32_t bar_xfull = bar_wempty + 8 * kW;
  const uint32_t bar_xempty = bar_xfull + 8 * kX;
  const uint32_t bar_afull = bar_xempty + 8 * kX;
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
    for (int s = 0; s < kW; ++s) {
      mbar_init(bar_wfull + 8 * s, 1);   // arrive.expect_tx of the issuing lane; the bulk copy completes the bytes
      mbar_init(bar_wempty + 8 * s, 4);  // the 4 convert warps of the tile
    }
    for (int s = 0; s < kX; ++s) {
      mbar_init(bar_xfull + 8 * s, 32);  // the 32 producer lanes (cp.async arrive-on-completion)
      mbar_init(bar_xempty + 8 * s, 1);  // tcgen05.commit
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

  if (warp == 0) {
    // =============================== weight producer ===============================
    // ONE bulk copy per 8576-byte record, issued by one lane.  (Sixteen-byte LDGSTS requests top
    // out near 4.3 TB/s chip-wide, and several bulk requests per tile exceed the bulk engine's
    // ~22 B/clk/SM; one request per tile stays under it.)  Weights never depend on the previous
    // kernel, so this role does not execute griddepcontrol.wait.
    if (lane == 0) {
      const uint64_t pol_w = policy_evict_first();
      const uint8_t* src = tiles + int64_t(t_begin) * kTileBytes;
      uint32_t s = 0, ph = 0;
      for (int it = 0; it < nk; ++it) {
        if (it >= kW) mbar_wait_susp(bar_wempty + 8 * s, ph ^ 1);
        mbar_arrive_expect_tx(bar_wfull + 8 * s, kTileBytes);
        bulk_g2s_hint(w_base + s * kTileBytes, src + int64_t(it) * kTileBytes, kTileBytes, bar_wfull + 8 * s, pol_w);
        TCT(0, it);
        if (++s == uint32_t(kW)) {
          s = 0;
          ph ^= 1;
        }
      }
    }
  } else if (warp < kTcProducers) {
    // =============================== activation producers ===============================
    // 16-byte cp.async (LDGSTS, LSU path) into the 128-byte-swizzled K-major B tile; only the M
    // real rows of a tile are fetched (rows >= M of the B operand may hold anything: they only
    // reach accumulator columns that are never read).  Producer p takes tiles p - 1, p + 1, ...
    const int M = geo.M;
    const __half* xg = reinterpret_cast<const __half*>(xsrc);
    auto issue_x = [&](int kt, uint32_t st) {
      // item = (row j, 16-byte chunk cc of the row's 256 bytes): half h = cc >> 3, chunk c = cc & 7
      for (int item = lane; item < 16 * M; item += 32) {
        const int j = item >> 4, cc = item & 15, h = cc >> 3, c = cc & 7;
        cp_async16_plain(st + uint32_t(h * kXHalf + j * 128 + 16 * (c ^ (j & 7))), xg + int64_t(j) * geo.K + kt * kTileK + 8 * cc);
      }
    };
    pdl_wait();  // x belongs to the previous kernel until here
    int it = warp - 1;
    int kt = (t_begin + it) % KT;
    const int kstep = 2 % KT;
    uint32_t s = uint32_t(it), ph = 0;
    for (; it < nk; it += 2) {
      if (it >= kX) mbar_wait_susp(bar_xempty + 8 * s, ph ^ 1);
      issue_x(kt, x_base + s * kXBytes);
      cp_async_mbar_arrive_noinc(bar_xfull + 8 * s);
      kt += kstep;
      if (kt >= KT) kt -= KT;
      s += 2;
      if (s >= uint32_t(kX)) {
        s -= kX;
        ph ^= 1;
      }
    }
  } else if (warp == kTcFirstMma || warp == kTcFirstMma + 1) {
    // =============================== MMA issuers ===============================
    // Two issuers take alternate tiles: their accumulators and A buffers are disjoint, so no
    // ordering between them is needed, and the per-tile barrier round trip of one overlaps the
    // issue of the other.  afull[a] is only signalled by convert warps that have also seen
    // dempty[d] (accumulator d drained); xfull[s] says the activations of the tile landed.
    const uint64_t desc0 = tc::smem_desc_k_sw128(x_base);
    int it = warp - kTcFirstMma;
    uint32_t s = uint32_t(it), xph = 0, a = uint32_t(it), q = 0;  // s = it % kX, a = it % 3, q = it / 3
    for (; it < nk; it += 2) {
      const uint32_t af = a + kTcGroups * (q & 1);          // = it % 6
      const uint32_t ab = af;                               // A buffer of tile it
      mbar_wait_susp(bar_xfull + 8 * s, xph);
      mbar_wait_susp(bar_afull + 8 * af, (q >> 1) & 1);
      fence_proxy_async_smem();  // the activations were written through the generic proxy (cp.async)
      tc::fence_after_thread_sync();
      if (tc::elect_one()) {
        TCT(4, it);
        const uint32_t d = uint32_t(it) & (kND - 1);
        const uint64_t bd = desc0 + uint64_t((s * kXBytes) >> 4);
        const uint32_t ta = tmem_a + ab * 64, td = tmem_d + d * BN;
#pragma unroll
        for (int j = 0; j < 8; ++j)
          tc::mma_ts(td, ta + 8 * j, bd + uint64_t(((j >> 2) * kXHalf + (j & 3) * 32) >> 4), kIdesc, j != 0);
        tc::commit(bar_xempty + 8 * s);
        tc::commit(bar_aempty + 8 * ab);
        tc::commit(bar_dfull + 8 * d);
        TCT(5, it);
      }
      __syncwarp();
      s += 2;
      if (s >= uint32_t(kX)) {
        s -= kX;
        xph ^= 1;
      }
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
    if (grp < nk) mbar_wait_susp(bar_wfull + 8 * s, 0);
    for (int it = grp; it < nk; it += kTcGroups, ++j) {
      // the record of tile `it` is in shared memory
      const uint32_t wbase = w_base + s * kTileBytes;
      if (quarter == 0 && lane == 0) TCT(1, it);
      uint4 gq[4];
#pragma unroll
      for (int c = 0; c < 4; ++c) gq[c] = lds128(wbase + row_off + 16u * (uint32_t(c) ^ swz));
      const uint32_t sbits = lds16(wbase + s_off);
      const uint32_t z = lds8(wbase + z_off);
      // group scale (x 2^24: the operands carry 2^-24) for the drain warp of this row.  Slot reuse
      // distance 16 > 11: this group waited for the drain of tile it - 11 before its previous tile.
      sts32(sc_base + uint32_t(((it & (kTcScaleSlots - 1)) * 128 + r) * 4),
            __float_as_uint(bits16_as_half_to_float(sbits) * 16777216.f));
      // The record is in registers: give the stage back before anything else (the ring has to
      // cover the loaded DRAM latency, not the conversion).  The store above needs `sbits`, the
      // loads return in issue order, and the volatile asm statements keep their order, so every
      // read of the stage has completed when the arrive issues.
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_wempty + 8 * s);
      // A buffer free (the MMAs of tile it - 6 retired) and, on behalf of the MMA issuer,
      // accumulator it % kND drained
      const uint32_t ab = uint32_t(grp) + kTcGroups * (j & 1);  // = it % kNA
      const uint32_t use = j >> 1;                              // = it / kNA
      const uint32_t a_addr = a_addr0 + (ab - uint32_t(grp)) * 64;
      mbar_wait_susp(bar_aempty + 8 * ab, (use & 1) ^ 1);
      mbar_wait_susp(bar_dempty + 8 * (it & (kND - 1)), ((it / kND) & 1) ^ 1);
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
      if (quarter == 0 && lane == 0) TCT(9, it);
      tc::wait_st();
      if (quarter == 0 && lane == 0) TCT(10, it);
      tc::fence_before_thread_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_afull + 8 * (grp + kTcGroups * (j & 1)));
      if (quarter == 0 && lane == 0) TCT(3, it);
      s += kTcGroups;
      if (s >= uint32_t(kW)) {
        s -= kW;
        ph_s ^= 1;
      }
      if (it + kTcGroups < nk) mbar_wait_susp(bar_wfull + 8 * s, ph_s);
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
      mbar_wait_susp(bar_dfull + 8 * d0, par);
      if (two) mbar_wait_susp(bar_dfull + 8 * (d0 + 1), par);
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

__host__ __device__ constexpr int tc_smem_bytes(int xr, int kw, int kx) {
  return kx * tc_x_bytes(xr) + kw * kTileBytes + 1024 /*align slack*/ + kTcScaleSlots * 128 * 4 +
         8 * (2 * kw + 2 * kx + kTcNAF + kTcNAMax + 2 * kTcNDMax) + 16;
}

#ifdef AWQ_TRACE
extern "C" void awq_b200_debug_set_tc_trace(void* buf) {
  long long* p = (long long*)buf;
  cudaMemcpyToSymbol(g_tc_trace, &p, sizeof(p));
}
#endif

static int g_tc_ctas = 0;     // debug: grid cap (0 = one CTA per SM)
void gemv_tc_debug(int ctas, int variant) {
  (void)variant;  // one instantiation per M range
  g_tc_ctas = ctas;
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

template <int XR, int BN, int MV, int kW, int kX>
static int gemv_tc_launch(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K, int64_t N,
                          void* workspace, int64_t workspace_bytes, cudaStream_t st) {
  auto kern = gemv_tc_kernel<XR, BN, MV, kW, kX>;
  constexpr int smem = tc_smem_bytes(XR, kW, kX);
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
  // the CTA owns the SM: ~200 KB of weight records in flight (the loaded DRAM latency is 2 - 3 us)
  if (M <= 8) return gemv_tc_launch<8, 16, 8, 24, 6>(x, tiles, bias, y, M, K, N, workspace, workspace_bytes, st);
  if (M <= 16) return gemv_tc_launch<16, 16, 16, 21, 8>(x, tiles, bias, y, M, K, N, workspace, workspace_bytes, st);
  return gemv_tc_launch<32, 32, 32, 18, 8>(x, tiles, bias, y, M, K, N, workspace, workspace_bytes, st);
}

}  // namespace awq
