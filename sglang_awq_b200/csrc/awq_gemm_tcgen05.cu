// awq_gemm_tcgen05.cu -- kernel 3: y[M,N] = x[M,K] @ W (+ bias) on the 5th-gen tensor cores.
//
// Replaces awq_dequantize + torch.matmul (cuBLAS) + add_ of the reference's
// AWQLinearMethod.apply (python/sglang/srt/layers/quantization/awq.py:434-451) for M where the
// op is a real contraction: no [K,N] fp16 matrix is ever written to or re-read from HBM.
//
// "Swap-AB" formulation: D^T[n, m] = sum_k W[k, n] x[m, k].
//   * A operand = dequantised weights, 128 output features (TMEM lanes) x 16 k per MMA, written
//     by the dequant warps DIRECTLY into tensor memory with tcgen05.st: thread = one output
//     row, so it needs one scale and one zero per tile, and the tile layout (awq_repack.cu)
//     makes each 16-byte granule of its row come out as 16 consecutive TMEM columns.  The
//     fp16/bf16 values are bit-identical to the reference's dequantised weights.
//   * B operand = activations, BN tokens x 16 k, K-major with 128-byte swizzle, fetched by TMA
//     (cp.async.bulk.tensor.2d) straight from x[M, K]; rows past M are zero-filled by TMA.
//   * D = fp32 accumulator, 128 lanes x BN columns of TMEM; tcgen05.mma.kind::f16 issued by ONE
//     thread; completion tracked with tcgen05.commit -> mbarrier.
// Warp roles (14 warps): 0 = producer (one bulk copy of the 8576-byte tile record + two TMA
// boxes per 128-k stage), 1 = TMEM allocator + MMA issuer, 2..13 = three dequant groups of four
// warps (one per TMEM lane quarter).  Group g converts tiles it % 3 == g into its own TMEM A
// buffer, so three tiles are in flight and the st -> mbarrier -> mma -> commit latency chain
// of one tile overlaps the conversion of the next two.  At the end of a segment the same
// warps run the epilogue (tcgen05.ld -> bias -> fp16/bf16 -> y), column chunks split by group.
// Work split: the flattened (n-tile, m-tile, k-tile) sequence is cut into one contiguous range
// per CTA (stream-K).  With many (n,m) segments (prefill) ranges are aligned to whole segments;
// with few (decode-sized M) they cut through K and partial accumulators are closed through the
// fp32 workspace in RANK order (deterministic), as in awq_gemv.cu.
#include "tcgen05.cuh"
#include "host_common.h"

#include <cuda.h>

namespace awq {

constexpr int kGemmGroups = 3;                      // dequant warp groups; group g converts tiles it % 3 == g
constexpr int kGemmDequantWarps = 4 * kGemmGroups;  // one warp per TMEM lane quarter per group
constexpr int kGemmThreads = 32 * (2 + kGemmDequantWarps);
constexpr int kGemmSemInts = 8192;
constexpr int kABufCols = 64;   // one 128-k tile of A = 128 lanes x 64 columns (2 x 16-bit per column)
constexpr int kNumABufs = kGemmGroups;              // each group owns one A buffer in TMEM

// sized for the largest record (group size 32: 9728 B) so that one instantiation serves every group size
__host__ __device__ constexpr int gemm_stage_bytes(int bn) { return ((bn * 256 + kTileBytesMax + 1023) / 1024) * 1024; }
__host__ __device__ constexpr int gemm_stages(int bn) {
  return bn <= 16 ? 15 : (bn <= 32 ? 12 : (bn <= 64 ? 8 : (bn <= 128 ? 5 : 3)));
}
__host__ __device__ constexpr int gemm_tmem_cols(int bn) {
  return (bn + kNumABufs * kABufCols) <= 256 ? 256 : 512;  // 256 + 3 * 64 = 448 <= 512
}

struct GemmGeom {
  int M, K, N, KT, MT, NT, S;   // S = segments = NT * MT; segment seg = (m-tile seg / NT, n-tile seg % NT)
  int ttot;                 // S * KT
  int align;                // 1: CTA ranges are whole segments
  int silu;                 // gate_up tiles: y[M, N / 2] = silu(gate) * up (rows r, r + 8 of a 16-row slab)
  int sub;                  // metadata sections per record: 1 (group size >= 128), 2 (64), 4 (32)
};

// first tile of CTA i (32-bit math; host guarantees no overflow)
__host__ __device__ __forceinline__ int gemm_begin(int i, int grid, const GemmGeom& g) {
  if (g.align) return int(unsigned(i) * unsigned(g.S) / unsigned(grid)) * g.KT;
  return int(unsigned(i) * unsigned(g.ttot) / unsigned(grid));
}

template <typename T, int BN>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmap_x, const uint8_t* __restrict__ tiles,
            const T* __restrict__ bias, T* __restrict__ y, float* __restrict__ ws_partial,
            int* __restrict__ ws_sem, const GemmGeom geo) {
  constexpr int kStages = gemm_stages(BN);
  constexpr int kStageBytes = gemm_stage_bytes(BN);
  constexpr int kXBytes = BN * 256;
  constexpr int kTmemCols = gemm_tmem_cols(BN);
  constexpr uint32_t kIdesc = tc::instr_desc_f16(Elem<T>::kDtype, 128, BN);
  const int rec_bytes = tile_bytes_sub(geo.sub);
  const uint32_t kTxBytes = uint32_t(rec_bytes + kXBytes);

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // SWIZZLE_128B needs 1024-byte aligned stage bases; dynamic smem is only guaranteed 16
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar_base = smem_base + kStages * kStageBytes;
  const uint32_t bar_full = bar_base;
  const uint32_t bar_empty = bar_full + 8 * kStages;
  const uint32_t bar_afull = bar_empty + 8 * kStages;
  const uint32_t bar_aempty = bar_afull + 8 * kNumABufs;
  const uint32_t bar_accfull = bar_aempty + 8 * kNumABufs;
  const uint32_t bar_accempty = bar_accfull + 8;
  const uint32_t tmem_holder = bar_accempty + 8;

  // the shuffle tells the compiler that the warp index (and every role decision made from it) is
  // warp-uniform: loop control, barrier addresses and descriptors stay in the uniform datapath
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
  const int grid = gridDim.x;
  const int t_begin = gemm_begin(blockIdx.x, grid, geo);
  const int t_end = gemm_begin(blockIdx.x + 1, grid, geo);
  const int nk = t_end - t_begin;
  const int KT = geo.KT, NT = geo.NT;

  if (tid == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(bar_full + 8 * s, 1);
      mbar_init(bar_empty + 8 * s, 1 + 4);  // tcgen05.commit + the 4 warps of the group that converts it
    }
    for (int a = 0; a < kNumABufs; ++a) {
      mbar_init(bar_afull + 8 * a, 4);
      mbar_init(bar_aempty + 8 * a, 1);
    }
    mbar_init(bar_accfull, 1);
    mbar_init(bar_accempty, kGemmDequantWarps);
    mbar_fence_init();
    tc::prefetch_tmap(&tmap_x);
  }
  if (warp == 1) {
    tc::tmem_alloc(tmem_holder, kTmemCols);
    tc::tmem_relinquish();
  }
  tc::fence_before_thread_sync();
  __syncthreads();
  tc::fence_after_thread_sync();
  const uint32_t tmem_base = lds32(tmem_holder);
  const uint32_t tmem_acc = tmem_base;              // columns [0, BN)
  const uint32_t tmem_a = tmem_base + uint32_t(BN); // columns [BN, BN + 128)
  pdl_launch_dependents();

  if (warp == 0) {
    // =============================== producer ===============================
    if (lane == 0 && nk > 0) {
      const uint64_t pol_w = policy_evict_first();
      // coordinates of the next tile to issue, advanced incrementally (no divisions in the loop)
      int seg0 = t_begin / KT;
      // Segments are ordered n-tile fastest: CTAs that run side by side work on the same token
      // tile, so the x tiles (7/8 of the bytes a CTA pulls from L2, which is what bounds this
      // kernel: ~6300 B/clk for the whole chip) are requested by many SMs at about the same time.
      int p_kt = t_begin - seg0 * KT, p_mt = seg0 / NT, p_nt = seg0 - (seg0 / NT) * NT;
      const uint8_t* wp = tiles + (int64_t(p_nt) * KT + p_kt) * rec_bytes;  // record of (p_nt, p_kt)
      auto advance = [&]() {
        ++p_kt;
        wp += rec_bytes;
        if (p_kt == KT) {  // next segment: the next n-tile (records continue) or the first one of the next token tile
          p_kt = 0;
          ++p_nt;
          if (p_nt == NT) {
            p_nt = 0;
            ++p_mt;
            wp = tiles;
          }
        }
      };
      auto issue_w = [&](int s) {
        bulk_g2s_hint(smem_base + s * kStageBytes + kXBytes, wp, uint32_t(rec_bytes), bar_full + 8 * s, pol_w);
      };
      auto issue_x = [&](int s, int kt_, int mt_) {
        const uint32_t st = smem_base + s * kStageBytes;
        tc::tma_load_2d(st, &tmap_x, kt_ * kTileK, mt_ * BN, bar_full + 8 * s);
        tc::tma_load_2d(st + BN * 128, &tmap_x, kt_ * kTileK + 64, mt_ * BN, bar_full + 8 * s);
      };
      const int pre = nk < kStages ? nk : kStages;
      int pre_kt[kStages], pre_mt[kStages];
      for (int it = 0; it < pre; ++it) {  // weights do not depend on the previous kernel
        mbar_arrive_expect_tx(bar_full + 8 * it, kTxBytes);
        issue_w(it);
        pre_kt[it] = p_kt;
        pre_mt[it] = p_mt;
        advance();
      }
      pdl_wait();
      for (int it = 0; it < pre; ++it) issue_x(it, pre_kt[it], pre_mt[it]);
      int s = pre == kStages ? 0 : pre;
      uint32_t ph = 1;  // parity of the "empty" phase that frees stage s for its second use ... (flips per ring pass)
      for (int it = pre; it < nk; ++it) {
        mbar_wait(bar_empty + 8 * s, ph ^ 1);
        mbar_arrive_expect_tx(bar_full + 8 * s, kTxBytes);
        issue_w(s);
        issue_x(s, p_kt, p_mt);
        advance();
        if (++s == kStages) {
          s = 0;
          ph ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    // =============================== MMA issuer ===============================
    // The whole warp runs the loop (warp-uniform control flow); one elected lane issues the eight
    // MMAs and the commits of a tile back to back.  (A loop owned by `lane == 0` made the compiler
    // wrap every tcgen05.mma in its own election loop: ~150 instructions and ~1250 clocks per tile,
    // more than the 1024 clocks the eight MMAs of a 256-token tile take.)
    int kt = t_begin % KT;
    uint32_t seg_idx = 0;
    int s = 0, a = 0;
    uint32_t sph = 0, aph = 0;
    const uint64_t desc0 = tc::smem_desc_k_sw128(smem_base);
    for (int it = 0; it < nk; ++it) {
      const bool seg_first = (it == 0) || (kt == 0);
      if (seg_first) mbar_wait(bar_accempty, (seg_idx & 1) ^ 1);  // epilogue has drained the accumulator
      mbar_wait(bar_full + 8 * s, sph);                            // x tile landed
      mbar_wait(bar_afull + 8 * a, aph);                           // A tile written to TMEM
      tc::fence_after_thread_sync();
      ++kt;
      const bool seg_last = (kt == KT) || (it + 1 == nk);
      if (tc::elect_one()) {
        const uint64_t bd = desc0 + uint64_t((uint32_t(s) * kStageBytes) >> 4);
        const uint32_t ta = tmem_a + uint32_t(a) * kABufCols;
#pragma unroll
        for (int j = 0; j < 8; ++j)
          tc::mma_ts(tmem_acc, ta + 8 * j, bd + uint64_t(((j >> 2) * (BN * 128) + (j & 3) * 32) >> 4), kIdesc,
                     !(seg_first && j == 0));
        tc::commit(bar_empty + 8 * s);   // stage's x may be overwritten once these MMAs retire
        tc::commit(bar_aempty + 8 * a);  // ... and so may the A buffer
        if (seg_last) tc::commit(bar_accfull);  // segment (or this CTA's part of it) accumulated
      }
      __syncwarp();
      if (seg_last) {
        ++seg_idx;
        if (kt == KT) kt = 0;
      }
      if (++s == kStages) {
        s = 0;
        sph ^= 1;
      }
      if (++a == kNumABufs) {
        a = 0;
        aph ^= 1;
      }
    }
  } else {
    // =============================== dequant + epilogue ===============================
    pdl_wait();  // y / workspace belong to the previous kernel until here (cheap: no-op if done)
    const int dw = warp - 2;           // 0 .. 4*groups-1
    const int quarter = warp & 3;      // TMEM lane quarter this warp may touch
    const int grp = dw >> 2;           // dequant group: converts tiles it % kGemmGroups == grp into A buffer grp
    const int r = 32 * quarter + lane; // output row inside the n-tile = TMEM lane
    const uint32_t lane_base = uint32_t(32 * quarter) << 16;
    const uint32_t row_off = uint32_t(r) * 64u;
    const uint32_t swz = uint32_t((r >> 1) & 3);
    const uint32_t s_off = uint32_t(tile_scale_off(r));
    const uint32_t z_off = uint32_t(tile_zero_off(r));

    int seg = t_begin / KT;
    int seg_idx = 0;
    // Events in tile order: the group's own tiles (grp, grp + G, ...) are converted, and every
    // group takes part in closing a segment after its last tile.  `close_it` = index of the tile
    // that ends the current segment (or this CTA's part of it).
    int own = grp;
    int close_it = KT - (t_begin - seg * KT) - 1;
    if (close_it > nk - 1) close_it = nk - 1;
    int s = grp % kStages;
    uint32_t sph = uint32_t(grp / kStages) & 1u, aph = 1;  // aph: a fresh "A empty" barrier passes parity 1 at once
    while (nk > 0) {
      if (own <= close_it) {
        const int it = own;
        mbar_wait(bar_full + 8 * s, sph);
        const uint32_t wbase = smem_base + s * kStageBytes + kXBytes;
        uint4 gq[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) gq[c] = lds128(wbase + row_off + 16u * (uint32_t(c) ^ swz));
        // group size >= 128: one (scale, zero) per row per tile; 64 / 32: one per 64- / 32-k sub-group, i.e. per
        // pair of granules / per granule (granule c = k 32c .. 32c + 31), in metadata section (c * sub) / 4
        uint32_t sbits4[4], z4[4];
        sbits4[0] = lds16(wbase + s_off);
        z4[0] = lds8(wbase + z_off);
        if (geo.sub > 1) {  // uniform
#pragma unroll
          for (int c = 1; c < 4; ++c) {
            const uint32_t sec = uint32_t(kTileMetaSection) * uint32_t((c * geo.sub) >> 2);
            sbits4[c] = lds16(wbase + s_off + sec);
            z4[c] = lds8(wbase + z_off + sec);
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_empty + 8 * s);  // this warp no longer needs the record

        // All 128 weights of the row are converted BEFORE waiting for the A buffer: the arithmetic
        // overlaps the MMAs of the group's previous tile, and the buffer is only held for the four
        // stores (64 result registers; 448 threads leave 144 per thread).
        uint32_t av[4][16];
        auto convert = [&](int c, uint32_t s2, uint32_t zc_e, uint32_t zc_o) {  // granule c = TMEM columns 16c .. 16c+15
          const uint32_t words[4] = {gq[c].x, gq[c].y, gq[c].z, gq[c].w};
#pragma unroll
          for (int w = 0; w < 4; ++w) {
            uint32_t raw[4];
            Unpack<T>::raw(words[w], raw);
            av[c][4 * w + 0] = Unpack<T>::mul(Unpack<T>::template diff<0>(raw[0], zc_e), s2);
            av[c][4 * w + 1] = Unpack<T>::mul(Unpack<T>::template diff<1>(raw[1], zc_o), s2);
            av[c][4 * w + 2] = Unpack<T>::mul(Unpack<T>::template diff<2>(raw[2], zc_e), s2);
            av[c][4 * w + 3] = Unpack<T>::mul(Unpack<T>::template diff<3>(raw[3], zc_o), s2);
          }
        };
        if (geo.sub == 1) {
          const uint32_t s2 = sbits4[0] | (sbits4[0] << 16);
          uint32_t zc_e, zc_o;
          Unpack<T>::zero_consts_scalar(z4[0], zc_e, zc_o);
#pragma unroll
          for (int c = 0; c < 4; ++c) convert(c, s2, zc_e, zc_o);
        } else {
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            uint32_t zc_e, zc_o;
            Unpack<T>::zero_consts_scalar(z4[c], zc_e, zc_o);
            convert(c, sbits4[c] | (sbits4[c] << 16), zc_e, zc_o);
          }
        }
        // the MMAs that read this group's A buffer (tile it - groups) must have retired
        mbar_wait(bar_aempty + 8 * grp, aph);
        aph ^= 1;
        tc::fence_after_thread_sync();
        const uint32_t a_addr = tmem_a + lane_base + uint32_t(grp * kABufCols);
#pragma unroll
        for (int c = 0; c < 4; ++c) tc::st_32x32b_x16(a_addr + 16 * c, av[c]);
        tc::wait_st();
        tc::fence_before_thread_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_afull + 8 * grp);
        (void)it;
        own += kGemmGroups;
        s += kGemmGroups;
        while (s >= kStages) {  // kStages may be smaller than the group count
          s -= kStages;
          sph ^= 1;
        }
        continue;
      }
      {
        // ---------------- close this CTA's part of segment `seg` ----------------
        const bool seg_complete = (close_it + t_begin + 1) % KT == 0;
        const int mt = seg / NT, nt = seg - mt * NT;
        const int lo = seg * KT, hi = lo + KT;
        int first, last;
        if (geo.align) {
          first = last = int(blockIdx.x);
        } else {
          first = int(unsigned(lo) * unsigned(grid) / unsigned(geo.ttot));
          while (first > 0 && gemm_begin(first, grid, geo) > lo) --first;
          while (gemm_begin(first + 1, grid, geo) <= lo) ++first;
          last = int(unsigned(hi - 1) * unsigned(grid) / unsigned(geo.ttot));
          if (last >= grid) last = grid - 1;
          while (last + 1 < grid && gemm_begin(last + 1, grid, geo) < hi) ++last;
          while (gemm_begin(last, grid, geo) >= hi) --last;
        }
        const int ncontrib = last - first + 1;
        const int rank = int(blockIdx.x) - first;

        mbar_wait(bar_accfull, seg_idx & 1);
        tc::fence_after_thread_sync();

        const int n = nt * kTileN + r;
        const float bias_v = (bias != nullptr && n < geo.N) ? Elem<T>::to_float(bias[n]) : 0.f;
        constexpr int kChunks = BN / 16;
        // partial slot: compact index first + seg + rank (unique per (segment, contributor))
        float4* part = reinterpret_cast<float4*>(ws_partial) +
                       (int64_t(first + seg + rank) * kTileN * (BN / 4)) + int64_t(r) * 4;
        bool publish = ncontrib > 1;
        // SiluAndMul epilogue (python/sglang/srt/models/llama.py:94-106): the up row of gate row r is TMEM
        // lane r + 8 = lane + 8 of the same warp; values are rounded to T first like the unfused sequence
        auto store_silu = [&](int ch, const float (&val)[16]) {
          const int NH = geo.N >> 1, col = nt * (kTileN / 2) + (r >> 4) * 8 + (r & 7);
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int m = mt * BN + 16 * ch + j;
            const float gq = Elem<T>::to_float(Elem<T>::from_float(val[j]));
            const float uq = __shfl_xor_sync(0xffffffffu, gq, 8);
            if ((lane & 8) == 0 && m < geo.M && col < NH)
              y[int64_t(m) * NH + col] = Elem<T>::from_float(gq / (1.f + __expf(-gq)) * uq);
          }
        };
#pragma unroll 1
        for (int ch = grp; ch < kChunks; ch += kGemmGroups) {
          uint32_t v[16];
          tc::ld_32x32b_x16(tmem_acc + lane_base + uint32_t(16 * ch), v);
          tc::wait_ld();
          if (publish) {
            // layout [chunk][row][16 floats]: a lane writes 64 contiguous bytes
            float4* p = part + int64_t(ch) * kTileN * 4;
#pragma unroll
            for (int q = 0; q < 4; ++q)
              __stcg(p + q, make_float4(__uint_as_float(v[4 * q]), __uint_as_float(v[4 * q + 1]),
                                        __uint_as_float(v[4 * q + 2]), __uint_as_float(v[4 * q + 3])));
          } else if (geo.silu) {
            store_silu(ch, reinterpret_cast<const float(&)[16]>(v));
          } else if (n < geo.N) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const int m = mt * BN + 16 * ch + j;
              if (m < geo.M) y[int64_t(m) * geo.N + n] = Elem<T>::from_float(__uint_as_float(v[j]) + bias_v);
            }
          }
        }
        tc::fence_before_thread_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_accempty);  // accumulator may be overwritten

        if (publish) {
          __syncwarp();
          unsigned ticket = 0;
          int* sem = ws_sem + seg * kGemmDequantWarps + dw;
          if (lane == 0)
            asm volatile("atom.add.acq_rel.gpu.global.u32 %0, [%1], 1;" : "=r"(ticket) : "l"(sem) : "memory");
          ticket = __shfl_sync(0xffffffffu, ticket, 0);
          if (int(ticket) == ncontrib - 1) {
            // last arriver for this (segment, warp): sum the partials in rank order
            const float4* base = reinterpret_cast<const float4*>(ws_partial) +
                                 (int64_t(first + seg) * kTileN * (BN / 4)) + int64_t(r) * 4;
#pragma unroll 1
            for (int ch = grp; ch < kChunks; ch += kGemmGroups) {
              float accv[16];
#pragma unroll
              for (int j = 0; j < 16; ++j) accv[j] = 0.f;
#pragma unroll 2
              for (int rr = 0; rr < ncontrib; ++rr) {
                const float4* p = base + int64_t(rr) * kTileN * (BN / 4) + int64_t(ch) * kTileN * 4;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                  const float4 f = __ldcg(p + q);
                  accv[4 * q] += f.x;
                  accv[4 * q + 1] += f.y;
                  accv[4 * q + 2] += f.z;
                  accv[4 * q + 3] += f.w;
                }
              }
              if (geo.silu) {
                store_silu(ch, accv);
              } else if (n < geo.N) {
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                  const int m = mt * BN + 16 * ch + j;
                  if (m < geo.M) y[int64_t(m) * geo.N + n] = Elem<T>::from_float(accv[j] + bias_v);
                }
              }
            }
            __syncwarp();
            if (lane == 0) *sem = 0;
          }
        }
        ++seg_idx;
        if (seg_complete) ++seg;
        if (close_it == nk - 1) break;
        close_it += KT;
        if (close_it > nk - 1) close_it = nk - 1;
      }
    }
  }

  // ---- teardown ----
  tc::fence_before_thread_sync();
  __syncthreads();
  if (warp == 1) {
    tc::fence_after_thread_sync();
    tc::tmem_dealloc(tmem_base, kTmemCols);
  }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode_fn() {
  static PFN_encodeTiled fn = nullptr;
  if (fn == nullptr) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_encodeTiled>(p);
  }
  return fn;
}

int make_x_tmap(CUtensorMap* map, const void* x, int64_t M, int64_t K, int bn, int dtype) {
  PFN_encodeTiled enc = get_encode_fn();
  AWQ_CHECK(enc != nullptr, "awq_b200_gemm: cuTensorMapEncodeTiled is not available from the driver");
  const cuuint64_t dims[2] = {cuuint64_t(K), cuuint64_t(M)};
  const cuuint64_t strides[1] = {cuuint64_t(K) * 2};
  const cuuint32_t box[2] = {64u, cuuint32_t(bn)};
  const cuuint32_t estr[2] = {1u, 1u};
  const CUresult r = enc(map, dtype == AWQ_B200_F16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16,
                         2, const_cast<void*>(x), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  AWQ_CHECK(r == CUDA_SUCCESS, "awq_b200_gemm: cuTensorMapEncodeTiled failed (%d)", int(r));
  return 0;
}

static int gemm_bn_for(int64_t M) { return M <= 16 ? 16 : (M <= 32 ? 32 : (M <= 64 ? 64 : (M <= 128 ? 128 : 256))); }

struct GemmPlan {
  GemmGeom geo;
  int bn, grid, smem;
  int64_t ws_need;
};

static int g_gemm_bn = 0;      // debug: forced token tile (0 = planner)
static int g_gemm_align = -1;  // debug: forced work split (-1 = planner, 0 = stream-K through K, 1 = whole segments)

static void gemm_geom(GemmPlan& p, int64_t M, int64_t K, int64_t N, int bn, int align) {
  p.bn = bn;
  GemmGeom& g = p.geo;
  g.M = int(M);
  g.K = int(K);
  g.N = int(N);
  g.KT = int(K / kTileK);
  g.MT = int((M + bn - 1) / bn);
  g.NT = int((N + kTileN - 1) / kTileN);
  g.S = g.NT * g.MT;
  g.ttot = g.S * g.KT;
  const int sms = sm_count();
  g.align = align;
  g.silu = 0;
  g.sub = 1;
  p.grid = g.align ? (sms < g.S ? sms : g.S) : (sms < g.ttot ? sms : g.ttot);
  p.smem = gemm_stages(bn) * gemm_stage_bytes(bn) + 1024 /*align slack*/ + 8 * (2 * gemm_stages(bn) + 2 * kNumABufs + 2) + 16;
  p.ws_need = g.align ? 0 : int64_t(kGemmSemInts) * 4 + int64_t(p.grid + g.S + 1) * kTileN * bn * 4;
}

// Token tile and work split.  Measured on B200 (scripts/gemm_matrix.py, profiles/r02_gemm_plans.txt):
// a 128-k tile costs ~0.80 / 0.57 / 0.47 us at 256 / 128 / 64 tokens (L2 -> SM traffic for the x tile
// bounds the first, the conversion pipeline the others); whole segments per CTA ("aligned") cost
// waves x KT tiles; cutting through K balances any problem but pays 13-35 us for the fp32 partials.
static GemmPlan gemm_plan(int64_t M, int64_t K, int64_t N) {
  GemmPlan p;
  const int sms = sm_count();
  const int64_t KT = K / kTileK, NT = (N + kTileN - 1) / kTileN;
  auto split_ok = [&](int64_t S) { return S < 2 * sms && S * kGemmDequantWarps <= kGemmSemInts; };
  if (g_gemm_bn > 0 || M <= 64) {
    const int bn = g_gemm_bn > 0 ? g_gemm_bn : gemm_bn_for(M);
    const int64_t S = NT * ((M + bn - 1) / bn);
    int align = split_ok(S) ? 0 : 1;
    if (g_gemm_align >= 0 && (g_gemm_align == 1 || split_ok(S))) align = g_gemm_align;
    gemm_geom(p, M, K, N, bn, align);
    return p;
  }
  static const int kBn[3] = {256, 128, 64};
  static const float kTile[3] = {0.80f, 0.57f, 0.47f}, kFix[3] = {35.f, 28.f, 13.f};
  int best_bn = 256, best_align = 1;
  float best = 1e30f;
  for (int c = 0; c < 3; ++c) {
    const int bn = kBn[c];
    if (bn >= 2 * M) continue;  // more than half of every MMA would be padding
    const int64_t S = NT * ((M + bn - 1) / bn);
    const float aligned = float((S + sms - 1) / sms) * float(KT) * kTile[c];
    if (aligned < best) {
      best = aligned;
      best_bn = bn;
      best_align = 1;
    }
    if (split_ok(S)) {
      const float split = float((S * KT + sms - 1) / sms) * kTile[c] + kFix[c];
      if (split < best) {
        best = split;
        best_bn = bn;
        best_align = 0;
      }
    }
  }
  gemm_geom(p, M, K, N, best_bn, best_align);
  return p;
}

template <typename T, int BN>
static int gemm_launch(const GemmPlan& p, const void* x, const void* tiles, const void* bias, void* y, int dtype,
                       void* workspace, cudaStream_t st) {
  auto kern = gemm_kernel<T, BN>;
  static bool attr_done[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && !attr_done[dev]) {
    AWQ_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, p.smem),
                   "awq_b200_gemm: cudaFuncSetAttribute");
    attr_done[dev] = true;
  }
  CUtensorMap tmap;
  if (int rc = make_x_tmap(&tmap, x, p.geo.M, p.geo.K, BN, dtype)) return rc;
  int* sem = reinterpret_cast<int*>(workspace);
  float* partial = workspace ? reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(workspace) + kGemmSemInts * 4) : nullptr;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(unsigned(p.grid));
  cfg.blockDim = dim3(kGemmThreads);
  cfg.dynamicSmemBytes = size_t(p.smem);
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  AWQ_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, tmap, (const uint8_t*)tiles, (const T*)bias, (T*)y, partial, sem, p.geo),
                 "awq_b200_gemm: launch");
  return 0;
}

int64_t gemm_workspace_bytes(int64_t max_M, int64_t max_N) {
  // split mode only when segments < 2 * SMs: slots <= grid + S + 1 < 3 * SMs + 1; 148-SM parts: 445
  const int bn = gemm_bn_for(max_M);
  (void)max_N;
  return int64_t(kGemmSemInts) * 4 + int64_t(3 * 160 + 1) * kTileN * bn * 4;
}

}  // namespace awq

using namespace awq;

static int gemm_any(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K, int64_t N, int dtype,
                    void* workspace, int64_t workspace_bytes, void* stream, int silu, int sub = 1) {
  if (M == 0 || N == 0) return 0;
  AWQ_CHECK(M >= 1 && M < (1 << 24), "awq_b200_gemm: M (%lld) out of range", (long long)M);
  AWQ_CHECK(K > 0 && K % kTileK == 0, "awq_b200_gemm: K (%lld) must be a positive multiple of 128", (long long)K);
  AWQ_CHECK(N > 0 && N % 8 == 0, "awq_b200_gemm: N (%lld) must be a positive multiple of 8", (long long)N);
  AWQ_CHECK(dtype == AWQ_B200_F16 || dtype == AWQ_B200_BF16, "awq_b200_gemm: bad dtype %d", dtype);
  AWQ_CHECK(x && tiles && y, "awq_b200_gemm: null pointer");
  AWQ_CHECK(((uintptr_t)x % 16) == 0 && ((uintptr_t)tiles % 16) == 0, "awq_b200_gemm: x and tiles must be 16-byte aligned");
  if (int rc = awq_b200_check_device()) return rc;
  GemmPlan p = gemm_plan(M, K, N);
  p.geo.silu = silu;
  p.geo.sub = sub;
  AWQ_CHECK(int64_t(p.geo.ttot) * (p.grid + 1) < (int64_t(1) << 31), "awq_b200_gemm: problem too large");
  if (!p.geo.align) {
    AWQ_CHECK(workspace != nullptr && workspace_bytes >= p.ws_need,
              "awq_b200_gemm: workspace too small (%lld < %lld bytes); size it with awq_b200_workspace_bytes",
              (long long)workspace_bytes, (long long)p.ws_need);
    AWQ_CHECK(((uintptr_t)workspace % 16) == 0, "awq_b200_gemm: workspace must be 16-byte aligned");
  }
  cudaStream_t st = (cudaStream_t)stream;
#define AWQ_GEMM_CASE(TT, BB) return gemm_launch<TT, BB>(p, x, tiles, bias, y, dtype, workspace, st)
  if (dtype == AWQ_B200_F16) {
    switch (p.bn) {
      case 16: AWQ_GEMM_CASE(__half, 16);
      case 32: AWQ_GEMM_CASE(__half, 32);
      case 64: AWQ_GEMM_CASE(__half, 64);
      case 128: AWQ_GEMM_CASE(__half, 128);
      default: AWQ_GEMM_CASE(__half, 256);
    }
  } else {
    switch (p.bn) {
      case 16: AWQ_GEMM_CASE(__nv_bfloat16, 16);
      case 32: AWQ_GEMM_CASE(__nv_bfloat16, 32);
      case 64: AWQ_GEMM_CASE(__nv_bfloat16, 64);
      case 128: AWQ_GEMM_CASE(__nv_bfloat16, 128);
      default: AWQ_GEMM_CASE(__nv_bfloat16, 256);
    }
  }
#undef AWQ_GEMM_CASE
}

extern "C" int awq_b200_gemm(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K,
                             int64_t N, int dtype, void* workspace, int64_t workspace_bytes, void* stream) {
  return gemm_any(x, tiles, bias, y, M, K, N, dtype, workspace, workspace_bytes, stream, 0);
}

extern "C" void awq_b200_debug_set_gemm_plan(int bn, int align) {
  g_gemm_bn = (bn == 16 || bn == 32 || bn == 64 || bn == 128 || bn == 256) ? bn : 0;
  g_gemm_align = align < 0 ? -1 : (align ? 1 : 0);
}

static int g_linear_gemv_max_m = 64;

extern "C" int awq_b200_linear_route(int64_t M, int64_t K, int64_t N) {
  // decode kernels up to M = 32; in (32, 64] the mma.sync stream-K kernel only wins on small
  // matrices (measured: o_proj 16.3 vs 18.8 us, gate_up 63.4 vs 38.6 us at M = 64)
  if (M > 32 && M <= g_linear_gemv_max_m && K * N >= (int64_t(24) << 20)) return 1;
  return M <= g_linear_gemv_max_m ? 0 : 1;
}

extern "C" void awq_b200_set_linear_gemv_max_m(int m) { g_linear_gemv_max_m = m < 0 ? 0 : (m > 64 ? 64 : m); }

extern "C" int awq_b200_linear(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K,
                               int64_t N, int dtype, void* workspace, int64_t workspace_bytes, void* stream) {
  if (awq_b200_linear_route(M, K, N) == 0)
    return awq_b200_gemv(x, tiles, bias, y, M, K, N, dtype, workspace, workspace_bytes, stream);
  return awq_b200_gemm(x, tiles, bias, y, M, K, N, dtype, workspace, workspace_bytes, stream);
}

namespace awq {
int gemv_silu_mul(const void* x, const void* tiles, void* y, int64_t M, int64_t K, int64_t N, int dtype, void* workspace,
                  int64_t workspace_bytes, cudaStream_t st);
}

// y[M, N / 2] = silu(x @ W_gate) * (x @ W_up) on tiles made by awq_b200_repack_gate_up: the gate_up_proj linear and
// the SiluAndMul that follows it in the reference MLP (python/sglang/srt/models/llama.py:94-106) as ONE launch.
extern "C" int awq_b200_linear_silu_mul(const void* x, const void* tiles, void* y, int64_t M, int64_t K, int64_t N,
                                        int dtype, void* workspace, int64_t workspace_bytes, void* stream) {
  if (M == 0 || N == 0) return 0;
  AWQ_CHECK(M >= 1 && M < (1 << 24), "awq_b200_linear_silu_mul: M (%lld) out of range", (long long)M);
  AWQ_CHECK(K > 0 && K % kTileK == 0, "awq_b200_linear_silu_mul: K (%lld) must be a positive multiple of 128", (long long)K);
  AWQ_CHECK(N > 0 && N % 16 == 0, "awq_b200_linear_silu_mul: N (%lld) must be 2 x a multiple of 8", (long long)N);
  AWQ_CHECK(dtype == AWQ_B200_F16 || dtype == AWQ_B200_BF16, "awq_b200_linear_silu_mul: bad dtype %d", dtype);
  AWQ_CHECK(x && tiles && y, "awq_b200_linear_silu_mul: null pointer");
  AWQ_CHECK(((uintptr_t)x % 16) == 0 && ((uintptr_t)tiles % 16) == 0 && ((uintptr_t)workspace % 16) == 0,
            "awq_b200_linear_silu_mul: x, tiles and workspace must be 16-byte aligned");
  if (int rc = awq_b200_check_device()) return rc;
  if (M <= 32 && awq_b200_linear_route(M, K, N) == 0) {
    const int rc = gemv_silu_mul(x, tiles, y, M, K, N, dtype, workspace, workspace_bytes, (cudaStream_t)stream);
    if (rc >= 0) return rc;
  }
  return gemm_any(x, tiles, nullptr, y, M, K, N, dtype, workspace, workspace_bytes, stream, 1);
}

namespace awq {
bool gemv_sub_fits(int64_t M, int64_t K, int64_t group, int dtype);
int gemv_sub(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K, int64_t N, int64_t group,
             int dtype, cudaStream_t st);
void gemv_sub_debug(int force_pw);
static int g_sub_decode = 1;  // debug: 0 sends every group-32 / group-64 problem to the tcgen05 kernel
}  // namespace awq
extern "C" void awq_b200_debug_set_sub_decode(int on) {
  awq::g_sub_decode = on;
  awq::gemv_sub_debug(on == 2);
}

// AWQLinearMethod.apply for any group size of the reference op (awq_triton.py:10): records from awq_b200_repack_g.
// 32 / 64: decode batches (M <= 8) take the sub-group decode kernels
// (awq_gemv_sub.cu); everything else the tcgen05 kernel (its dequant warps scale every weight, so a (scale, zero) pair
// per 32- / 64-k sub-group costs them nothing); multiples of 128: plain records, the usual routing.
extern "C" int awq_b200_linear_g(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K, int64_t N,
                                 int64_t group_size, int dtype, void* workspace, int64_t workspace_bytes, void* stream) {
  if (group_size >= kTileK && group_size % kTileK == 0)
    return awq_b200_linear(x, tiles, bias, y, M, K, N, dtype, workspace, workspace_bytes, stream);
  AWQ_CHECK(group_size == 32 || group_size == 64, "awq_b200_linear_g: group size (%lld) must be 32, 64 or a multiple of 128",
            (long long)group_size);
  AWQ_CHECK(((uintptr_t)workspace % 16) == 0, "awq_b200_linear_g: workspace must be 16-byte aligned");
  if (g_sub_decode && N > 0 && N % 8 == 0 && gemv_sub_fits(M, K, group_size, dtype)) {
    AWQ_CHECK(x && tiles && y, "awq_b200_linear_g: null pointer");
    AWQ_CHECK(((uintptr_t)x % 16) == 0 && ((uintptr_t)tiles % 16) == 0, "awq_b200_linear_g: x and tiles must be 16-byte aligned");
    AWQ_CHECK(K * N / 2 < (int64_t(1) << 40), "awq_b200_linear_g: problem too large");
    if (int rc = awq_b200_check_device()) return rc;
    return gemv_sub(x, tiles, bias, y, M, K, N, group_size, dtype, (cudaStream_t)stream);
  }
  return gemm_any(x, tiles, bias, y, M, K, N, dtype, workspace, workspace_bytes, stream, 0, int(kTileK / group_size));
}
