// awq_dequant.cu -- kernel 1 (standalone dequantise, reference layout) and its tile-layout twin.
//
// Replaces sgl-kernel/csrc/gemm/awq_kernel.cu:126-221 of the reference (dequantize_weights +
// awq_dequantize).  Same result bit-for-bit: out[k,n] = RN_T((w - z) * s) with (w - z) formed
// exactly in T and a single mul.rn (fp16) / mul.rn.bf16x2 (bf16).
//
// B200 mapping (HBM-bound: 0.5 B read + 2 B written per weight):
//   * a warp owns 32 consecutive packed words of a row: one coalesced 128 B read, and each
//     lane's 8 outputs go out as ONE 128-bit streaming store -> the warp writes 512 contiguous
//     bytes, a block up to 4 KB of one output row (wide blocks: DRAM page locality of the writes);
//   * a thread walks R = 8 rows of the same column block, issuing the 8 word loads up front
//     (8 independent requests in flight per thread), and keeps the group's zero constants and
//     128-bit scale vector in registers across those rows (metadata is read once per group,
//     not once per row as in the reference);
//   * zero-point subtraction is folded into the int4->fp16 conversion (13 ALU ops per 8
//     weights instead of 22), so the kernel stays far below the issue limit at HBM rate.
#include "common.cuh"
#include "host_common.h"

namespace awq {

constexpr int kDqRows = 8;      // rows per thread
constexpr int kDqThreads = 256; // block = bx packed columns x (256 / bx) row groups; bx = 128 where the row allows:
                                // a block then writes 2 KB contiguous per output row
static int g_dq_variant = 0;    // debug: bits 0-1 = block width (0 = 128, 1 = 32, 2 = 64, 3 = 256), bit 2 = default-cached stores

__device__ __forceinline__ uint32_t ldg_stream_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ uint4 ldg_nc_u128(const void* p) {
  uint4 v;
  asm volatile("ld.global.nc.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}

// the output is written once and never read by this kernel: streaming (evict-first) stores
__device__ __forceinline__ void stg_stream_u128(void* p, const uint4& v) {
  asm volatile("st.global.cs.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

template <typename T, bool kStreamStores>
__global__ void __launch_bounds__(kDqThreads)
dequant_kernel(const uint32_t* __restrict__ qweight, const T* __restrict__ scales,
               const uint32_t* __restrict__ qzeros, T* __restrict__ out, int K, int NP, int G) {
  const int col = blockIdx.x * blockDim.x + threadIdx.x;
  const int row0 = (blockIdx.y * blockDim.y + threadIdx.y) * kDqRows;
  if (col >= NP || row0 >= K) return;
  const int64_t N = int64_t(NP) * 8;

  uint32_t w[kDqRows];
#pragma unroll
  for (int i = 0; i < kDqRows; ++i) {
    const int r = row0 + i;
    w[i] = r < K ? ldg_stream_u32(qweight + int64_t(r) * NP + col) : 0u;
  }

  int g = row0 / G;
  int left = G - (row0 - g * G);  // rows left in the current group
  uint32_t zc[4];
  uint4 s;
  bool need = true;
#pragma unroll
  for (int i = 0; i < kDqRows; ++i) {
    const int r = row0 + i;
    if (r >= K) break;
    if (need) {
      uint32_t zr[4];
      Unpack<T>::raw(__ldg(qzeros + int64_t(g) * NP + col), zr);
      Unpack<T>::zero_consts(zr, zc);
      s = ldg_nc_u128(scales + int64_t(g) * N + 8 * col);
      need = false;
    }
    uint32_t wr[4];
    Unpack<T>::raw(w[i], wr);
    uint4 o;
    o.x = Unpack<T>::mul(Unpack<T>::template diff<0>(wr[0], zc[0]), s.x);
    o.y = Unpack<T>::mul(Unpack<T>::template diff<1>(wr[1], zc[1]), s.y);
    o.z = Unpack<T>::mul(Unpack<T>::template diff<2>(wr[2], zc[2]), s.z);
    o.w = Unpack<T>::mul(Unpack<T>::template diff<3>(wr[3], zc[3]), s.w);
    if constexpr (kStreamStores) {
      stg_stream_u128(out + int64_t(r) * N + 8 * col, o);
    } else {
      *reinterpret_cast<uint4*>(out + int64_t(r) * N + 8 * col) = o;
    }
    if (--left == 0) {
      ++g;
      left = G;
      need = true;
    }
  }
}

// Tile layout -> dense [K, N].  One block per tile, one thread per output feature (row of the
// tile).  Not a hot path: used by consumers that need the dense matrix once at load time.
template <typename T>
__global__ void __launch_bounds__(128)
dequant_tiled_kernel(const uint8_t* __restrict__ tiles, T* __restrict__ out, int K, int N, int KT, int sub) {
  const int kt = blockIdx.x, nt = blockIdx.y;
  const int r = threadIdx.x;
  const int n = nt * kTileN + r;
  if (n >= N) return;
  const uint8_t* rec = tiles + (int64_t(nt) * KT + kt) * tile_bytes_sub(sub);
  const uint32_t* wt = reinterpret_cast<const uint32_t*>(rec) + r * 16;
  uint16_t* o16 = reinterpret_cast<uint16_t*>(out);
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    // granule c = k 32c .. 32c + 31 of the tile = sub-group (c * sub) / 4
    const int sec = kTileMetaSection * ((c * sub) >> 2);
    const uint32_t sbits = *reinterpret_cast<const uint16_t*>(rec + tile_scale_off(r) + sec);
    const uint32_t s2 = sbits | (sbits << 16);
    const uint32_t z = rec[tile_zero_off(r) + sec];
    uint32_t zc_even, zc_odd;
    Unpack<T>::zero_consts_scalar(z, zc_even, zc_odd);
    const uint4 gran = *reinterpret_cast<const uint4*>(wt + 4 * (c ^ ((r >> 1) & 3)));
    const uint32_t words[4] = {gran.x, gran.y, gran.z, gran.w};
#pragma unroll
    for (int sp = 0; sp < 4; ++sp) {
      uint32_t wr[4];
      Unpack<T>::raw(words[sp], wr);
      uint32_t v[4];
      v[0] = Unpack<T>::mul(Unpack<T>::template diff<0>(wr[0], zc_even), s2);
      v[1] = Unpack<T>::mul(Unpack<T>::template diff<1>(wr[1], zc_odd), s2);
      v[2] = Unpack<T>::mul(Unpack<T>::template diff<2>(wr[2], zc_even), s2);
      v[3] = Unpack<T>::mul(Unpack<T>::template diff<3>(wr[3], zc_odd), s2);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        // quad q = nibbles (q, q+4) = consecutive k offsets (2q, 2q+1) of word 4c+sp
        const int k0 = kt * kTileK + 8 * (4 * c + sp) + 2 * q;
        o16[int64_t(k0) * N + n] = uint16_t(v[q] & 0xFFFFu);
        o16[int64_t(k0 + 1) * N + n] = uint16_t(v[q] >> 16);
      }
    }
  }
}

}  // namespace awq

using namespace awq;

extern "C" int awq_b200_dequantize(const int32_t* qweight, const void* scales, const int32_t* qzeros,
                                   void* out, int64_t K, int64_t N, int64_t scale_rows, int dtype,
                                   void* stream) {
  AWQ_CHECK(K >= 0 && N >= 0, "awq_b200_dequantize: negative size");
  if (K == 0 || N == 0) return 0;
  AWQ_CHECK(N % 8 == 0, "awq_b200_dequantize: N (%lld) must be a multiple of 8", (long long)N);
  AWQ_CHECK(scale_rows > 0 && K % scale_rows == 0,
            "awq_b200_dequantize: K (%lld) must be a multiple of scales rows (%lld)", (long long)K,
            (long long)scale_rows);
  AWQ_CHECK(dtype == AWQ_B200_F16 || dtype == AWQ_B200_BF16, "awq_b200_dequantize: bad dtype %d", dtype);
  AWQ_CHECK(qweight && scales && qzeros && out, "awq_b200_dequantize: null pointer");
  AWQ_CHECK(((uintptr_t)scales % 16) == 0 && ((uintptr_t)out % 16) == 0 && ((uintptr_t)qweight % 4) == 0,
            "awq_b200_dequantize: scales/out must be 16-byte aligned");
  AWQ_CHECK(K < (1ll << 31) && N < (1ll << 31), "awq_b200_dequantize: dimension too large");
  if (int rc = awq_b200_check_device()) return rc;
  const int NP = int(N / 8), G = int(K / scale_rows);
  int bx = 128;  // measured best or equal on every model shape (profiles/r02_dequant_variants.txt)
  const int cap = g_dq_variant & 3;
  if (cap) bx = cap == 3 ? 256 : 16 << cap;  // 32, 64, 256
  while (bx > 32 && bx / 2 >= NP) bx >>= 1;
  const int by = kDqThreads / bx;
  dim3 block(bx, by);
  dim3 grid((NP + bx - 1) / bx, (unsigned)((K + kDqRows * by - 1) / (kDqRows * by)));
  AWQ_CHECK(grid.y <= 65535u, "awq_b200_dequantize: K too large");
  cudaStream_t st = (cudaStream_t)stream;
  const bool stream_stores = (g_dq_variant & 4) == 0;
#define AWQ_DQ_LAUNCH(TT, SS)                                                                                      \
  dequant_kernel<TT, SS><<<grid, block, 0, st>>>((const uint32_t*)qweight, (const TT*)scales, (const uint32_t*)qzeros, \
                                                 (TT*)out, int(K), NP, G)
  if (dtype == AWQ_B200_F16) {
    if (stream_stores) AWQ_DQ_LAUNCH(__half, true); else AWQ_DQ_LAUNCH(__half, false);
  } else {
    if (stream_stores) AWQ_DQ_LAUNCH(__nv_bfloat16, true); else AWQ_DQ_LAUNCH(__nv_bfloat16, false);
  }
#undef AWQ_DQ_LAUNCH
  AWQ_CUDA_LAUNCH_CHECK("awq_b200_dequantize");
  return 0;
}

extern "C" void awq_b200_debug_set_dequant_variant(int v) { g_dq_variant = v; }

static int dequantize_tiled_any(const char* who, const void* tiles, void* out, int64_t K, int64_t N, int64_t group, int dtype,
                                void* stream) {
  if (K == 0 || N == 0) return 0;
  AWQ_CHECK(K > 0 && N > 0 && K % kTileK == 0 && N % 8 == 0, "%s: need K %% 128 == 0, N %% 8 == 0", who);
  AWQ_CHECK(group == 32 || group == 64 || (group >= kTileK && group % kTileK == 0), "%s: bad group size %lld", who, (long long)group);
  AWQ_CHECK(dtype == AWQ_B200_F16 || dtype == AWQ_B200_BF16, "%s: bad dtype %d", who, dtype);
  AWQ_CHECK(tiles && out, "%s: null pointer", who);
  AWQ_CHECK(((uintptr_t)tiles % 16) == 0, "%s: tiles must be 16-byte aligned", who);
  if (int rc = awq_b200_check_device()) return rc;
  const int KT = int(K / kTileK), NT = int((N + kTileN - 1) / kTileN);
  AWQ_CHECK(NT <= 65535, "%s: N too large", who);
  const int sub = group < kTileK ? int(kTileK / group) : 1;
  dim3 grid(KT, NT);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == AWQ_B200_F16)
    dequant_tiled_kernel<__half><<<grid, 128, 0, st>>>((const uint8_t*)tiles, (__half*)out, int(K), int(N), KT, sub);
  else
    dequant_tiled_kernel<__nv_bfloat16><<<grid, 128, 0, st>>>((const uint8_t*)tiles, (__nv_bfloat16*)out, int(K), int(N),
                                                              KT, sub);
  AWQ_CUDA_LAUNCH_CHECK(who);
  return 0;
}

extern "C" int awq_b200_dequantize_tiled(const void* tiles, void* out, int64_t K, int64_t N, int dtype,
                                         void* stream) {
  return dequantize_tiled_any("awq_b200_dequantize_tiled", tiles, out, K, N, kTileK, dtype, stream);
}

extern "C" int awq_b200_dequantize_tiled_g(const void* tiles, void* out, int64_t K, int64_t N, int64_t group_size, int dtype,
                                           void* stream) {
  return dequantize_tiled_any("awq_b200_dequantize_tiled_g", tiles, out, K, N, group_size, dtype, stream);
}
