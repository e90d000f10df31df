// awq_repack.cu -- kernel 4: one-time load-time repack of AWQ tensors into the B200 tile layout.
//
// Called from AWQLinearMethod.process_weights_after_loading (reference hook:
// python/sglang/srt/layers/quantization/awq.py:429-432, which only re-wraps the tensors).
//
// Why a new layout: the reference layout packs 8 OUTPUT features per word (row = one k), so a
// streaming kernel needs 8 scales + 8 zeros per word and sweeps the matrix in 4-byte columns.
// The tile layout packs 8 consecutive INPUT features of ONE output feature per word:
//   * one tile (128 n x 128 k = one AWQ group per row) is ONE contiguous 8576-byte record
//     (weights | per-16-row metadata): a warp's 16 rows are 1 KB + one 48-byte block, a whole
//     tile is a single bulk copy, consecutive k-tiles of an n-tile are adjacent in memory;
//   * a thread that owns output row r needs ONE scale and ONE zero per tile;
//   * the nibble order inside a word is the AWQ interleave applied along k, so the four lop3
//     "quads" of the int4->fp16 trick come out as (k, k+1) pairs: exactly the register
//     fragments of mma.m16n8k16 (GEMV; the activations' matching fragment is one 128-bit
//     load of 8 consecutive k) and the 32-bit TMEM columns of a tcgen05 A operand (GEMM) --
//     no shuffles, no shared-memory transpose;
//   * the 16-byte granules of a row are XOR-swizzled so both access patterns (warp = 16 rows x
//     4 granules, or thread = 1 row x 4 granules) are bank-conflict free.
// Byte-for-byte spec: oracle/awq_oracle.c:awq_oracle_repack (checked bit-exactly in tests).
#include "common.cuh"
#include "host_common.h"

namespace awq {

template <typename T>
__global__ void __launch_bounds__(256)
repack_kernel(const uint32_t* __restrict__ qweight, const T* __restrict__ scales,
              const uint32_t* __restrict__ qzeros, uint8_t* __restrict__ tiles, int K, int N, int KT, int inter, int group) {
  __shared__ uint32_t src[kTileK][17];  // [k][packed col], +1 pad against bank conflicts
  const int kt = blockIdx.x, nt = blockIdx.y;
  const int NP = N / 8;
  const int tid = threadIdx.x;
  // group = 32 / 64: `sub` metadata sections per record (one per sub-group); group = a multiple of 128: one section,
  // scale row (128 kt) / group (the group's metadata is repeated in every tile it spans)
  const int sub = group < kTileK ? kTileK / group : 1;
  uint8_t* rec = tiles + (int64_t(nt) * KT + kt) * tile_bytes_sub(sub);

  // inter > 0 (merged gate_up weight, awq_b200_repack_gate_up): the tile's 128 physical rows are 8 slabs of
  // [8 gate rows | 8 up rows]: packed columns 8 nt .. + 7 of the gate half and of the up half
  const int IP = inter / 8;
  for (int i = tid; i < kTileK * 16; i += 256) {
    const int k = i >> 4, c = i & 15;
    int col = nt * 16 + c;
    bool ok = col < NP;
    if (inter > 0) {
      const int h = nt * 8 + (c & 7);
      col = (c & 8 ? IP : 0) + h;
      ok = h < IP;
    }
    src[k][c] = ok ? qweight[int64_t(kt * kTileK + k) * NP + col] : 0u;
  }
  __syncthreads();

  const int rev[8] = {0, 4, 1, 5, 2, 6, 3, 7};
  const int order[8] = {0, 2, 4, 6, 1, 3, 5, 7};
  uint32_t* wt = reinterpret_cast<uint32_t*>(rec);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int o = tid + 256 * i;
    const int r = o >> 4, pos = o & 15;
    const int c = (pos >> 2) ^ ((r >> 1) & 3), sp = pos & 3;
    const int shift = 4 * rev[r & 7];
    const int pc = inter > 0 ? (r & 8) + (r >> 4) : r >> 3;  // packed column of src holding physical row r
    const int k0 = 8 * (4 * c + sp);
    uint32_t word = 0;
#pragma unroll
    for (int p = 0; p < 8; ++p) word |= ((src[k0 + order[p]][pc] >> shift) & 0xFu) << (4 * p);
    wt[o] = word;  // rows n >= N read zero-filled src -> zero words
  }

  // logical column of physical row r of this tile (-1: past the end)
  auto logical = [&](int r) {
    if (inter == 0) {
      const int n = nt * kTileN + r;
      return n < N ? n : -1;
    }
    const int h = nt * 64 + (r >> 4) * 8 + (r & 7);
    return h < inter ? (r & 8 ? inter : 0) + h : -1;
  };
  for (int j = 0; j < sub; ++j) {
    const int srow = sub > 1 ? kt * sub + j : (kt * kTileK) / group;  // row of scales / qzeros
    if (tid < kTileN) {
      const int r = tid, n = logical(r);
      uint16_t sv = 0;
      if (n >= 0) sv = reinterpret_cast<const uint16_t*>(scales)[int64_t(srow) * N + n];
      *reinterpret_cast<uint16_t*>(rec + tile_scale_off(r) + kTileMetaSection * j) = sv;
    } else {
      const int r = tid - kTileN, n = logical(r);
      uint32_t z = 0;
      if (n >= 0) z = (qzeros[int64_t(srow) * NP + n / 8] >> (4 * rev[n & 7])) & 0xFu;
      rec[tile_zero_off(r) + kTileMetaSection * j] = uint8_t(z);
    }
  }
}

}  // namespace awq

using namespace awq;

static int repack_any(const char* who, const int32_t* qweight, const void* scales, const int32_t* qzeros, void* tiles,
                      int64_t K, int64_t N, int64_t inter, int64_t group, int dtype, void* stream) {
  if (K == 0 || N == 0) return 0;
  AWQ_CHECK(K > 0 && N > 0, "%s: negative size", who);
  AWQ_CHECK(K % kTileK == 0, "%s: K (%lld) must be a multiple of 128", who, (long long)K);
  AWQ_CHECK(group == 32 || group == 64 || (group >= kTileK && group % kTileK == 0 && K % group == 0),
            "%s: group size (%lld) must be 32, 64 or a multiple of 128 that divides K", who, (long long)group);
  AWQ_CHECK(N % 8 == 0, "%s: N (%lld) must be a multiple of 8", who, (long long)N);
  AWQ_CHECK(inter == 0 || (N == 2 * inter && inter % 8 == 0), "%s: N (%lld) must be 2 x a multiple of 8", who, (long long)N);
  AWQ_CHECK(dtype == AWQ_B200_F16 || dtype == AWQ_B200_BF16, "%s: bad dtype %d", who, dtype);
  AWQ_CHECK(qweight && scales && qzeros && tiles, "%s: null pointer", who);
  AWQ_CHECK(((uintptr_t)tiles % 16) == 0, "%s: tiles must be 16-byte aligned", who);
  if (int rc = awq_b200_check_device()) return rc;
  const int KT = int(K / kTileK), NT = int((N + kTileN - 1) / kTileN);
  AWQ_CHECK(NT <= 65535, "%s: N too large", who);
  dim3 grid(KT, NT);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == AWQ_B200_F16)
    repack_kernel<__half><<<grid, 256, 0, st>>>((const uint32_t*)qweight, (const __half*)scales,
                                                (const uint32_t*)qzeros, (uint8_t*)tiles, int(K), int(N), KT, int(inter), int(group));
  else
    repack_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const uint32_t*)qweight, (const __nv_bfloat16*)scales,
                                                       (const uint32_t*)qzeros, (uint8_t*)tiles, int(K), int(N), KT, int(inter), int(group));
  AWQ_CUDA_LAUNCH_CHECK(who);
  return 0;
}

extern "C" int awq_b200_repack(const int32_t* qweight, const void* scales, const int32_t* qzeros, void* tiles,
                               int64_t K, int64_t N, int dtype, void* stream) {
  return repack_any("awq_b200_repack", qweight, scales, qzeros, tiles, K, N, 0, kTileK, dtype, stream);
}

// any group size the reference op accepts on this path (awq_triton.py:10: -1 = K, 32, 64, 128): 32 / 64 give records
// of awq_b200_tile_bytes(group) bytes that only awq_b200_linear_g / awq_b200_dequantize_tiled_g read; multiples of 128
// give the plain 8576-byte records (usable with every entry point)
extern "C" int awq_b200_repack_g(const int32_t* qweight, const void* scales, const int32_t* qzeros, void* tiles,
                                 int64_t K, int64_t N, int64_t group_size, int dtype, void* stream) {
  return repack_any("awq_b200_repack_g", qweight, scales, qzeros, tiles, K, N, 0, group_size, dtype, stream);
}

extern "C" int64_t awq_b200_tile_bytes(int64_t group_size) {
  if (group_size == 32) return tile_bytes_sub(4);
  if (group_size == 64) return tile_bytes_sub(2);
  return group_size >= kTileK && group_size % kTileK == 0 ? kTileBytes : -1;
}

// merged gate_up weight (logical columns [0, N/2) = gate, [N/2, N) = up: the per-rank layout of
// MergedColumnParallelLinear, python/sglang/srt/layers/linear.py:461-775) -> tiles whose 16-row slabs hold
// 8 gate rows followed by the 8 up rows of the same output columns, for awq_b200_linear_silu_mul
extern "C" int awq_b200_repack_gate_up(const int32_t* qweight, const void* scales, const int32_t* qzeros, void* tiles,
                                       int64_t K, int64_t N, int dtype, void* stream) {
  AWQ_CHECK(N % 16 == 0, "awq_b200_repack_gate_up: N (%lld) must be 2 x a multiple of 8", (long long)N);
  return repack_any("awq_b200_repack_gate_up", qweight, scales, qzeros, tiles, K, N, N / 2, kTileK, dtype, stream);
}
