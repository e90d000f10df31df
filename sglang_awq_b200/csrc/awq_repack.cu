// awq_repack.cu -- kernel 4: one-time load-time repack of AWQ tensors into the B200 tile layout.
//
// Called from AWQLinearMethod.process_weights_after_loading (reference hook:
// python/sglang/srt/layers/quantization/awq.py:429-432, which only re-wraps the tensors).
//
// Why a new layout: the reference layout packs 8 OUTPUT features per word (row = one k), so a
// streaming kernel needs 8 scales + 8 zeros per word and the matrix is swept in 4-byte wide
// columns.  The tile layout packs 8 INPUT features of ONE output feature per word, ordered so
// that
//   * one tile (128 n x 128 k = one AWQ group per row) is 8192 contiguous bytes -> a GEMV /
//     GEMM pipeline stage is a single cp.async.bulk (UBLKCP) with no address arithmetic;
//   * a thread that owns output row r needs ONE scale and ONE zero per tile;
//   * nibble order inside a word makes the 4 lop3 "quads" of the int4->fp16 trick come out as
//     (k, k+1) pairs that are exactly the register fragments of mma.m16n8k16 (GEMV) and the
//     32-bit TMEM columns of a tcgen05 A operand (GEMM) -- no shuffles, no smem transpose;
//   * the 16-byte granules of a row are XOR-swizzled so both access patterns (warp = 16 rows x
//     4 granules, or thread = 1 row x 4 granules) are bank-conflict free.
// Byte-for-byte spec: oracle/awq_oracle.c:awq_oracle_repack (checked bit-exactly in tests).
#include "common.cuh"
#include "host_common.h"

namespace awq {

template <typename T>
__global__ void __launch_bounds__(256)
repack_kernel(const uint32_t* __restrict__ qweight, const T* __restrict__ scales,
              const uint32_t* __restrict__ qzeros, uint32_t* __restrict__ qweight_t,
              T* __restrict__ scales_t, uint8_t* __restrict__ qzeros_t, int K, int N, int KT) {
  __shared__ uint32_t src[kTileK][17];  // [k][packed col], +1 pad against bank conflicts
  const int kt = blockIdx.x, nt = blockIdx.y;
  const int NP = N / 8;
  const int tid = threadIdx.x;
  const int64_t tile = int64_t(nt) * KT + kt;

  for (int i = tid; i < kTileK * 16; i += 256) {
    const int k = i >> 4, c = i & 15;
    const int col = nt * 16 + c;
    src[k][c] = col < NP ? qweight[int64_t(kt * kTileK + k) * NP + col] : 0u;
  }
  __syncthreads();

  const int rev[8] = {0, 4, 1, 5, 2, 6, 3, 7};
  uint32_t* wt = qweight_t + tile * kTileWWords;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int o = tid + 256 * i;
    const int r = o >> 4, pos = o & 15;
    const int c = (pos >> 2) ^ ((r >> 1) & 3), sp = pos & 3;
    const int shift = 4 * rev[r & 7];
    const int pc = r >> 3;
    uint32_t word = 0;
#pragma unroll
    for (int p = 0; p < 8; ++p) {
      const int q = p & 3;
      const int k = 32 * sp + 16 * (q >> 1) + 8 * (q & 1) + 2 * c + (p >> 2);
      word |= ((src[k][pc] >> shift) & 0xFu) << (4 * p);
    }
    wt[o] = word;  // rows n >= N read zero-filled src -> zero words
  }

  if (tid < kTileN) {
    const int r = tid, n = nt * kTileN + r;
    const int pr = (r & ~15) | ((r & 7) << 1) | ((r >> 3) & 1);
    T s;
    if (n < N) {
      s = scales[int64_t(kt) * N + n];
    } else {
      const uint16_t zero = 0;
      s = *reinterpret_cast<const T*>(&zero);
    }
    scales_t[tile * kTileN + pr] = s;
  } else if (tid < kTileN + kTileZBytes) {
    const int b = tid - kTileN;
    const int rlo = 16 * (b >> 3) + (b & 7), rhi = rlo + 8;
    uint32_t byte = 0;
    const int nlo = nt * kTileN + rlo, nhi = nt * kTileN + rhi;
    if (nlo < N) byte |= (qzeros[int64_t(kt) * NP + nlo / 8] >> (4 * rev[nlo & 7])) & 0xFu;
    if (nhi < N) byte |= ((qzeros[int64_t(kt) * NP + nhi / 8] >> (4 * rev[nhi & 7])) & 0xFu) << 4;
    qzeros_t[tile * kTileZBytes + b] = uint8_t(byte);
  }
}

}  // namespace awq

using namespace awq;

extern "C" int awq_b200_repack(const int32_t* qweight, const void* scales, const int32_t* qzeros,
                               int32_t* qweight_t, void* scales_t, uint8_t* qzeros_t, int64_t K,
                               int64_t N, int dtype, void* stream) {
  if (K == 0 || N == 0) return 0;
  AWQ_CHECK(K > 0 && N > 0, "awq_b200_repack: negative size");
  AWQ_CHECK(K % kTileK == 0, "awq_b200_repack: K (%lld) must be a multiple of 128 (group size 128)", (long long)K);
  AWQ_CHECK(N % 8 == 0, "awq_b200_repack: N (%lld) must be a multiple of 8", (long long)N);
  AWQ_CHECK(dtype == AWQ_B200_F16 || dtype == AWQ_B200_BF16, "awq_b200_repack: bad dtype %d", dtype);
  AWQ_CHECK(qweight && scales && qzeros && qweight_t && scales_t && qzeros_t, "awq_b200_repack: null pointer");
  if (int rc = awq_b200_check_device()) return rc;
  const int KT = int(K / kTileK), NT = int((N + kTileN - 1) / kTileN);
  AWQ_CHECK(NT <= 65535, "awq_b200_repack: N too large");
  dim3 grid(KT, NT);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == AWQ_B200_F16)
    repack_kernel<__half><<<grid, 256, 0, st>>>((const uint32_t*)qweight, (const __half*)scales,
                                                (const uint32_t*)qzeros, (uint32_t*)qweight_t, (__half*)scales_t,
                                                qzeros_t, int(K), int(N), KT);
  else
    repack_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const uint32_t*)qweight, (const __nv_bfloat16*)scales,
                                                       (const uint32_t*)qzeros, (uint32_t*)qweight_t,
                                                       (__nv_bfloat16*)scales_t, qzeros_t, int(K), int(N), KT);
  AWQ_CUDA_LAUNCH_CHECK("awq_b200_repack");
  return 0;
}
