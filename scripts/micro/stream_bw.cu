// stream_bw.cu -- development microbenchmark: what HBM bandwidth do the decode kernel's LOAD
// patterns reach on B200 when nothing but the loads (and an XOR) is executed?
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o stream_bw stream_bw.cu && ./stream_bw
//
// Patterns (all LDG.128, L1 no-allocate, kDepth x 1 KB per warp in flight):
//   slab   : CTA = one 16-row slab over the whole K (8 warps split K); a warp reads 1 KB pieces
//            at a stride of one 8576-byte tile record        (awq_gemv_direct.cu)
//   contig : every warp streams its own contiguous byte range in 1 KB pieces
//   tile   : CTA streams a contiguous range of 8 KB tiles, warp w takes KB w of each tile
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x)                                                                      \
  do {                                                                             \
    cudaError_t e = (x);                                                           \
    if (e != cudaSuccess) {                                                        \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); \
      exit(1);                                                                     \
    }                                                                              \
  } while (0)

__device__ __forceinline__ uint4 ldg128_stream(const void* p) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p));
  return v;
}

constexpr int kRec = 8576;

// generic: warp reads `n` pieces of 1 KB: piece j at p0 + j * stride
template <int kDepth>
__device__ __forceinline__ uint32_t read_pieces(const uint8_t* p, int n, int64_t stride) {
  uint4 a[kDepth], b[kDepth];
#pragma unroll
  for (int s = 0; s < kDepth; ++s)
    if (s < n) {
      a[s] = ldg128_stream(p);
      b[s] = ldg128_stream(p + 512);
      p += stride;
    }
  uint32_t acc = 0;
  for (int i = 0; i < n; i += kDepth) {
#pragma unroll
    for (int s = 0; s < kDepth; ++s) {
      if (i + s < n) {
        acc ^= a[s].x ^ a[s].y ^ a[s].z ^ a[s].w ^ b[s].x ^ b[s].y ^ b[s].z ^ b[s].w;
        if (i + s + kDepth < n) {
          a[s] = ldg128_stream(p);
          b[s] = ldg128_stream(p + 512);
          p += stride;
        }
      }
    }
  }
  return acc;
}

template <int kDepth, int kMinB>
__global__ void __launch_bounds__(256, kMinB) k_slab(const uint8_t* base, int KT, uint32_t* out) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int slab = blockIdx.x, nt = slab >> 3, sl = slab & 7;
  const int kb = warp * KT / 8, ke = (warp + 1) * KT / 8;
  const uint8_t* p = base + (int64_t(nt) * KT + kb) * kRec + sl * 1024 + 16 * lane;
  uint32_t acc = read_pieces<kDepth>(p, ke - kb, kRec);
  if (acc == 0x12345678u) out[0] = acc;
}

template <int kDepth, int kMinB>
__global__ void __launch_bounds__(256, kMinB) k_contig(const uint8_t* base, int64_t pieces_total, uint32_t* out) {
  const int lane = threadIdx.x & 31;
  const int64_t gw = int64_t(blockIdx.x) * 8 + (threadIdx.x >> 5), nw = int64_t(gridDim.x) * 8;
  const int64_t b = gw * pieces_total / nw, e = (gw + 1) * pieces_total / nw;
  uint32_t acc = read_pieces<kDepth>(base + b * 1024 + 16 * lane, int(e - b), 1024);
  if (acc == 0x12345678u) out[0] = acc;
}

template <int kDepth, int kMinB>
__global__ void __launch_bounds__(256, kMinB) k_tile(const uint8_t* base, int tiles_total, uint32_t* out) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = int(int64_t(blockIdx.x) * tiles_total / gridDim.x), e = int(int64_t(blockIdx.x + 1) * tiles_total / gridDim.x);
  uint32_t acc = read_pieces<kDepth>(base + int64_t(b) * kRec + warp * 1024 + 16 * lane, e - b, kRec);
  if (acc == 0x12345678u) out[0] = acc;
}

// like k_tile, but the CTA has 8*kWide warps that take kWide CONSECUTIVE tiles per step: same bytes in
// flight per SM at equal warps/SM, longer contiguous spans
template <int kDepth, int kWide, int kMinB>
__global__ void __launch_bounds__(256 * kWide, kMinB) k_tile_wide(const uint8_t* base, int tiles_total, uint32_t* out) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = int(int64_t(blockIdx.x) * tiles_total / gridDim.x), e = int(int64_t(blockIdx.x + 1) * tiles_total / gridDim.x);
  const int sub = warp >> 3;  // which of the kWide tiles of a step
  const int n = (e - b - sub + kWide - 1) / kWide;
  uint32_t acc = read_pieces<kDepth>(base + int64_t(b + sub) * kRec + (warp & 7) * 1024 + 16 * lane, n > 0 ? n : 0, int64_t(kRec) * kWide);
  if (acc == 0x12345678u) out[0] = acc;
}

struct Shape {
  const char* name;
  int K, N;
};

template <typename F>
static float time_launches(F launch, int windows, int reps) {
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  for (int w = 0; w < windows; ++w) launch(w);
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  for (int r = 0; r < reps; ++r)
    for (int w = 0; w < windows; ++w) launch(w);
  CK(cudaEventRecord(e1));
  CK(cudaDeviceSynchronize());
  float ms = 0;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  return ms * 1e3f / (reps * windows);
}

int main() {
  const size_t total = size_t(1) << 30;
  uint8_t* buf;
  uint32_t* out;
  CK(cudaMalloc(&buf, total + (1 << 20)));
  CK(cudaMalloc(&out, 64));
  CK(cudaMemset(buf, 1, total));
  int sms = 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  const Shape shapes[] = {{"o 4096x4096", 4096, 4096},
                          {"qkv 4096x6144", 4096, 6144},
                          {"down 14336x4096", 14336, 4096},
                          {"gate_up 4096x28672", 4096, 28672},
                          {"70b gate_up 8192x57344", 8192, 57344}};
  for (const Shape& s : shapes) {
    const int KT = s.K / 128, NT = s.N / 128, tiles = KT * NT;
    const size_t bytes = size_t(tiles) * kRec;
    const int windows = int(total / bytes) < 1 ? 1 : (int(total / bytes) > 16 ? 16 : int(total / bytes));
    const int reps = bytes < (64u << 20) ? 20 : 5;
    printf("== %s: %.1f MB, %d windows\n", s.name, bytes / 1e6, windows);
#define RUN(label, KERNEL, GRID, ...)                                                          \
  {                                                                                            \
    float us = time_launches([&](int w) { KERNEL<<<GRID, 256>>>(buf + size_t(w) * bytes, __VA_ARGS__, out); }, windows, reps); \
    CK(cudaGetLastError());                                                                    \
    printf("  %-34s grid %5d  %8.2f us  %7.1f GB/s\n", label, int(GRID), us, bytes / us / 1e3); \
  }
    RUN("slab   D=4 occ2", (k_slab<4, 2>), s.N / 16, KT);
    RUN("slab   D=4 occ4", (k_slab<4, 4>), s.N / 16, KT);
    RUN("slab   D=8 occ2", (k_slab<8, 2>), s.N / 16, KT);
    RUN("slab   D=8 occ4", (k_slab<8, 4>), s.N / 16, KT);
    RUN("contig D=4 occ2 grid=2*sm", (k_contig<4, 2>), 2 * sms, int64_t(bytes / 1024));
    RUN("contig D=8 occ2 grid=2*sm", (k_contig<8, 2>), 2 * sms, int64_t(bytes / 1024));
    RUN("contig D=8 occ4 grid=4*sm", (k_contig<8, 4>), 4 * sms, int64_t(bytes / 1024));
    RUN("tile   D=4 occ2 grid=2*sm", (k_tile<4, 2>), 2 * sms, tiles);
    RUN("tile   D=8 occ2 grid=2*sm", (k_tile<8, 2>), 2 * sms, tiles);
    RUN("tile   D=8 occ4 grid=4*sm", (k_tile<8, 4>), 4 * sms, tiles);
    RUN("tile   D=4 occ4 grid=4*sm", (k_tile<4, 4>), 4 * sms, tiles);
#define RUNW(label, KERNEL, GRID, THREADS, ...)                                                \
  {                                                                                            \
    float us = time_launches([&](int w) { KERNEL<<<GRID, THREADS>>>(buf + size_t(w) * bytes, __VA_ARGS__, out); }, windows, reps); \
    CK(cudaGetLastError());                                                                    \
    printf("  %-34s grid %5d  %8.2f us  %7.1f GB/s\n", label, int(GRID), us, bytes / us / 1e3); \
  }
    RUNW("wide2  D=4 512thr x2/sm", (k_tile_wide<4, 2, 2>), 2 * sms, 512, tiles);
    RUNW("wide4  D=4 1024thr x1/sm", (k_tile_wide<4, 4, 1>), 1 * sms, 1024, tiles);
    RUNW("wide4  D=6 1024thr x1/sm", (k_tile_wide<6, 4, 1>), 1 * sms, 1024, tiles);
    RUNW("wide2  D=8 512thr x2/sm", (k_tile_wide<8, 2, 2>), 2 * sms, 512, tiles);
    RUNW("wide2  D=6 512thr x2/sm", (k_tile_wide<6, 2, 2>), 2 * sms, 512, tiles);
    RUN("tile   D=6 occ4 grid=4*sm", (k_tile<6, 4>), 4 * sms, tiles);
    RUN("tile   D=4 occ6 grid=6*sm", (k_tile<4, 6>), 6 * sms, tiles);
    RUN("tile   D=4 occ8 grid=8*sm", (k_tile<4, 8>), 8 * sms, tiles);
  }
  return 0;
}
