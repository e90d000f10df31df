// Microbenchmark: how fast can cp.async.bulk (UBLKCP) rings stream HBM on B200?
// Each CTA streams a contiguous slab through an S-stage smem ring; "consumers" only wait and release.
// Variants: copy size per stage, #copies per stage, stages, CTAs, L2 hint, plus an LDG.128 baseline.
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include <cuda_runtime.h>
#include "../../sglang_awq_b200/csrc/common.cuh"
using namespace awq;

// mode 0: one copy of `stage_bytes`; mode 1: split into `pieces` copies issued by lane 0;
// mode 2: pieces issued by different lanes
__global__ void __launch_bounds__(288) ring_kernel(const uint8_t* __restrict__ src, size_t per_cta, int stage_bytes, int stages,
                                                   int pieces, int mode, int hint, unsigned long long* sink) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t base = smem_u32(smem);
  const uint32_t bar_full = base + stages * stage_bytes, bar_empty = bar_full + 8 * stages;
  if (tid == 0) {
    for (int s = 0; s < stages; ++s) { mbar_init(bar_full + 8 * s, 1); mbar_init(bar_empty + 8 * s, 8); }
    mbar_fence_init();
  }
  __syncthreads();
  const int n = int(per_cta / stage_bytes);
  const uint8_t* p = src + size_t(blockIdx.x) * per_cta;
  if (warp == 8) {
    const uint64_t pol = policy_evict_first();
    const int piece = stage_bytes / pieces;
    for (int it = 0; it < n; ++it) {
      const int s = it % stages;
      if (lane == 0) {
        if (it >= stages) mbar_wait(bar_empty + 8 * s, ((it / stages) & 1) ^ 1);
        mbar_arrive_expect_tx(bar_full + 8 * s, stage_bytes);
      }
      __syncwarp();
      const uint32_t dst = base + s * stage_bytes;
      const uint8_t* sp = p + size_t(it) * stage_bytes;
      if (mode == 2) {
        for (int i = lane; i < pieces; i += 32) {
          if (hint) bulk_g2s_hint(dst + i * piece, sp + i * piece, piece, bar_full + 8 * s, pol);
          else bulk_g2s(dst + i * piece, sp + i * piece, piece, bar_full + 8 * s);
        }
      } else if (lane == 0) {
        for (int i = 0; i < pieces; ++i) {
          if (hint) bulk_g2s_hint(dst + i * piece, sp + i * piece, piece, bar_full + 8 * s, pol);
          else bulk_g2s(dst + i * piece, sp + i * piece, piece, bar_full + 8 * s);
        }
      }
    }
    return;
  }
  unsigned long long acc = 0;
  for (int it = 0; it < n; ++it) {
    const int s = it % stages;
    mbar_wait(bar_full + 8 * s, (it / stages) & 1);
    acc += lds32(base + s * stage_bytes + 4 * tid);
    __syncwarp();
    if (lane == 0) mbar_arrive(bar_empty + 8 * s);
  }
  if (acc == 0x123456789ull) sink[0] = acc;
}

// baseline: plain 128-bit loads, UNROLL in flight per thread
template <int UNROLL>
__global__ void __launch_bounds__(256) ldg_kernel(const uint4* __restrict__ src, size_t n_vec, unsigned long long* sink) {
  size_t i = size_t(blockIdx.x) * blockDim.x * UNROLL + threadIdx.x;
  const size_t stride = size_t(gridDim.x) * blockDim.x * UNROLL;
  uint32_t acc = 0;
  for (; i + (UNROLL - 1) * blockDim.x < n_vec; i += stride) {
    uint4 v[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u)
      asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "l"(src + i + u * blockDim.x));
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) acc ^= v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
  }
  if (acc == 0x12345678u) sink[0] = acc;
}

int main() {
  const size_t total = size_t(1) << 30;  // 1 GiB source (>> L2)
  uint8_t* src; unsigned long long* sink;
  cudaMalloc(&src, total); cudaMalloc(&sink, 8);
  cudaMemset(src, 1, total);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  auto run_ring = [&](int ctas, int stage_bytes, int stages, int pieces, int mode, int hint) {
    size_t per_cta = (total / ctas) / stage_bytes * stage_bytes;
    if (per_cta > (size_t(4) << 20)) per_cta = (size_t(4) << 20) / stage_bytes * stage_bytes;  // 4 MiB per CTA max
    int smem = stages * stage_bytes + 16 * stages;
    cudaFuncSetAttribute(ring_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    float best = 1e9;
    for (int rep = 0; rep < 3; ++rep) {
      cudaEventRecord(e0);
      ring_kernel<<<ctas, 288, smem>>>(src, per_cta, stage_bytes, stages, pieces, mode, hint, sink);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    cudaError_t err = cudaGetLastError();
    printf("ring ctas=%4d stage=%6d stages=%2d pieces=%2d mode=%d hint=%d  MB=%7.1f  us=%8.1f  GB/s=%7.1f %s\n", ctas, stage_bytes, stages,
           pieces, mode, hint, per_cta * ctas / 1e6, best * 1e3, per_cta * ctas / (best * 1e-3) / 1e9, err == cudaSuccess ? "" : cudaGetErrorString(err));
  };
  // sweep
  for (int ctas : {148, 296, 444, 592}) run_ring(ctas, 8192, 8, 1, 0, 1);
  for (int ctas : {148, 296}) run_ring(ctas, 8192, 8, 1, 0, 0);
  for (int ctas : {148, 296}) run_ring(ctas, 16384, 6, 1, 0, 1);
  for (int ctas : {148, 296}) run_ring(ctas, 32768, 3, 1, 0, 1);
  for (int ctas : {148}) run_ring(ctas, 32768, 6, 1, 0, 1);
  for (int ctas : {148}) run_ring(ctas, 65536, 3, 1, 0, 1);
  for (int ctas : {296}) run_ring(ctas, 8192, 12, 1, 0, 1);
  for (int ctas : {296, 592}) run_ring(ctas, 4096, 8, 1, 0, 1);
  for (int ctas : {296}) run_ring(ctas, 8192, 8, 4, 0, 1);   // 4 pieces by lane 0
  for (int ctas : {296}) run_ring(ctas, 8192, 8, 4, 2, 1);   // 4 pieces by 4 lanes
  for (int ctas : {296}) run_ring(ctas, 8192, 8, 32, 2, 1);  // 32 x 256B pieces by 32 lanes
  for (int ctas : {296}) run_ring(ctas, 8192, 8, 32, 0, 1);  // 32 x 256B pieces by lane 0
  // baseline LDG
  for (int blocks : {148 * 4, 148 * 8, 148 * 16}) {
    float best = 1e9;
    for (int rep = 0; rep < 3; ++rep) {
      cudaEventRecord(e0);
      ldg_kernel<8><<<blocks, 256>>>((const uint4*)src, total / 16, sink);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    printf("ldg blocks=%5d unroll=8 GB/s=%7.1f\n", blocks, total / (best * 1e-3) / 1e9);
  }
  return 0;
}
