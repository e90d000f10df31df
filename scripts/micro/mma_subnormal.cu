// mma_subnormal.cu -- does mma.sync.m16n8k16 (f16 in, f32 accumulate) on sm_100a treat fp16
// SUBNORMAL A operands exactly?  (A = n * 2^-24 for n in 0..255, B = arbitrary fp16.)
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
__global__ void k(const uint32_t* a, const uint32_t* b, float* d) {
  const int lane = threadIdx.x;
  float acc[4] = {0, 0, 0, 0};
  uint32_t A[4] = {a[lane * 4 + 0], a[lane * 4 + 1], a[lane * 4 + 2], a[lane * 4 + 3]};
  uint32_t B[2] = {b[lane * 2 + 0], b[lane * 2 + 1]};
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(acc[0]), "+f"(acc[1]), "+f"(acc[2]), "+f"(acc[3])
               : "r"(A[0]), "r"(A[1]), "r"(A[2]), "r"(A[3]), "r"(B[0]), "r"(B[1]));
  for (int i = 0; i < 4; ++i) d[lane * 4 + i] = acc[i];
}
int main() {
  // A[row][k] (16x16), B[k][n] (16x8)
  uint16_t Ah[16][16];
  __half Bh[16][8];
  srand(1);
  for (int r = 0; r < 16; ++r)
    for (int c = 0; c < 16; ++c) Ah[r][c] = (c & 1) ? uint16_t((rand() % 16) << 4) : uint16_t(rand() % 16);  // subnormal bit patterns
  for (int r = 0; r < 16; ++r)
    for (int c = 0; c < 8; ++c) Bh[r][c] = __float2half((rand() % 2001 - 1000) / 250.0f);
  uint32_t ha[128], hb[64];
  for (int lane = 0; lane < 32; ++lane) {
    int g = lane >> 2, c = lane & 3;
    auto pa = [&](int r, int k) { return uint32_t(Ah[r][k]) | (uint32_t(Ah[r][k + 1]) << 16); };
    ha[lane * 4 + 0] = pa(g, 2 * c);
    ha[lane * 4 + 1] = pa(g + 8, 2 * c);
    ha[lane * 4 + 2] = pa(g, 2 * c + 8);
    ha[lane * 4 + 3] = pa(g + 8, 2 * c + 8);
    auto pb = [&](int k, int n) { return uint32_t(*(uint16_t*)&Bh[k][n]) | (uint32_t(*(uint16_t*)&Bh[k + 1][n]) << 16); };
    hb[lane * 2 + 0] = pb(2 * c, g);
    hb[lane * 2 + 1] = pb(2 * c + 8, g);
  }
  uint32_t *da, *db;
  float* dd;
  cudaMalloc(&da, sizeof(ha));
  cudaMalloc(&db, sizeof(hb));
  cudaMalloc(&dd, 128 * 4);
  cudaMemcpy(da, ha, sizeof(ha), cudaMemcpyHostToDevice);
  cudaMemcpy(db, hb, sizeof(hb), cudaMemcpyHostToDevice);
  k<<<1, 32>>>(da, db, dd);
  float hd[128];
  cudaMemcpy(hd, dd, sizeof(hd), cudaMemcpyDeviceToHost);
  double maxrel = 0;
  int bad = 0;
  for (int lane = 0; lane < 32; ++lane) {
    int g = lane >> 2, c = lane & 3;
    for (int i = 0; i < 4; ++i) {
      int r = g + 8 * (i >> 1), n = 2 * c + (i & 1);
      double ref = 0;
      for (int kk = 0; kk < 16; ++kk) ref += double(Ah[r][kk]) * ldexp(1.0, -24) * double(__half2float(Bh[kk][n]));
      double got = hd[lane * 4 + i];
      double rel = fabs(got - ref) / (fabs(ref) + 1e-30);
      if (rel > maxrel) maxrel = rel;
      if (rel > 1e-6) ++bad;
    }
  }
  printf("subnormal-A mma: max rel err %.3e, bad %d / 128  (%s)\n", maxrel, bad, bad ? "NOT exact" : "exact to fp32 rounding");
  return 0;
}
