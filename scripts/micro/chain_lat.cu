// chain_lat.cu -- development microbenchmark: what does a kernel BOUNDARY cost on B200 inside a
// CUDA graph, with and without programmatic dependent launch (PDL), and do two consecutive
// kernels overlap when each leaves room for the other on the SM?
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o chain_lat chain_lat.cu && ./chain_lat
//
// Kernel body: [launch_dependents] -> spin `pre` ns (stands for the weight prefetch that does not
// depend on the previous kernel) -> [griddepcontrol.wait] -> read the previous kernel's output,
// spin `post` ns (stands for the dependent part), write own output.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x)                                                                        \
  do {                                                                               \
    cudaError_t e = (x);                                                             \
    if (e != cudaSuccess) {                                                          \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); \
      exit(1);                                                                       \
    }                                                                                \
  } while (0)

__device__ __forceinline__ unsigned long long gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__global__ void chain_kernel(const float* __restrict__ in, float* __restrict__ out, int pre_ns, int post_ns, int pdl) {
  extern __shared__ uint8_t smem[];
  if (pdl) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  if (pre_ns > 0) {
    const unsigned long long t0 = gtime();
    while (gtime() - t0 < (unsigned long long)pre_ns) {
    }
  }
  if (pdl) asm volatile("griddepcontrol.wait;" ::: "memory");
  float v = in[(blockIdx.x * blockDim.x + threadIdx.x) & 1023];
  if (post_ns > 0) {
    const unsigned long long t0 = gtime();
    while (gtime() - t0 < (unsigned long long)post_ns) {
    }
  }
  if (threadIdx.x == 0) smem[0] = 1;
  out[(blockIdx.x * blockDim.x + threadIdx.x) & 1023] = v + 1.f;
}

static float run(int grid, int threads, int smem, int pre, int post, int pdl, int nk) {
  float *a, *b;
  CK(cudaMalloc(&a, 4096));
  CK(cudaMalloc(&b, 4096));
  CK(cudaMemset(a, 0, 4096));
  CK(cudaMemset(b, 0, 4096));
  CK(cudaFuncSetAttribute(chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  cudaStream_t st;
  CK(cudaStreamCreate(&st));
  cudaGraph_t graph;
  cudaGraphExec_t exec;
  CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeGlobal));
  for (int i = 0; i < nk; ++i) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(threads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    CK(cudaLaunchKernelEx(&cfg, chain_kernel, (const float*)(i & 1 ? b : a), (i & 1 ? a : b), pre, post, pdl));
  }
  CK(cudaStreamEndCapture(st, &graph));
  CK(cudaGraphInstantiate(&exec, graph, 0));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  for (int i = 0; i < 3; ++i) CK(cudaGraphLaunch(exec, st));
  CK(cudaStreamSynchronize(st));
  CK(cudaEventRecord(e0, st));
  const int reps = 10;
  for (int i = 0; i < reps; ++i) CK(cudaGraphLaunch(exec, st));
  CK(cudaEventRecord(e1, st));
  CK(cudaStreamSynchronize(st));
  float ms;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  CK(cudaGraphExecDestroy(exec));
  CK(cudaGraphDestroy(graph));
  CK(cudaStreamDestroy(st));
  CK(cudaFree(a));
  CK(cudaFree(b));
  return ms * 1e3f / (reps * nk);
}

int main() {
  const int nk = 256;
  printf("# grid threads smem_KB pre_ns post_ns pdl -> us per kernel (graph of %d kernels)\n", nk);
  const int grids[] = {148, 296, 1792};
  for (int gi = 0; gi < 3; ++gi)
    for (int pdl = 0; pdl < 2; ++pdl)
      printf("grid=%4d thr=256 smem=  8 pre=    0 post=    0 pdl=%d  %.3f us\n", grids[gi], pdl, run(grids[gi], 256, 8 << 10, 0, 0, pdl, nk));
  // overlap: does pre(N+1) run under post(N)?  one CTA per SM, each using < half / > half of the SM's shared memory
  const int smems[] = {100 << 10, 200 << 10};
  for (int si = 0; si < 2; ++si)
    for (int pdl = 0; pdl < 2; ++pdl) {
      printf("grid= 148 thr=256 smem=%3d pre= 3000 post= 3000 pdl=%d  %.3f us\n", smems[si] >> 10, pdl,
             run(148, 256, smems[si], 3000, 3000, pdl, nk));
      printf("grid= 148 thr=256 smem=%3d pre= 3000 post= 1000 pdl=%d  %.3f us\n", smems[si] >> 10, pdl,
             run(148, 256, smems[si], 3000, 1000, pdl, nk));
      printf("grid= 148 thr=256 smem=%3d pre= 1000 post= 3000 pdl=%d  %.3f us\n", smems[si] >> 10, pdl,
             run(148, 256, smems[si], 1000, 3000, pdl, nk));
    }
  // many short CTAs (non-persistent grid), 4 resident per SM
  for (int pdl = 0; pdl < 2; ++pdl)
    printf("grid=1792 thr=256 smem= 50 pre= 1000 post= 1000 pdl=%d  %.3f us\n", pdl, run(1792, 256, 50 << 10, 1000, 1000, pdl, nk));
  return 0;
}
