// awq_allreduce.cu -- one-shot all-reduce (sum) over peer memory for the row-parallel decode path.
//
// Replaces, for decode-sized messages, the collective that closes RowParallelLinear in the
// reference: tensor_model_parallel_all_reduce (python/sglang/srt/distributed/communication_op.py:11-13)
// -> GroupCoordinator.all_reduce (parallel_state.py:544-623), whose fast path is the P2P custom
// all-reduce (sgl-kernel/csrc/allreduce/custom_all_reduce.cuh:168-273: one-shot below 512 KB,
// fp32 accumulation, rank-order-invariant sums) and whose fallback is ncclAllReduce.
//
// One process per GPU.  Every rank owns a STAGING BUFFER of identical layout that all peers have
// mapped (NVLink 5 / NVSwitch peer access; the host side gets the mapping from PyTorch's
// symmetric-memory rendezvous -- plumbing -- and hands this library a device array of the
// world's base pointers):
//
//   [0, 64)            control: epoch (calls completed), done (blocks finished), error
//   [4096, 8192)       flags[block][src rank]  (u32, monotonically increasing epochs)
//   [8192, ...)        two data slots of `slot_bytes` each (call n uses slot n & 1)
//   [8192 + 2 slots, + 16 MB) packet region of the low-latency variant (below)
//
// Per call, block b of every rank owns the same 16-byte-vector range of the message:
//   1. copy its range of the partial sums into the rank's own slot (local HBM),
//   2. bar.sync, then one thread per peer: fence.acq_rel.sys + st.relaxed.sys of the new epoch into
//      flags[b][rank] of THAT peer's buffer (a 4-byte NVLink write),
//   3. one thread per peer spins (ld.acquire.sys) on its OWN flags[b][peer] until the epoch shows,
//   4. every thread sums its vectors over the world's slots in RANK order in fp32 -- the same
//      order on every rank, so all ranks hold bit-identical results -- and writes the output.
// One barrier per call: a rank re-enters slot n & 1 only in call n + 2, after the flag exchange
// of call n + 1, which every peer joins only after its reads of call n have retired (kernels on
// a stream do not overlap).  Epochs live in device memory and are advanced by the kernel itself,
// so the call is CUDA-graph capturable and replayable.  A spin that exceeds ~35 s sets the error
// word and traps (the stream fails; no output is ever produced from incomplete inputs).
#include "common.cuh"
#include "gemv_common.cuh"
#include "host_common.h"

namespace awq {

constexpr int kArThreads = 512;
constexpr int kArMaxBlocks = 64;
constexpr int kArMaxWorld = 8;
constexpr int kArFlagsOff = 4096;
constexpr int kArDataOff = 8192;
// A rank waits this many SM clocks (~35 s at 1.9 GHz) for its peers before it gives up.  Giving
// up means __trap(): the launch fails and every later CUDA call on the context reports it.  The
// kernel never writes `out` from incomplete inputs (the reference's custom all-reduce and NCCL
// simply wait; skews of seconds are legitimate: lazy module loading, host GC pauses, debuggers).
constexpr long long kArSpinLimit = 1ll << 36;

struct ArCtrl {
  unsigned int epoch;
  unsigned int done;
  unsigned int error;
};

__device__ __forceinline__ void st_flag_sys(unsigned int* p, unsigned int v) {
  asm volatile("fence.acq_rel.sys;\n\tst.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_flag_sys(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// peer data: written by another GPU since this kernel started -> no read-only / L1 path
__device__ __forceinline__ uint4 ld_peer128(const void* p) {
  uint4 v;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}

template <typename T>
__device__ __forceinline__ void acc8(float (&s)[8], const uint4& v) {
  const T* h = reinterpret_cast<const T*>(&v);
#pragma unroll
  for (int i = 0; i < 8; ++i) s[i] += Elem<T>::to_float(h[i]);
}

template <typename T, int kWorld>
__device__ __forceinline__ uint4 sum_world(const uint4 (&v)[kWorld]) {
  float s[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int r = 0; r < kWorld; ++r) acc8<T>(s, v[r]);  // rank order
  uint4 o;
  T* oh = reinterpret_cast<T*>(&o);
#pragma unroll
  for (int j = 0; j < 8; ++j) oh[j] = Elem<T>::from_float(s[j]);
  return o;
}

template <typename T, int kWorld>
__global__ void __launch_bounds__(kArThreads, 1)
allreduce_oneshot_kernel(const uint4* in, uint4* out, void* const* __restrict__ peers, int rank,
                         int nvec, int vec_per_block, long long slot_bytes) {
  __shared__ uint8_t* s_peer[kArMaxWorld];
  const int tid = threadIdx.x, b = blockIdx.x;
  if (tid < kWorld) s_peer[tid] = reinterpret_cast<uint8_t*>(peers[tid]);
  __syncthreads();
  uint8_t* mine = s_peer[rank];
  ArCtrl* ctrl = reinterpret_cast<ArCtrl*>(mine);
  const int v0 = b * vec_per_block;
  const int v1 = v0 + vec_per_block < nvec ? v0 + vec_per_block : nvec;

  // the next kernel (a linear whose weights do not depend on this collective) may start NOW: its blocks become resident,
  // issue their first weight requests and then sit in griddepcontrol.wait until this grid has completed; the
  // collective's NVLink latency and the next launch's ramp overlap instead of adding up
  pdl_launch_dependents();
  pdl_wait();  // the partial sums (and the epoch word) belong to the previous kernel until here
  // every block of this launch sees the same value: the last block to finish advances it
  const unsigned int epoch = *reinterpret_cast<volatile unsigned int*>(&ctrl->epoch) + 1u;
  const long long slot_off = kArDataOff + (long long)(epoch & 1u) * slot_bytes;
  // 1. publish this rank's partial sums
  uint4* my_slot = reinterpret_cast<uint4*>(mine + slot_off);
  for (int i = v0 + tid; i < v1; i += kArThreads) my_slot[i] = in[i];
  __syncthreads();
  // 2. tell every peer (and ourselves) that block b's range of this rank is in place
  if (tid < kWorld)
    st_flag_sys(reinterpret_cast<unsigned int*>(s_peer[tid] + kArFlagsOff) + b * kArMaxWorld + rank, epoch);
  // 3. wait for every rank's range
  if (tid < kWorld) {
    const unsigned int* f = reinterpret_cast<const unsigned int*>(mine + kArFlagsOff) + b * kArMaxWorld + tid;
    const long long t0 = clock64();
    while (int(ld_flag_sys(f) - epoch) < 0) {
      if (clock64() - t0 > kArSpinLimit) {  // a peer is gone: abort the stream, never return partial sums
        atomicExch(&ctrl->error, 1u);
        __threadfence_system();
        __trap();
      }
    }
  }
  __syncthreads();
  // 4. sum in rank order, fp32
  // two vectors per thread per round: all 2 * world NVLink reads of a round are in flight together
  for (int i = v0 + tid; i < v1; i += 2 * kArThreads) {
    const int i2 = i + kArThreads;
    const bool two = i2 < v1;
    uint4 va[kWorld], vb[kWorld];
#pragma unroll
    for (int r = 0; r < kWorld; ++r) va[r] = ld_peer128(reinterpret_cast<const uint4*>(s_peer[r] + slot_off) + i);
    if (two) {
#pragma unroll
      for (int r = 0; r < kWorld; ++r) vb[r] = ld_peer128(reinterpret_cast<const uint4*>(s_peer[r] + slot_off) + i2);
    }
    out[i] = sum_world<T, kWorld>(va);
    if (two) out[i2] = sum_world<T, kWorld>(vb);
  }
  // the last block of this launch to finish advances the epoch for the next call
  __syncthreads();
  if (tid == 0) {
    if (gridDim.x == 1) {  // single-block launch (messages <= 8 KB): nothing to count
      *reinterpret_cast<volatile unsigned int*>(&ctrl->epoch) = epoch;
    } else {
      __threadfence();
      if (atomicAdd(&ctrl->done, 1u) == gridDim.x - 1) {
        ctrl->done = 0;
        __threadfence();
        *reinterpret_cast<volatile unsigned int*>(&ctrl->epoch) = epoch;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Low-latency variant for messages <= kArLLMaxBytes (512 KB): no fence, no flag round trip.  Every 8 bytes
// that cross NVLink carry their own validity: a packet is {4 data bytes, 4-byte epoch}, written
// as 16-byte stores of two packets straight into the RECEIVER's buffer (region [slot][source
// rank]); the receiver polls its local copy until both epochs of a 16-byte word match the
// current call, then sums in rank order.  One NVLink one-way latency per call instead of a
// system fence plus a flag write plus a read round trip.
//
// Work split: SMALL blocks (128 threads, one 16-byte vector per thread), one block per SM.  A thread
// pushes 2 x (world - 1) 16-byte stores and polls as many loads per round, so a 512-thread block
// at 8 ranks pushed / polled 114 KB through ONE SM's load-store path per round (~1 us each at
// ~100 B/clk) while 146 SMs idled -- that, not NVLink, is what made 8 ranks cost 8-9 us.
//
// Epochs are PER BLOCK (block b always owns vectors [128 b, 128 b + 128) of a message, whatever
// its length, so the epoch sequence of a vector position is the same on every rank): no block of a
// launch waits for another, there is no "last block" counter at the tail, and two slots suffice as
// before -- a peer's block b can run at most one call ahead of ours, because its next call needs
// this rank's block-b packets of that call, which are sent only after this call's kernel retired.
//
// kNorm: the fused epilogue of the reference's flashinfer_allreduce_residual_rmsnorm
// (python/sglang/srt/layers/flashinfer_comm_fusion.py:126-200, called from
// RMSNorm.forward_with_allreduce_fusion, layernorm.py:255-285): a = T(sum over ranks);
// r = a + residual (fp32); residual <- T(r); out <- T(r * rsqrt(mean(r^2) + eps) * weight)
// (RMSNorm.forward_native, layernorm.py:178-222).  One CLUSTER of hidden / 1024 blocks owns one
// token row; the row's sum of squares is closed through distributed shared memory.
// kWorld == 1 is the same epilogue without a collective (fused_add_rmsnorm, layernorm.py:122-124).
// ---------------------------------------------------------------------------------------------
constexpr int kArLLThreads = 128;
constexpr int kArBlockEpochOff = 1024;   // u32 per block of the one-shot packet kernel, [1024, 2048)
constexpr int kArBlockEpoch2Off = 2048;  // u32 per block of the two-shot packet kernel, [2048, 4096)
constexpr int kArLLMaxBlocks = (kArBlockEpoch2Off - kArBlockEpochOff) / 4;  // 256

__device__ __forceinline__ void st_peer128(void* p, const uint4& v) {
  asm volatile("st.relaxed.sys.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint32_t ar_cluster_nctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ float ar_ld_cluster_f32(uint32_t smem_addr, uint32_t cta) {
  uint32_t a;
  float v;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(a) : "r"(smem_addr), "r"(cta));
  asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ void ar_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

struct ArNormArgs {
  uint4* residual;      // T [rows, hidden], in / out
  const uint4* weight;  // T [hidden]
  int vec_per_row;      // hidden / 8
  float eps;
  float inv_hidden;
};

template <typename T, int kWorld, bool kNorm>
__global__ void __launch_bounds__(kArLLThreads)
allreduce_ll_kernel(const uint4* in, uint4* out, void* const* __restrict__ peers, int rank, int nvec, long long ll_off,
                    long long ll_src_stride, const ArNormArgs norm, int pushed) {
  __shared__ uint8_t* s_peer[kArMaxWorld];
  __shared__ float s_part[kArLLThreads / 32];
  __shared__ float s_row;  // this block's share of the row's sum of squares (read by the whole cluster)
  const int tid = threadIdx.x;
  if constexpr (kWorld > 1) {
    if (tid < kWorld) s_peer[tid] = reinterpret_cast<uint8_t*>(peers[tid]);  // written once at start-up: no dependency
    __syncthreads();
  }
  const int i = blockIdx.x * kArLLThreads + tid;  // this thread's 16-byte vector of the message
  uint4 w4 = make_uint4(0u, 0u, 0u, 0u);
  if constexpr (kNorm) {
    if (i < nvec) w4 = norm.weight[i % norm.vec_per_row];  // a parameter: not produced by the previous kernel
  }

  // the next kernel (a linear whose weights do not depend on this collective) may start NOW: its blocks become resident,
  // issue their first weight requests and then sit in griddepcontrol.wait until this grid has completed; the
  // collective's NVLink latency and the next launch's ramp overlap instead of adding up
  pdl_launch_dependents();
  pdl_wait();  // the partial sums, the residual and the epoch word belong to the previous kernel until here
  uint4 sum4 = make_uint4(0u, 0u, 0u, 0u);
  unsigned int epoch = 0;
  unsigned int* my_epoch = nullptr;
  if constexpr (kWorld == 1) {
    if (i < nvec) sum4 = in[i];
  } else {
    uint8_t* mine = s_peer[rank];
    ArCtrl* ctrl = reinterpret_cast<ArCtrl*>(mine);
    my_epoch = reinterpret_cast<unsigned int*>(mine + kArBlockEpochOff) + blockIdx.x;
    epoch = *reinterpret_cast<volatile unsigned int*>(my_epoch) + 1u;
    const long long slot = ll_off + (long long)(epoch & 1u) * (ll_src_stride * kArMaxWorld);
    if (i < nvec) {
      const uint4 mine_v = in[i];
      // push: vector i becomes two 16-byte words {d0, e, d1, e} {d2, e, d3, e} in every peer's [slot][rank]
      const uint4 p0 = make_uint4(mine_v.x, epoch, mine_v.y, epoch), p1 = make_uint4(mine_v.z, epoch, mine_v.w, epoch);
      // staggered destinations (rank + 1, rank + 2, ...): at any instant the world's pushes fan out
      // over all receivers instead of converging on one NVLink ingress
      // (pushed: the linear that produced `in` already sent these packets from its own epilogue, awq_b200_linear_allreduce)
      if (!pushed) {
#pragma unroll
        for (int j = 1; j < kWorld; ++j) {
          const int r = (rank + j) & (kWorld - 1);
          uint4* dst = reinterpret_cast<uint4*>(s_peer[r] + slot + rank * ll_src_stride) + 2 * i;
          st_peer128(dst, p0);
          st_peer128(dst + 1, p1);
        }
      }
      // pull from the local copies: the loads of ALL sources are issued before any is examined, so a
      // poll round costs one L2 latency, not world - 1 of them
      uint4 a[kWorld], c[kWorld];
      bool ok[kWorld];
#pragma unroll
      for (int r = 0; r < kWorld; ++r) {
        ok[r] = r == rank;
        a[r] = c[r] = make_uint4(0u, 0u, 0u, 0u);
      }
      const uint4* src0 = reinterpret_cast<const uint4*>(mine + slot) + 2 * i;
      const long long t0 = clock64();
      bool all;
      do {
#pragma unroll
        for (int r = 0; r < kWorld; ++r) {
          if (!ok[r]) {
            const uint4* src = reinterpret_cast<const uint4*>(reinterpret_cast<const uint8_t*>(src0) + r * ll_src_stride);
            a[r] = ld_peer128(src);
            c[r] = ld_peer128(src + 1);
          }
        }
        all = true;
#pragma unroll
        for (int r = 0; r < kWorld; ++r) {
          if (!ok[r]) {
            ok[r] = a[r].y == epoch && a[r].w == epoch && c[r].y == epoch && c[r].w == epoch;
            all = all && ok[r];
          }
        }
        if (!all && clock64() - t0 > kArSpinLimit) {  // a peer is gone: abort the stream, never return partial sums
          atomicExch(&ctrl->error, 1u);
          __threadfence_system();
          __trap();
        }
      } while (!all);
      uint4 v[kWorld];  // a / c keep the words that passed the check
#pragma unroll
      for (int r = 0; r < kWorld; ++r) v[r] = r == rank ? mine_v : make_uint4(a[r].x, a[r].z, c[r].x, c[r].z);
      sum4 = sum_world<T, kWorld>(v);
    }
  }

  if constexpr (!kNorm) {
    if (i < nvec) out[i] = sum4;
  } else {
    // r = T(sum) + residual in fp32; residual <- T(r); the row's mean square over the cluster
    float r[8];
    float ss = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) r[j] = 0.f;
    if (i < nvec) {
      const uint4 res4 = norm.residual != nullptr ? norm.residual[i] : make_uint4(0u, 0u, 0u, 0u);  // null: plain RMSNorm
      const T* sh = reinterpret_cast<const T*>(&sum4);
      const T* rh = reinterpret_cast<const T*>(&res4);
      uint4 o;
      T* oh = reinterpret_cast<T*>(&o);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        r[j] = Elem<T>::to_float(sh[j]) + Elem<T>::to_float(rh[j]);
        ss = fmaf(r[j], r[j], ss);
        oh[j] = Elem<T>::from_float(r[j]);
      }
      if (norm.residual != nullptr) norm.residual[i] = o;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, d);
    if ((tid & 31) == 0) s_part[tid >> 5] = ss;
    __syncthreads();
    if (tid == 0) {
      float t = 0.f;
#pragma unroll
      for (int w = 0; w < kArLLThreads / 32; ++w) t += s_part[w];
      s_row = t;
    }
    ar_cluster_sync();
    float tot = 0.f;
    const uint32_t ncta = ar_cluster_nctarank(), s_row_addr = smem_u32(&s_row);
    for (uint32_t cta = 0; cta < ncta; ++cta) tot += ar_ld_cluster_f32(s_row_addr, cta);  // fixed order: same bits in every block
    const float rs = rsqrtf(tot * norm.inv_hidden + norm.eps);
    if (i < nvec) {
      const T* wh = reinterpret_cast<const T*>(&w4);
      uint4 o;
      T* oh = reinterpret_cast<T*>(&o);
#pragma unroll
      for (int j = 0; j < 8; ++j) oh[j] = Elem<T>::from_float(r[j] * rs * Elem<T>::to_float(wh[j]));
      out[i] = o;
    }
    ar_cluster_sync();  // nobody exits while a peer block may still read its s_row
  }
  if constexpr (kWorld > 1) {
    __syncthreads();  // every thread of the block is past its poll loop
    if (tid == 0) *reinterpret_cast<volatile unsigned int*>(my_epoch) = epoch;
  }
}

constexpr int64_t kArLLMaxBytes = 512 << 10;                              // largest message on the packet path
constexpr int64_t kArLLSrcStride = 2 * kArLLMaxBytes;                     // one source rank's packets of one slot
constexpr int64_t kArLLRegion = 2 * kArMaxWorld * kArLLSrcStride;         // two slots x eight sources = 16 MB
static_assert(kArLLMaxBytes / 16 / kArLLThreads <= kArLLMaxBlocks, "per-block epoch words");

template <typename T, int kWorld, bool kNorm>
static int ar_launch_ll(const void* in, void* out, void* const* peers, int rank, int nvec, int64_t ll_off, cudaStream_t st,
                        const ArNormArgs& norm, int pushed = 0) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(unsigned((nvec + kArLLThreads - 1) / kArLLThreads));
  cfg.blockDim = dim3(kArLLThreads);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (pdl_enabled()) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  if (kNorm) {  // one cluster per token row
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = unsigned(norm.vec_per_row / kArLLThreads);
    attr[na].val.clusterDim.y = 1;
    attr[na].val.clusterDim.z = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = unsigned(na);
  AWQ_CUDA_CHECK(cudaLaunchKernelEx(&cfg, allreduce_ll_kernel<T, kWorld, kNorm>, (const uint4*)in, (uint4*)out, peers, rank,
                                    nvec, (long long)ll_off, (long long)kArLLSrcStride, norm, pushed),
                 "awq_b200_allreduce: launch (packets)");
  return 0;
}

// ---------------------------------------------------------------------------------------------
// Two-shot packet variant for the larger decode messages at 4 / 8 ranks (M = 16 .. 64 tokens x hidden): a
// reduce-scatter and an all-gather, both with self-validating packets.  Rank r owns chunk r (nvec / world
// vectors): every rank pushes its share of chunk p straight into rank p's buffer, rank r sums its chunk in rank
// order in fp32, rounds once, and pushes the result into every peer's buffer.  2 x (world - 1) / world of the
// message crosses each NVLink direction instead of (world - 1) x (one-shot): at 8 ranks and 256 KB that is
// 0.9 MB per rank instead of 3.5 MB (packets) or 1.8 MB (peer reads), for one more one-way latency.  Every rank
// ends up with the owner's bits, which are the same fp32 rank-order sum the one-shot kernels produce.
// Thread j of block b handles vector j of EVERY chunk; epochs per block, two slots, own region: the protocol
// argument is the one given for the one-shot packet kernel, applied to each phase.
// ---------------------------------------------------------------------------------------------
constexpr int64_t kAr2MaxBytes = 1 << 20;                              // largest message
constexpr int64_t kAr2SrcStride = 2 * (kAr2MaxBytes / 4);              // packets of one source: chunk <= message / 4 (world >= 4)
constexpr int64_t kAr2PhaseStride = kArMaxWorld * kAr2SrcStride;       // reduce-scatter | all-gather
constexpr int64_t kAr2Region = 2 * 2 * kAr2PhaseStride;                // two slots x two phases x eight sources = 16 MB
static_assert(kAr2MaxBytes / 16 / 4 / kArLLThreads <= (kArFlagsOff - kArBlockEpoch2Off) / 4, "per-block epoch words");

// polls the local copies of vector `idx` of every source but `rank` until all carry `epoch`; a[r] / c[r] = the two
// 16-byte words {d0, e, d1, e} {d2, e, d3, e}
template <int kWorld>
__device__ __forceinline__ void ar_poll_packets(const uint8_t* region, long long src_stride, int idx, unsigned int epoch,
                                                int rank, ArCtrl* ctrl, uint4 (&a)[kWorld], uint4 (&c)[kWorld]) {
  bool ok[kWorld];
#pragma unroll
  for (int r = 0; r < kWorld; ++r) {
    ok[r] = r == rank;
    a[r] = c[r] = make_uint4(0u, 0u, 0u, 0u);
  }
  const long long t0 = clock64();
  bool all;
  do {
#pragma unroll
    for (int r = 0; r < kWorld; ++r) {
      if (!ok[r]) {
        const uint4* src = reinterpret_cast<const uint4*>(region + r * src_stride) + 2 * idx;
        a[r] = ld_peer128(src);
        c[r] = ld_peer128(src + 1);
      }
    }
    all = true;
#pragma unroll
    for (int r = 0; r < kWorld; ++r) {
      if (!ok[r]) {
        ok[r] = a[r].y == epoch && a[r].w == epoch && c[r].y == epoch && c[r].w == epoch;
        all = all && ok[r];
      }
    }
    if (!all && clock64() - t0 > kArSpinLimit) {  // a peer is gone: abort the stream, never return partial sums
      atomicExch(&ctrl->error, 1u);
      __threadfence_system();
      __trap();
    }
  } while (!all);
}

template <typename T, int kWorld>
__global__ void __launch_bounds__(kArLLThreads)
allreduce_twoshot_kernel(const uint4* in, uint4* out, void* const* __restrict__ peers, int rank, int chunk, long long two_off) {
  __shared__ uint8_t* s_peer[kArMaxWorld];
  const int tid = threadIdx.x;
  if (tid < kWorld) s_peer[tid] = reinterpret_cast<uint8_t*>(peers[tid]);
  __syncthreads();
  uint8_t* mine = s_peer[rank];
  ArCtrl* ctrl = reinterpret_cast<ArCtrl*>(mine);
  unsigned int* my_epoch = reinterpret_cast<unsigned int*>(mine + kArBlockEpoch2Off) + blockIdx.x;
  const int j = blockIdx.x * kArLLThreads + tid;  // this thread's vector of every chunk

  // the next kernel (a linear whose weights do not depend on this collective) may start NOW: its blocks become resident,
  // issue their first weight requests and then sit in griddepcontrol.wait until this grid has completed; the
  // collective's NVLink latency and the next launch's ramp overlap instead of adding up
  pdl_launch_dependents();
  pdl_wait();  // the partial sums (and the epoch word) belong to the previous kernel until here
  const unsigned int epoch = *reinterpret_cast<volatile unsigned int*>(my_epoch) + 1u;
  const long long rs = two_off + (long long)(epoch & 1u) * (2 * kAr2PhaseStride), ag = rs + kAr2PhaseStride;
  if (j < chunk) {
    // ---- reduce-scatter: my share of chunk p goes to rank p ----
    uint4 v[kWorld];
#pragma unroll
    for (int p = 0; p < kWorld; ++p) v[p] = in[p * chunk + j];
#pragma unroll
    for (int jj = 1; jj < kWorld; ++jj) {
      const int p = (rank + jj) & (kWorld - 1);
      uint4* dst = reinterpret_cast<uint4*>(s_peer[p] + rs + rank * kAr2SrcStride) + 2 * j;
      st_peer128(dst, make_uint4(v[p].x, epoch, v[p].y, epoch));
      st_peer128(dst + 1, make_uint4(v[p].z, epoch, v[p].w, epoch));
    }
    uint4 a[kWorld], c[kWorld];
    ar_poll_packets<kWorld>(mine + rs, kAr2SrcStride, j, epoch, rank, ctrl, a, c);
    const uint4 own = v[rank];
#pragma unroll
    for (int r = 0; r < kWorld; ++r) v[r] = r == rank ? own : make_uint4(a[r].x, a[r].z, c[r].x, c[r].z);
    const uint4 red = sum_world<T, kWorld>(v);  // rank order, fp32, one rounding: the bits every rank will hold
    // ---- all-gather: the reduced vector goes to everybody ----
    const uint4 q0 = make_uint4(red.x, epoch, red.y, epoch), q1 = make_uint4(red.z, epoch, red.w, epoch);
#pragma unroll
    for (int jj = 1; jj < kWorld; ++jj) {
      const int p = (rank + jj) & (kWorld - 1);
      uint4* dst = reinterpret_cast<uint4*>(s_peer[p] + ag + rank * kAr2SrcStride) + 2 * j;
      st_peer128(dst, q0);
      st_peer128(dst + 1, q1);
    }
    out[rank * chunk + j] = red;
    ar_poll_packets<kWorld>(mine + ag, kAr2SrcStride, j, epoch, rank, ctrl, a, c);
#pragma unroll
    for (int r = 0; r < kWorld; ++r)
      if (r != rank) out[r * chunk + j] = make_uint4(a[r].x, a[r].z, c[r].x, c[r].z);
  }
  __syncthreads();  // every thread of the block is past its poll loops
  if (tid == 0) *reinterpret_cast<volatile unsigned int*>(my_epoch) = epoch;
}

template <typename T, int kWorld>
static int ar_launch_twoshot(const void* in, void* out, void* const* peers, int rank, int nvec, int64_t two_off, cudaStream_t st) {
  const int chunk = nvec / kWorld;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(unsigned((chunk + kArLLThreads - 1) / kArLLThreads));
  cfg.blockDim = dim3(kArLLThreads);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  AWQ_CUDA_CHECK(cudaLaunchKernelEx(&cfg, allreduce_twoshot_kernel<T, kWorld>, (const uint4*)in, (uint4*)out, peers, rank, chunk,
                                    (long long)two_off),
                 "awq_b200_allreduce: launch (two-shot packets)");
  return 0;
}

template <typename T, int kWorld>
static int ar_launch(const void* in, void* out, void* const* peers, int rank, int nvec, int64_t slot_bytes, cudaStream_t st) {
  // one 16-byte vector per thread up to 512 KB (one NVLink round trip), two at 1 MB
  int blocks = (nvec + kArThreads - 1) / kArThreads;
  if (blocks < 1) blocks = 1;
  if (blocks > kArMaxBlocks) blocks = kArMaxBlocks;
  const int vpb = (nvec + blocks - 1) / blocks;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(unsigned(blocks));
  cfg.blockDim = dim3(kArThreads);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  AWQ_CUDA_CHECK(cudaLaunchKernelEx(&cfg, allreduce_oneshot_kernel<T, kWorld>, (const uint4*)in, (uint4*)out, peers, rank, nvec,
                                    vpb, (long long)slot_bytes),
                 "awq_b200_allreduce: launch");
  return 0;
}

// debug: 0 = flag-exchange kernel for every size, 1 = default gates, 2 = one-shot packets whenever they fit,
// 3 = two-shot packets whenever they fit
static int g_ar_ll = 1;

// the one-shot packet path moves 2 x bytes x (world - 1) per rank: 512 KB messages at 2 ranks, 256 KB at 4, 128 KB at 8
static bool ar_use_packets(int world, int nvec) {
  const int64_t bytes = int64_t(nvec) * 16;
  if (g_ar_ll == 0 || g_ar_ll == 3 || bytes > kArLLMaxBytes) return false;
  return g_ar_ll == 2 || bytes * world <= 2 * kArLLMaxBytes;
}
static int64_t g_ar_two_min[2] = {192 << 10, 96 << 10};  // smallest two-shot message at 4 / 8 ranks (measured cross-over)
static bool ar_use_twoshot(int world, int nvec) {
  const int64_t bytes = int64_t(nvec) * 16;
  if (world < 4 || nvec % world != 0 || bytes > kAr2MaxBytes || g_ar_ll == 0 || g_ar_ll == 2) return false;
  return g_ar_ll == 3 || bytes >= g_ar_two_min[world == 4 ? 0 : 1];
}

template <typename T>
static int ar_world(const void* in, void* out, void* const* peers, int rank, int world, int nvec, int64_t slot_bytes,
                    cudaStream_t st) {
  if (ar_use_twoshot(world, nvec)) {
    const int64_t two_off = kArDataOff + 2 * slot_bytes + kArLLRegion;
    if (world == 4) return ar_launch_twoshot<T, 4>(in, out, peers, rank, nvec, two_off, st);
    return ar_launch_twoshot<T, 8>(in, out, peers, rank, nvec, two_off, st);
  }
  if (ar_use_packets(world, nvec)) {
    const int64_t ll_off = kArDataOff + 2 * slot_bytes;
    const ArNormArgs none = {nullptr, nullptr, 0, 0.f, 0.f};
    switch (world) {
      case 2: return ar_launch_ll<T, 2, false>(in, out, peers, rank, nvec, ll_off, st, none);
      case 4: return ar_launch_ll<T, 4, false>(in, out, peers, rank, nvec, ll_off, st, none);
      case 8: return ar_launch_ll<T, 8, false>(in, out, peers, rank, nvec, ll_off, st, none);
      default: return set_error(1, "awq_b200_allreduce: world size %d (supported: 2, 4, 8)", world);
    }
  }
  switch (world) {
    case 2: return ar_launch<T, 2>(in, out, peers, rank, nvec, slot_bytes, st);
    case 4: return ar_launch<T, 4>(in, out, peers, rank, nvec, slot_bytes, st);
    case 8: return ar_launch<T, 8>(in, out, peers, rank, nvec, slot_bytes, st);
    default: return set_error(1, "awq_b200_allreduce: world size %d (supported: 2, 4, 8)", world);
  }
}

// ---------------------------------------------------------------------------------------------
// residual add + RMSNorm for shapes the cluster epilogue above does not take (hidden not a multiple
// of 1024, or above 8192): one block per token row, two passes over the row (the second one hits L1 / L2).
// Same arithmetic: r = x + residual in fp32, residual <- T(r), out <- T(r * rsqrt(mean r^2 + eps) * w).
// ---------------------------------------------------------------------------------------------
constexpr int kNormThreads = 256;

template <typename T>
__global__ void __launch_bounds__(kNormThreads)
residual_rmsnorm_kernel(const uint4* x, uint4* residual, const uint4* __restrict__ weight, uint4* out, int vec_per_row,
                        float eps, float inv_hidden) {
  __shared__ float s_part[kNormThreads / 32];
  __shared__ float s_tot;
  const int tid = threadIdx.x;
  const int64_t row0 = int64_t(blockIdx.x) * vec_per_row;
  pdl_wait();
  float ss = 0.f;
  for (int v = tid; v < vec_per_row; v += kNormThreads) {
    const uint4 a = x[row0 + v], b = residual != nullptr ? residual[row0 + v] : make_uint4(0u, 0u, 0u, 0u);
    const T* ah = reinterpret_cast<const T*>(&a);
    const T* bh = reinterpret_cast<const T*>(&b);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float r = Elem<T>::to_float(ah[j]) + Elem<T>::to_float(bh[j]);
      ss = fmaf(r, r, ss);
    }
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, d);
  if ((tid & 31) == 0) s_part[tid >> 5] = ss;
  __syncthreads();
  if (tid == 0) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < kNormThreads / 32; ++w) t += s_part[w];
    s_tot = t;
  }
  __syncthreads();
  const float rs = rsqrtf(s_tot * inv_hidden + eps);
  for (int v = tid; v < vec_per_row; v += kNormThreads) {
    const uint4 a = x[row0 + v], b = residual != nullptr ? residual[row0 + v] : make_uint4(0u, 0u, 0u, 0u), w = weight[v];
    const T* ah = reinterpret_cast<const T*>(&a);
    const T* bh = reinterpret_cast<const T*>(&b);
    const T* wh = reinterpret_cast<const T*>(&w);
    uint4 ro, oo;
    T* rh = reinterpret_cast<T*>(&ro);
    T* oh = reinterpret_cast<T*>(&oo);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float r = Elem<T>::to_float(ah[j]) + Elem<T>::to_float(bh[j]);
      rh[j] = Elem<T>::from_float(r);
      oh[j] = Elem<T>::from_float(r * rs * Elem<T>::to_float(wh[j]));
    }
    if (residual != nullptr) residual[row0 + v] = ro;
    out[row0 + v] = oo;
  }
  pdl_launch_dependents();
}

// SiluAndMul for the unfused sequence: out[r, j] = silu(in[r, j]) * in[r, d + j], evaluated in fp32, one rounding
// (reference op sgl_kernel.silu_and_mul, sgl-kernel/csrc/common_extension.cc:78-79; python/sglang/srt/layers/activation.py:72-77)
template <typename T>
__global__ void __launch_bounds__(256)
silu_and_mul_kernel(const uint4* __restrict__ in, uint4* __restrict__ out, int vec_per_row /* d / 8 */, long long nvec) {
  pdl_wait();
  const long long i = blockIdx.x * 256ll + threadIdx.x;
  if (i < nvec) {
    const long long r = i / vec_per_row, v = i - r * vec_per_row;
    const uint4 g4 = in[r * 2 * vec_per_row + v], u4 = in[r * 2 * vec_per_row + vec_per_row + v];
    const T* gh = reinterpret_cast<const T*>(&g4);
    const T* uh = reinterpret_cast<const T*>(&u4);
    uint4 o;
    T* oh = reinterpret_cast<T*>(&o);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float g = Elem<T>::to_float(gh[j]);
      oh[j] = Elem<T>::from_float(g / (1.f + __expf(-g)) * Elem<T>::to_float(uh[j]));
    }
    out[i] = o;
  }
  pdl_launch_dependents();
}

static bool norm_cluster_shape(int64_t hidden) { return hidden % (8 * kArLLThreads) == 0 && hidden / (8 * kArLLThreads) <= 8; }

template <typename T>
static int norm_launch(const void* x, void* residual, const void* weight, void* out, int64_t rows, int64_t hidden, float eps,
                       cudaStream_t st) {
  const ArNormArgs norm = {(uint4*)residual, (const uint4*)weight, int(hidden / 8), eps, 1.0f / float(hidden)};
  if (norm_cluster_shape(hidden))
    return ar_launch_ll<T, 1, true>(x, out, nullptr, 0, int(rows * hidden / 8), 0, st, norm);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(unsigned(rows));
  cfg.blockDim = dim3(kNormThreads);
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  AWQ_CUDA_CHECK(cudaLaunchKernelEx(&cfg, residual_rmsnorm_kernel<T>, (const uint4*)x, (uint4*)residual, (const uint4*)weight,
                                    (uint4*)out, norm.vec_per_row, eps, norm.inv_hidden),
                 "awq_b200_residual_rmsnorm: launch");
  return 0;
}

template <typename T>
static int ar_norm_world(const void* in, void* residual, const void* weight, void* out, int64_t rows, int64_t hidden, float eps,
                         void* const* peers, int rank, int world, int64_t slot_bytes, cudaStream_t st) {
  const int nvec = int(rows * hidden / 8);
  if (norm_cluster_shape(hidden) && ar_use_packets(world, nvec) && !ar_use_twoshot(world, nvec)) {
    const ArNormArgs norm = {(uint4*)residual, (const uint4*)weight, int(hidden / 8), eps, 1.0f / float(hidden)};
    const int64_t ll_off = kArDataOff + 2 * slot_bytes;
    switch (world) {
      case 2: return ar_launch_ll<T, 2, true>(in, out, peers, rank, nvec, ll_off, st, norm);
      case 4: return ar_launch_ll<T, 4, true>(in, out, peers, rank, nvec, ll_off, st, norm);
      case 8: return ar_launch_ll<T, 8, true>(in, out, peers, rank, nvec, ll_off, st, norm);
      default: return set_error(1, "awq_b200_allreduce_residual_rmsnorm: world size %d (supported: 2, 4, 8)", world);
    }
  }
  // two launches: the collective into `out`, then the epilogue on it (same bits as the fused kernel when the
  // cluster epilogue takes the shape)
  if (int rc = ar_world<T>(in, out, peers, rank, world, nvec, slot_bytes, st)) return rc;
  return norm_launch<T>(out, residual, weight, out, rows, hidden, eps, st);
}

}  // namespace awq

using namespace awq;

extern "C" void awq_b200_debug_set_allreduce_packets(int mode) { g_ar_ll = mode < 0 ? 0 : (mode > 3 ? 3 : mode); }

extern "C" int64_t awq_b200_allreduce_buffer_bytes(int64_t max_message_bytes) {
  if (max_message_bytes < 16) max_message_bytes = 16;
  const int64_t slot = (max_message_bytes + 127) / 128 * 128;
  return kArDataOff + 2 * slot + kArLLRegion + kAr2Region;
}

static int ar_check_args(const char* who, const void* in, const void* out, int64_t numel, int dtype, void* const* peers,
                         int64_t buffer_bytes, int rank, int world, int64_t* slot_bytes) {
  AWQ_CHECK(dtype == AWQ_B200_F16 || dtype == AWQ_B200_BF16, "%s: bad dtype %d", who, dtype);
  AWQ_CHECK(in && out && peers, "%s: null pointer", who);
  AWQ_CHECK(world >= 2 && world <= kArMaxWorld && rank >= 0 && rank < world, "%s: bad rank %d / world %d", who, rank, world);
  AWQ_CHECK(numel > 0 && numel % 8 == 0, "%s: numel (%lld) must be a positive multiple of 8", who, (long long)numel);
  AWQ_CHECK(((uintptr_t)in % 16) == 0 && ((uintptr_t)out % 16) == 0, "%s: in / out must be 16-byte aligned", who);
  *slot_bytes = (buffer_bytes - kArDataOff - kArLLRegion - kAr2Region) / 2 / 128 * 128;
  AWQ_CHECK(buffer_bytes >= kArDataOff + kArLLRegion + kAr2Region + 256 && numel * 2 <= *slot_bytes,
            "%s: message of %lld bytes does not fit the staging buffer (%lld bytes per slot)", who, (long long)(numel * 2),
            (long long)*slot_bytes);
  return awq_b200_check_device();
}

extern "C" int awq_b200_allreduce(const void* in, void* out, int64_t numel, int dtype, void* const* peer_buffers_dev,
                                  int64_t buffer_bytes, int rank, int world, void* stream) {
  if (numel == 0) return 0;
  int64_t slot_bytes = 0;
  if (int rc = ar_check_args("awq_b200_allreduce", in, out, numel, dtype, peer_buffers_dev, buffer_bytes, rank, world, &slot_bytes))
    return rc;
  const int nvec = int(numel / 8);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == AWQ_B200_F16) return ar_world<__half>(in, out, peer_buffers_dev, rank, world, nvec, slot_bytes, st);
  return ar_world<__nv_bfloat16>(in, out, peer_buffers_dev, rank, world, nvec, slot_bytes, st);
}

static int norm_check_args(const char* who, const void* x, const void* residual, const void* weight, const void* out,
                           int64_t rows, int64_t hidden, float eps, int dtype, bool allow_no_residual) {
  AWQ_CHECK(dtype == AWQ_B200_F16 || dtype == AWQ_B200_BF16, "%s: bad dtype %d", who, dtype);
  AWQ_CHECK(x && weight && out && (residual || allow_no_residual), "%s: null pointer", who);
  AWQ_CHECK(rows > 0 && hidden > 0 && hidden % 8 == 0, "%s: hidden (%lld) must be a positive multiple of 8", who, (long long)hidden);
  AWQ_CHECK(rows * hidden / 8 < (int64_t(1) << 31), "%s: too many elements", who);
  AWQ_CHECK(eps >= 0.f, "%s: negative eps", who);
  AWQ_CHECK(((uintptr_t)x % 16) == 0 && ((uintptr_t)residual % 16) == 0 && ((uintptr_t)weight % 16) == 0 && ((uintptr_t)out % 16) == 0,
            "%s: pointers must be 16-byte aligned", who);
  return 0;
}

extern "C" int awq_b200_residual_rmsnorm(const void* x, void* residual, const void* weight, void* out, int64_t rows,
                                         int64_t hidden, float eps, int dtype, void* stream) {
  if (rows == 0) return 0;
  if (int rc = norm_check_args("awq_b200_residual_rmsnorm", x, residual, weight, out, rows, hidden, eps, dtype, true)) return rc;
  if (int rc = awq_b200_check_device()) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == AWQ_B200_F16) return norm_launch<__half>(x, residual, weight, out, rows, hidden, eps, st);
  return norm_launch<__nv_bfloat16>(x, residual, weight, out, rows, hidden, eps, st);
}

extern "C" int awq_b200_allreduce_residual_rmsnorm(const void* in, void* residual, const void* weight, void* out, int64_t rows,
                                                   int64_t hidden, float eps, int dtype, void* const* peer_buffers_dev,
                                                   int64_t buffer_bytes, int rank, int world, void* stream) {
  if (rows == 0) return 0;
  if (int rc = norm_check_args("awq_b200_allreduce_residual_rmsnorm", in, residual, weight, out, rows, hidden, eps, dtype, false)) return rc;
  int64_t slot_bytes = 0;
  if (int rc = ar_check_args("awq_b200_allreduce_residual_rmsnorm", in, out, rows * hidden, dtype, peer_buffers_dev, buffer_bytes,
                             rank, world, &slot_bytes))
    return rc;
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == AWQ_B200_F16)
    return ar_norm_world<__half>(in, residual, weight, out, rows, hidden, eps, peer_buffers_dev, rank, world, slot_bytes, st);
  return ar_norm_world<__nv_bfloat16>(in, residual, weight, out, rows, hidden, eps, peer_buffers_dev, rank, world, slot_bytes, st);
}

extern "C" int awq_b200_silu_and_mul(const void* in, void* out, int64_t rows, int64_t d, int dtype, void* stream) {
  if (rows == 0 || d == 0) return 0;
  AWQ_CHECK(rows > 0 && d > 0 && d % 8 == 0, "awq_b200_silu_and_mul: d (%lld) must be a positive multiple of 8", (long long)d);
  AWQ_CHECK(dtype == AWQ_B200_F16 || dtype == AWQ_B200_BF16, "awq_b200_silu_and_mul: bad dtype %d", dtype);
  AWQ_CHECK(in && out && ((uintptr_t)in % 16) == 0 && ((uintptr_t)out % 16) == 0, "awq_b200_silu_and_mul: pointers must be 16-byte aligned");
  if (int rc = awq_b200_check_device()) return rc;
  const long long nvec = rows * (d / 8);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(unsigned((nvec + 255) / 256));
  cfg.blockDim = dim3(256);
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  if (dtype == AWQ_B200_F16)
    AWQ_CUDA_CHECK(cudaLaunchKernelEx(&cfg, silu_and_mul_kernel<__half>, (const uint4*)in, (uint4*)out, int(d / 8), nvec),
                   "awq_b200_silu_and_mul: launch");
  else
    AWQ_CUDA_CHECK(cudaLaunchKernelEx(&cfg, silu_and_mul_kernel<__nv_bfloat16>, (const uint4*)in, (uint4*)out, int(d / 8), nvec),
                   "awq_b200_silu_and_mul: launch");
  return 0;
}

// ---------------------------------------------------------------------------------------------
// RowParallelLinear.forward in one call (python/sglang/srt/layers/linear.py:1388-1414: apply, then
// tensor_model_parallel_all_reduce): when the decode router picks a kernel with the packet-push epilogue and the message
// takes the one-shot packet path, the partial sums leave for the peers from the linear's own epilogue and the collective
// only polls its local copies and sums (identical bits: same partial values, same rank-order fp32 sum); otherwise the
// linear and the collective run back to back.
// ---------------------------------------------------------------------------------------------
namespace awq {
int gemv_pushed(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K, int64_t N, int dtype,
                const ArPush& push, cudaStream_t st);
}

static int g_ar_fuse_linear = 1;
extern "C" void awq_b200_debug_set_linear_allreduce_fusion(int on) { g_ar_fuse_linear = on ? 1 : 0; }

extern "C" int awq_b200_linear_allreduce(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K,
                                         int64_t N, int dtype, void* workspace, int64_t workspace_bytes,
                                         void* const* peer_buffers_dev, int64_t buffer_bytes, int rank, int world, void* stream) {
  if (M == 0 || N == 0) return 0;
  int64_t slot_bytes = 0;
  if (int rc = ar_check_args("awq_b200_linear_allreduce", y, y, M * N, dtype, peer_buffers_dev, buffer_bytes, rank, world, &slot_bytes))
    return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const int nvec = int(M * N / 8);
  if (g_ar_fuse_linear && ar_use_packets(world, nvec) && !ar_use_twoshot(world, nvec) && awq_b200_linear_route(M, K, N) == 0) {
    const int64_t ll_off = kArDataOff + 2 * slot_bytes;
    const ArPush push = {peer_buffers_dev, (long long)ll_off, (long long)kArLLSrcStride, (long long)(kArLLSrcStride * kArMaxWorld),
                         kArBlockEpochOff, rank, world};
    const int rc = gemv_pushed(x, tiles, bias, y, M, K, N, dtype, push, st);
    if (rc > 0) return rc;
    if (rc == 0) {
      const ArNormArgs none = {nullptr, nullptr, 0, 0.f, 0.f};
#define AWQ_AR_PUSHED(TT)                                                                                     \
  switch (world) {                                                                                            \
    case 2: return ar_launch_ll<TT, 2, false>(y, y, peer_buffers_dev, rank, nvec, ll_off, st, none, 1);       \
    case 4: return ar_launch_ll<TT, 4, false>(y, y, peer_buffers_dev, rank, nvec, ll_off, st, none, 1);       \
    case 8: return ar_launch_ll<TT, 8, false>(y, y, peer_buffers_dev, rank, nvec, ll_off, st, none, 1);       \
    default: return set_error(1, "awq_b200_linear_allreduce: world size %d (supported: 2, 4, 8)", world);     \
  }
      if (dtype == AWQ_B200_F16) {
        AWQ_AR_PUSHED(__half)
      } else {
        AWQ_AR_PUSHED(__nv_bfloat16)
      }
#undef AWQ_AR_PUSHED
    }
  }
  if (int rc = awq_b200_linear(x, tiles, bias, y, M, K, N, dtype, workspace, workspace_bytes, stream)) return rc;
  return awq_b200_allreduce(y, y, M * N, dtype, peer_buffers_dev, buffer_bytes, rank, world, stream);
}
