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
  pdl_launch_dependents();
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
// system fence plus a flag write plus a read round trip.  Two slots as above: a peer can run at
// most one call ahead, because its next call needs this rank's packets of that call.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void st_peer128(void* p, const uint4& v) {
  asm volatile("st.relaxed.sys.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

template <typename T, int kWorld>
__global__ void __launch_bounds__(kArThreads, 1)
allreduce_ll_kernel(const uint4* in, uint4* out, void* const* __restrict__ peers, int rank, int nvec, long long ll_off,
                    long long ll_src_stride) {
  __shared__ uint8_t* s_peer[kArMaxWorld];
  const int tid = threadIdx.x;
  if (tid < kWorld) s_peer[tid] = reinterpret_cast<uint8_t*>(peers[tid]);
  __syncthreads();
  uint8_t* mine = s_peer[rank];
  ArCtrl* ctrl = reinterpret_cast<ArCtrl*>(mine);
  const int i = blockIdx.x * kArThreads + tid;  // this thread's 16-byte vector of the message

  pdl_wait();  // the partial sums (and the epoch word) belong to the previous kernel until here
  const unsigned int epoch = *reinterpret_cast<volatile unsigned int*>(&ctrl->epoch) + 1u;
  const long long slot = ll_off + (long long)(epoch & 1u) * (ll_src_stride * kArMaxWorld);
  if (i < nvec) {
    const uint4 mine_v = in[i];
    // push: vector i becomes two 16-byte words {d0, e, d1, e} {d2, e, d3, e} in every peer's [slot][rank]
    const uint4 p0 = make_uint4(mine_v.x, epoch, mine_v.y, epoch), p1 = make_uint4(mine_v.z, epoch, mine_v.w, epoch);
    // staggered destinations (rank + 1, rank + 2, ...): at any instant the world's pushes fan out
    // over all receivers instead of converging on one NVLink ingress
#pragma unroll
    for (int j = 1; j < kWorld; ++j) {
      const int r = (rank + j) & (kWorld - 1);
      uint4* dst = reinterpret_cast<uint4*>(s_peer[r] + slot + rank * ll_src_stride) + 2 * i;
      st_peer128(dst, p0);
      st_peer128(dst + 1, p1);
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
    out[i] = sum_world<T, kWorld>(v);
  }
  pdl_launch_dependents();
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

constexpr int64_t kArLLMaxBytes = 512 << 10;                              // largest message on the packet path
constexpr int64_t kArLLSrcStride = 2 * kArLLMaxBytes;                     // one source rank's packets of one slot
constexpr int64_t kArLLRegion = 2 * kArMaxWorld * kArLLSrcStride;         // two slots x eight sources = 16 MB

template <typename T, int kWorld>
static int ar_launch_ll(const void* in, void* out, void* const* peers, int rank, int nvec, int64_t ll_off, cudaStream_t st) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(unsigned((nvec + kArThreads - 1) / kArThreads));
  cfg.blockDim = dim3(kArThreads);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  AWQ_CUDA_CHECK(cudaLaunchKernelEx(&cfg, allreduce_ll_kernel<T, kWorld>, (const uint4*)in, (uint4*)out, peers, rank, nvec,
                                    (long long)ll_off, (long long)kArLLSrcStride),
                 "awq_b200_allreduce: launch (packets)");
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

static int g_ar_ll = 1;  // debug: 0 forces the flag-exchange kernel for every size

template <typename T>
static int ar_world(const void* in, void* out, void* const* peers, int rank, int world, int nvec, int64_t slot_bytes,
                    cudaStream_t st) {
  // the packet path moves 2 x bytes x (world - 1) per rank: 512 KB messages at 2 ranks, 256 KB at 4;
  // at 8 ranks the flag exchange measured faster at every size (8.1 vs 9.1 us at 8 KB)
  if (g_ar_ll && world <= 4 && int64_t(nvec) * 16 * world <= 2 * kArLLMaxBytes) {
    const int64_t ll_off = kArDataOff + 2 * slot_bytes;
    switch (world) {
      case 2: return ar_launch_ll<T, 2>(in, out, peers, rank, nvec, ll_off, st);
      case 4: return ar_launch_ll<T, 4>(in, out, peers, rank, nvec, ll_off, st);
      case 8: return ar_launch_ll<T, 8>(in, out, peers, rank, nvec, ll_off, st);
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

}  // namespace awq

using namespace awq;

extern "C" void awq_b200_debug_set_allreduce_packets(int enabled) { g_ar_ll = enabled ? 1 : 0; }

extern "C" int64_t awq_b200_allreduce_buffer_bytes(int64_t max_message_bytes) {
  if (max_message_bytes < 16) max_message_bytes = 16;
  const int64_t slot = (max_message_bytes + 127) / 128 * 128;
  return kArDataOff + 2 * slot + kArLLRegion;
}

extern "C" int awq_b200_allreduce(const void* in, void* out, int64_t numel, int dtype, void* const* peer_buffers_dev,
                                  int64_t buffer_bytes, int rank, int world, void* stream) {
  if (numel == 0) return 0;
  AWQ_CHECK(dtype == AWQ_B200_F16 || dtype == AWQ_B200_BF16, "awq_b200_allreduce: bad dtype %d", dtype);
  AWQ_CHECK(in && out && peer_buffers_dev, "awq_b200_allreduce: null pointer");
  AWQ_CHECK(world >= 2 && world <= kArMaxWorld && rank >= 0 && rank < world, "awq_b200_allreduce: bad rank %d / world %d", rank, world);
  AWQ_CHECK(numel > 0 && numel % 8 == 0, "awq_b200_allreduce: numel (%lld) must be a positive multiple of 8", (long long)numel);
  AWQ_CHECK(((uintptr_t)in % 16) == 0 && ((uintptr_t)out % 16) == 0, "awq_b200_allreduce: in / out must be 16-byte aligned");
  const int64_t slot_bytes = (buffer_bytes - kArDataOff - kArLLRegion) / 2 / 128 * 128;
  AWQ_CHECK(buffer_bytes >= kArDataOff + kArLLRegion + 256 && numel * 2 <= slot_bytes,
            "awq_b200_allreduce: message of %lld bytes does not fit the staging buffer (%lld bytes per slot)",
            (long long)(numel * 2), (long long)slot_bytes);
  if (int rc = awq_b200_check_device()) return rc;
  const int nvec = int(numel / 8);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == AWQ_B200_F16) return ar_world<__half>(in, out, peer_buffers_dev, rank, world, nvec, slot_bytes, st);
  return ar_world<__nv_bfloat16>(in, out, peer_buffers_dev, rank, world, nvec, slot_bytes, st);
}
