/*
 * awq_b200.h -- C ABI of libawq_b200.so: the B200 (sm_100a) AWQ W4A16 quantized-linear path.
 *
 * This is the drop-in boundary.  Plain pointers + sizes, no torch types.  All pointers are
 * DEVICE pointers unless stated; `stream` is a cudaStream_t passed as void* (NULL = legacy
 * default stream).  Every call is asynchronous on `stream`, performs no allocation and no
 * host synchronisation, and is CUDA-graph capturable.  Return value: 0 on success, non-zero
 * on error (message via awq_b200_last_error(), thread-local).  There is NO CPU fallback: on
 * a machine without an sm_100 device every compute entry point fails with an error.
 *
 * Reference interfaces replaced (paths relative to the reference repo root):
 *   awq_b200_dequantize ......... torch::Tensor awq_dequantize(qweight, scales, qzeros)
 *                                 sgl-kernel/include/sgl_kernel_ops.h:201,
 *                                 sgl-kernel/csrc/gemm/awq_kernel.cu:186-221, registered at
 *                                 sgl-kernel/csrc/common_extension.cc:126-127
 *   awq_b200_repack ............. hook AWQLinearMethod.process_weights_after_loading,
 *                                 python/sglang/srt/layers/quantization/awq.py:429-432
 *                                 (the reference re-wraps only; the repack is new)
 *   awq_b200_gemv / _gemm /
 *   awq_b200_linear(_g) ......... the body of AWQLinearMethod.apply, awq.py:434-451:
 *                                 awq_dequantize + torch.matmul (+ out.add_(bias)), fused
 *   awq_b200_linear_silu_mul .... gate_up_proj + SiluAndMul, python/sglang/srt/models/llama.py:94-106
 *   awq_b200_silu_and_mul,
 *   awq_b200_residual_rmsnorm ... sgl_kernel silu_and_mul / rmsnorm / fused_add_rmsnorm,
 *                                 sgl-kernel/csrc/common_extension.cc:66-79
 *   awq_b200_allreduce .......... tensor_model_parallel_all_reduce / custom all-reduce,
 *                                 python/sglang/srt/distributed/parallel_state.py:544-623
 *   awq_b200_allreduce_residual_rmsnorm .. flashinfer_allreduce_residual_rmsnorm,
 *                                 python/sglang/srt/layers/flashinfer_comm_fusion.py:126-200
 *   awq_b200_moe_linear / _sum .. AWQMoEMethod.apply -> fused_marlin_moe, awq.py:661-852
 *
 * Tensor formats
 *   reference ("AWQ") layout, row-major contiguous:
 *     qweight int32 [K, N/8]   qzeros int32 [K/G, N/8]   scales T [K/G, N]   (T = fp16 | bf16)
 *     nibble p of a word holds logical column 8c + {0,2,4,6,1,3,5,7}[p]
 *   B200 tile layout (produced by awq_b200_repack; G must be 128; spec in DESIGN.md §3):
 *     T_tiles = ceil(N/128) * (K/128) self-contained records stored [n_tile][k_tile], each
 *     AWQ_B200_TILE_BYTES = 8576 (67 * 128) bytes: 8192 B packed weights | 8 x 48 B metadata
 *     blocks (16 scales + 16 4-bit zeros + pad per 16 rows).  `tiles` pointers below address
 *     that array and must be 16-byte aligned.
 */
#ifndef AWQ_B200_H_
#define AWQ_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AWQ_B200_ABI_VERSION 1

/* element type of scales / activations / outputs */
#define AWQ_B200_F16 0
#define AWQ_B200_BF16 1

#define AWQ_B200_TILE_N 128
#define AWQ_B200_TILE_K 128
#define AWQ_B200_TILE_BYTES 8576

/* ---- library ------------------------------------------------------------------------ */

int awq_b200_abi_version(void);
/* message of the last failing call on this host thread ("" if none) */
const char* awq_b200_last_error(void);
/* 0 if the current device is an sm_100 part and the kernels can run; non-zero + message otherwise */
int awq_b200_check_device(void);

/* ---- kernel 1: standalone dequantise (reference layout in, [K,N] out) ---------------- */

/* out[k, n] = RN_T((w - z) * s), bit-exact with the reference kernel.  scale_rows = K / G
 * (the reference infers G = K / scales.size(0), awq_kernel.cu:189).  N % 8 == 0. */
int awq_b200_dequantize(const int32_t* qweight, const void* scales, const int32_t* qzeros,
                        void* out, int64_t K, int64_t N, int64_t scale_rows, int dtype,
                        void* stream);

/* ---- kernel 4: load-time repack into the tile layout -------------------------------- */

/* number of tiles for a [K, N] weight: ceil(N/128) * (K/128); -1 if K % 128 != 0 */
int64_t awq_b200_num_tiles(int64_t K, int64_t N);

/* group_size must be 128, N % 8 == 0.  Rows n >= N of the last n-tile are zero filled. */
int awq_b200_repack(const int32_t* qweight, const void* scales, const int32_t* qzeros,
                    void* tiles, int64_t K, int64_t N, int dtype, void* stream);

/* same, for a MERGED gate_up weight (logical columns [0, N/2) = gate, [N/2, N) = up: the per-rank layout of
 * MergedColumnParallelLinear, python/sglang/srt/layers/linear.py:461-775): every 16-row slab of the tile
 * array holds 8 gate rows followed by the 8 up rows of the same output columns, which is what lets
 * awq_b200_linear_silu_mul apply the activation in the epilogue.  N % 16 == 0.  Tiles made by this call are
 * only for awq_b200_linear_silu_mul (awq_b200_linear on them yields the columns in slab order). */
int awq_b200_repack_gate_up(const int32_t* qweight, const void* scales, const int32_t* qzeros,
                            void* tiles, int64_t K, int64_t N, int dtype, void* stream);

/* Other group sizes of the reference op (python/sglang/srt/layers/quantization/awq_triton.py:10: -1 = K, 32, 64, 128):
 * group_size 32 / 64 -> records of awq_b200_tile_bytes(group_size) = 9728 / 8960 bytes (2 / 4 metadata sections, one per
 * sub-group) that only the *_g entry points read; a multiple of 128 that divides K (incl. K itself) -> plain records (the
 * group's scale and zero repeated in every tile it spans), usable with every entry point. */
int64_t awq_b200_tile_bytes(int64_t group_size);
int awq_b200_repack_g(const int32_t* qweight, const void* scales, const int32_t* qzeros, void* tiles, int64_t K, int64_t N,
                      int64_t group_size, int dtype, void* stream);
int awq_b200_dequantize_tiled_g(const void* tiles, void* out, int64_t K, int64_t N, int64_t group_size, int dtype,
                                void* stream);
/* y = x @ W (+ bias) for any of those group sizes.  32 / 64: decode batches (M <= 8, K >= 1024) run the sub-group
 * decode kernels (awq_gemv_sub.cu): fp16 with M <= 4 (group 64) / M <= 2 (group 32) the persistent kernel with the
 * sub-group in the MMA's token columns, the other batches and bf16 its per-weight variant; everything else the
 * tcgen05 kernel (its dequant warps scale every weight anyway); multiples of 128 route like awq_b200_linear. */
int awq_b200_linear_g(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K, int64_t N,
                      int64_t group_size, int dtype, void* workspace, int64_t workspace_bytes, void* stream);

/* inverse view for consumers that want the dense matrix (e.g. kv_b_proj absorption,
 * python/sglang/srt/models/deepseek_v2.py:3436): tile layout in, out T [K, N], same bits as
 * awq_b200_dequantize on the original tensors. */
int awq_b200_dequantize_tiled(const void* tiles, void* out, int64_t K, int64_t N, int dtype,
                              void* stream);

/* ---- kernels 2+3: y[M,N] = x[M,K] @ W (+ bias) --------------------------------------- */

/* bytes of zero-initialised device scratch the split-K paths need for problems up to
 * (max_M, max_N).  The caller allocates it once per (device, stream), zero-fills it ONCE, and
 * passes it to every call; the kernels leave it zeroed/consistent on exit. */
int64_t awq_b200_workspace_bytes(int64_t max_M, int64_t max_N);

/* decode path, 1 <= M <= 64 (intended for M <= 16): HBM-bound streaming kernel.
 * x T [M, K] row-major (x and its row pitch 16-byte aligned), y T [M, N], bias T [N] or NULL. */
int awq_b200_gemv(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K,
                  int64_t N, int dtype, void* workspace, int64_t workspace_bytes, void* stream);

/* Allocation-free decode call.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid
 * it).  It is the tcgen05 decode kernel of awq_b200_gemv (fp16, 1 <= M <= 32, same x / tiles / bias / y contract)
 * with its scratch -- arrival counters and fp32 partial slots of the split n-tiles -- in a file-scope
 * `static __device__` array of AWQ_B200_STATIC_SCRATCH_BYTES inside the library instead of a caller workspace:
 * the call allocates nothing (no cudaMalloc / cudaFree / allocating library routine), it clears the counters with
 * a cudaMemsetAsync on `stream` and launches.  Every GPU that loads the library holds the array until the process
 * exits.  The scratch a problem needs is 32 KiB + (min(SMs, n_tiles * k_tiles) + n_tiles + 1) * 128 * BN * 4
 * bytes, n_tiles = ceil(N / 128), BN = 16 for M <= 16 else 32 (148 SMs: N <= 242560 at M <= 16, N <= 111744 at
 * M <= 32); larger problems are rejected with an error, use awq_b200_gemv with a caller workspace for them.
 * Every call shares the one array: calls that overlap on the GPU (two streams, parallel graph branches, two host
 * threads on one device) RACE -- serialise them on one stream.  Results are bit-identical to awq_b200_gemv
 * routed to the same kernel. */
#define AWQ_B200_STATIC_SCRATCH_BYTES (16 << 20)
int awq_b200_gemv_static(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K,
                         int64_t N, int dtype, void* stream);

/* prefill path (tcgen05 / TMEM), any M >= 1; intended for M >= 64. */
int awq_b200_gemm(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K,
                  int64_t N, int dtype, void* workspace, int64_t workspace_bytes, void* stream);

/* what AWQLinearMethod.apply calls: picks gemv or gemm from M (threshold is internal). */
int awq_b200_linear(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K,
                    int64_t N, int dtype, void* workspace, int64_t workspace_bytes, void* stream);

/* y[M, N/2] = silu(x @ W_gate) * (x @ W_up) on tiles from awq_b200_repack_gate_up: gate_up_proj and the
 * SiluAndMul that follows it (python/sglang/srt/models/llama.py:94-106; activation.py SiluAndMul) as ONE
 * launch -- the [M, N] intermediate never reaches HBM.  gate and up are rounded to T before the activation
 * (the bits the unfused sequence produces); the activation is evaluated in fp32 and rounded once. */
int awq_b200_linear_silu_mul(const void* x, const void* tiles, void* y, int64_t M, int64_t K, int64_t N,
                             int dtype, void* workspace, int64_t workspace_bytes, void* stream);

/* which kernel awq_b200_linear would pick for M: 0 = gemv, 1 = gemm */
int awq_b200_linear_route(int64_t M, int64_t K, int64_t N);

/* which decode kernel awq_b200_gemv routes (M, K, N) to (reporting / profiling; 0 = invalid shape):
 * 1 = gemv_direct_kernel, 2 = gemv_kernel (legacy stream-K), 3 = gemv_persist_kernel,
 * 4 = gemv_cluster_kernel, 5 = gemv_tc_kernel (tcgen05) */
int awq_b200_gemv_route(int64_t M, int64_t K, int64_t N, int dtype);

/* One-shot all-reduce (sum, fp32 accumulation in rank order: bit-identical on every rank) over
 * peer-mapped staging buffers, for the decode-sized messages that close RowParallelLinear.
 * Replaces tensor_model_parallel_all_reduce -> GroupCoordinator.all_reduce
 * (python/sglang/srt/distributed/communication_op.py:11-13, parallel_state.py:544-623) and its
 * P2P fast path (sgl-kernel/csrc/allreduce/custom_all_reduce.cuh:168-273) for messages that fit
 * one slot; larger messages stay on ncclAllReduce.
 *   peer_buffers_dev  DEVICE array of `world` pointers: base of every rank's staging buffer as
 *                     mapped in THIS process (entry `rank` is the local one); all buffers have
 *                     `buffer_bytes` bytes, identical layout, and were zero-filled once before
 *                     the first call on any rank (the host synchronises the ranks after that)
 *   in / out          local T [numel] (numel % 8 == 0, 16-byte aligned), may alias
 * Every rank must issue the same sequence of calls.  Asynchronous on `stream`, graph-capturable. */
int64_t awq_b200_allreduce_buffer_bytes(int64_t max_message_bytes);
/* diagnostic: 0 = flag-exchange kernel for every size, 1 (default) = one-shot self-validating packets up to
 * 512 KB / 128 KB / 64 KB at 2 / 4 / 8 ranks (measured cross-overs), two-shot packets (reduce-scatter + all-gather) above that up to 1 MB
 * at 4 / 8 ranks, flag exchange for the rest; 2 = one-shot packets whenever they fit (512 KB); 3 = two-shot packets
 * whenever they fit */
void awq_b200_debug_set_allreduce_packets(int mode);
int awq_b200_allreduce(const void* in, void* out, int64_t numel, int dtype, void* const* peer_buffers_dev,
                       int64_t buffer_bytes, int rank, int world, void* stream);

/* RowParallelLinear.forward in one call (python/sglang/srt/layers/linear.py:1388-1414: quant_method.apply, then
 * tensor_model_parallel_all_reduce): y[M, N] = sum over ranks of x_rank[M, K] @ W_rank (+ bias), every rank gets the
 * same bits as awq_b200_linear followed by awq_b200_allreduce.  When the decode router picks a kernel with the
 * packet-push epilogue and the message takes the one-shot packet path, the partial sums leave for the peers from the
 * linear's own epilogue (the NVLink transfer overlaps the rest of the launch) and the collective only polls and sums;
 * otherwise the two run back to back.  Every rank must make the same sequence of calls with the same (M, K, N). */
int awq_b200_linear_allreduce(const void* x, const void* tiles, const void* bias, void* y, int64_t M, int64_t K, int64_t N,
                              int dtype, void* workspace, int64_t workspace_bytes, void* const* peer_buffers_dev,
                              int64_t buffer_bytes, int rank, int world, void* stream);
/* diagnostic: 0 = awq_b200_linear_allreduce always runs the two kernels back to back (default 1) */
void awq_b200_debug_set_linear_allreduce_fusion(int on);

/* ---- fused epilogues either side of the path ------------------ */

/* residual add + RMSNorm, the arithmetic of RMSNorm.forward_native with a residual
 * (python/sglang/srt/layers/layernorm.py:178-222; the CUDA path is sgl_kernel.fused_add_rmsnorm,
 * layernorm.py:122-124):  r = x + residual in fp32;  residual <- T(r);
 * out <- T(r * rsqrt(mean(r^2) + eps) * weight).   x, residual, out: T [rows, hidden] (hidden % 8 == 0,
 * 16-byte aligned; out may alias x); weight T [hidden].  hidden % 1024 == 0 and <= 8192: one thread-block
 * cluster per token row (sum of squares closed through distributed shared memory); else one block per row. */
int awq_b200_residual_rmsnorm(const void* x, void* residual, const void* weight, void* out, int64_t rows,
                              int64_t hidden, float eps, int dtype, void* stream);

/* (residual may be NULL: plain RMSNorm of x, reference sgl_kernel.rmsnorm, layernorm.py:125) */

/* out[r, j] = silu(in[r, j]) * in[r, d + j] for in T [rows, 2 d] (d % 8 == 0), fp32 evaluation, one rounding: the
 * reference op sgl_kernel.silu_and_mul (sgl-kernel/csrc/common_extension.cc:78-79) for the sequence that does not use
 * the fused awq_b200_linear_silu_mul */
int awq_b200_silu_and_mul(const void* in, void* out, int64_t rows, int64_t d, int dtype, void* stream);

/* all-reduce (as awq_b200_allreduce: fp32 rank-order sum rounded to T) + residual add + RMSNorm in ONE
 * kernel when the message takes the packet path and hidden fits a cluster, else the two kernels back to back
 * (identical bits).  Replaces flashinfer_allreduce_residual_rmsnorm
 * (python/sglang/srt/layers/flashinfer_comm_fusion.py:126-200), reached through
 * RMSNorm.forward_with_allreduce_fusion (layernorm.py:255-285) after a row-parallel linear that skipped
 * its own all-reduce (python/sglang/srt/layers/communicator.py).  in: this rank's partial sums. */
int awq_b200_allreduce_residual_rmsnorm(const void* in, void* residual, const void* weight, void* out, int64_t rows,
                                        int64_t hidden, float eps, int dtype, void* const* peer_buffers_dev,
                                        int64_t buffer_bytes, int rank, int world, void* stream);

/* ---- AWQ mixture-of-experts linears, decode batches -------------- */

/* The router's (token, k) pairs against their experts' matrices: replaces the grouped GEMMs of AWQMoEMethod.apply
 * (python/sglang/srt/layers/quantization/awq.py:661-852 -> fused_marlin_moe -> sgl-kernel/csrc/moe/marlin_moe_wna16/).
 *   tiles        num_experts tile arrays of a [K, N] weight, stacked (expert e at e * awq_b200_num_tiles(K, N) records);
 *                silu != 0: made by awq_b200_repack_gate_up (W13), the output is silu(gate) * up, N / 2 wide
 *   expert_ids   int32 [pairs] on the device (pairs <= 1024); ids outside [0, num_experts) zero-fill their row
 *                (expert-parallel masks)
 *   x            T [pairs / x_div, K]: row pair / x_div feeds pair (x_div = top_k for W13, 1 for W2)
 *   row_scale    float [pairs] or NULL: multiplied in fp32 before the single rounding (top-k weights on W2; silu == 0 only)
 *   y            T [pairs, N] (or [pairs, N / 2] with silu)
 * Every selected expert's weights stream once (its pairs are the token columns of the MMA); experts nobody selected
 * are not read.  No workspace, no host-side routing, deterministic, graph-capturable.  Batches of hundreds of tokens:
 * gather the tokens of each expert on the host and use awq_b200_linear(_silu_mul) (tcgen05). */
int awq_b200_moe_linear(const void* x, const void* tiles, const int32_t* expert_ids, const float* row_scale, void* y,
                        int64_t pairs, int64_t x_div, int64_t K, int64_t N, int64_t num_experts, int silu, int dtype,
                        void* stream);
/* out[token, :] = T(sum_k float(o[token * topk + k, :])), hidden % 8 == 0 */
int awq_b200_moe_sum(const void* o, void* out, int64_t tokens, int64_t topk, int64_t hidden, int dtype, void* stream);

/* enable (1, default) / disable (0) programmatic dependent launch on the GEMV/GEMM kernels */
void awq_b200_set_pdl(int enabled);

/* tuning knob: the decode kernels prefetch up to `mb` MiB of their weight array into L2 in the
 * prologue, before the programmatic-dependent-launch wait (default 0 = off: measured slower on
 * B200, profiles/r01_notes.md) */
void awq_b200_set_l2_prefetch_mb(int mb);

/* tuning knob: awq_b200_linear uses the decode kernel for M <= m (0..64, default 64) and the
 * tcgen05 kernel above */
void awq_b200_set_linear_gemv_max_m(int m);

/* tuning knob: cap the decode kernel's grid at n CTAs per SM (0 = as many as fit, default).
 * A cap below the hardware limit leaves room for the NEXT kernel's CTAs to become resident
 * early under programmatic dependent launch. */
void awq_b200_set_gemv_ctas_per_sm(int n);

/* tuning knob: the decode kernel shares activations through shared memory for M >= m
 * (default 1 = always); below m every warp reads its fragments straight from L1/L2. */
void awq_b200_set_gemv_shared_x_min_m(int m);

/* tuning knob: the one-slab-per-CTA decode kernel is considered for M <= m (0..4, default 4;
 * the router itself uses it for M <= 2). */
void awq_b200_set_gemv_direct_max_m(int m);

/* diagnostic knob (tests, profiling): force the decode kernel awq_b200_gemv uses.
 *   impl    0 = route by shape (default)   1 = one-slab-per-CTA ("direct", M <= 4)
 *           2 = legacy stream-K ring kernel (any M <= 64)   3 = persistent team kernel (M <= 16)
 *           4 = cluster / DSMEM split-K kernel (M <= 16)
 *           5 = tcgen05 decode kernel (fp16, M <= 32; `split` = grid cap, 0 = one CTA per SM)
 *           a forced kernel that cannot take the problem falls back to the legacy kernel
 *   split   cluster size (1, 2, 4, 8) for impl 4 / team size (8, 16, 32) for impl 3; 0 = heuristic
 *   variant instantiation (prefetch depth x occupancy) selector; -1 = heuristic */
void awq_b200_debug_set_gemv_impl(int impl, int split, int variant);
/* diagnostic (profiling): dequant kernel variant; bits 0-1 = block width (0 = 128 packed columns,
 * 1 = 32, 2 = 64, 3 = 256), bit 2 = default-cached instead of streaming stores.  Results are identical. */
void awq_b200_debug_set_dequant_variant(int v);
/* debug: force the prefill kernel's token tile (16..256, 0 = planner) and work split (-1 = planner) */
void awq_b200_debug_set_gemm_plan(int bn, int align);
/* development aid: device buffer of >= (16 + 64 * 16 + 4 * 160 * 8) uint64 (zero-filled) that the tcgen05 decode launches
 * fill with %globaltimer stamps of their phases (CTA 0 of every launch; every CTA of launches 4..7)
 * in -DAWQ_TC_TRACE builds of the library (scripts/build_trace.sh); a no-op in the shipped one (scripts/trace_tc.py); NULL (default) switches it off */
void awq_b200_debug_set_tc_trace(void* buf);
/* diagnostic knob (tests, profiling): 0 sends every group-32 / group-64 problem of awq_b200_linear_g to the tcgen05
 * kernel, 1 (default) lets decode batches take the sub-group decode kernels, 2 = 1 with the per-weight variant for
 * every batch it can take */
void awq_b200_debug_set_sub_decode(int on);
/* diagnostic knob (tests, profiling): 0 keeps long-K expert matrices on the whole-K streaming kernel (rounds of 2 / 4
 * pairs), 1 (default) sends them to the chunked kernel (8 pairs per pass at any K) when the average expert has more
 * pairs than one round of the whole-K kernel holds, 2 sends all of them there */
void awq_b200_debug_set_moe_chunk(int on);
/* diagnostic knob: blocks per SM the whole-K MoE kernel's grid is sized for (0 = default rule) */
void awq_b200_debug_set_moe_blocks_per_sm(int n);
/* diagnostic: 1 if forcing `impl` (1..5 above) would really run that kernel on (M, K, N) */
int awq_b200_debug_gemv_impl_fits(int impl, int64_t M, int64_t K, int64_t N, int dtype);

#ifdef __cplusplus
}
#endif
#endif /* AWQ_B200_H_ */
