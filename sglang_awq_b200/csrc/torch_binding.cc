// torch_binding.cc -- thin TORCH_LIBRARY registration over the C ABI (include/awq_b200.h).
//
// Keeps the reference's operator schema unchanged so callers drop in:
//   sgl_kernel::awq_dequantize(Tensor qweight, Tensor scales, Tensor qzeros) -> Tensor
//     (reference: sgl-kernel/csrc/common_extension.cc:126-127, sgl-kernel/include/sgl_kernel_ops.h:201)
// and adds the fused ops AWQLinearMethod uses on B200:
//   sgl_kernel::awq_b200_repack(Tensor qweight, Tensor scales, Tensor qzeros) -> Tensor tiles
//       tiles: int32 [T, 2144] = T records of 8576 bytes (weights | scales | zeros)
//   sgl_kernel::awq_b200_linear(Tensor x, Tensor tiles, Tensor? bias, int out_features) -> Tensor
//   sgl_kernel::awq_b200_dequantize_tiled(Tensor tiles, int in_features, int out_features,
//                                         ScalarType dtype) -> Tensor
//   sgl_kernel::awq_b200_repack_gate_up(Tensor qweight, Tensor scales, Tensor qzeros) -> Tensor tiles
//   sgl_kernel::awq_b200_linear_silu_mul(Tensor x, Tensor tiles, int out_features) -> Tensor [.., out_features / 2]
//       gate_up_proj + SiluAndMul (reference python/sglang/srt/models/llama.py:94-106) in one launch
//   sgl_kernel::fused_add_rmsnorm(Tensor! input, Tensor! residual, Tensor weight, float eps, bool enable_pdl) -> ()
//       the reference's own schema (sgl-kernel/csrc/common_extension.cc:69-70)
// No device code here: tensors are validated (the reference validates nothing,
// awq_kernel.cu:186-221; we follow the TORCH_CHECK convention of awq_marlin_repack.cu:193-223)
// and their pointers forwarded on the current CUDA stream.  CUDA dispatch key only: there is
// no CPU implementation, a CPU tensor raises NotImplementedError from the dispatcher.
#include <ATen/ATen.h>
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <torch/library.h>

#include <map>
#include <mutex>
#include <tuple>
#include <vector>

#include <cuda_runtime_api.h>

#include "awq_b200.h"

namespace {

int dtype_code(const at::Tensor& t, const char* what) {
  if (t.scalar_type() == at::ScalarType::Half) return AWQ_B200_F16;
  if (t.scalar_type() == at::ScalarType::BFloat16) return AWQ_B200_BF16;
  TORCH_CHECK(false, what, ": expected float16 or bfloat16, got ", t.scalar_type());
}

void check_rc(int rc) { TORCH_CHECK(rc == 0, awq_b200_last_error()); }

void* current_stream() { return (void*)at::cuda::getCurrentCUDAStream().stream(); }

// One zero-initialised split-K scratch per (device, stream).  A CUDA graph that captured a launch
// keeps the scratch POINTER (its ticket semaphores must be zero at kernel entry), so a buffer that
// was ever handed out is never freed: growth retires the old tensor into `retired` instead of
// dropping it, and growing while the stream is being captured is an error (the new buffer's
// zero-fill would be captured into the graph and the earlier nodes would keep the old pointer).
// The first allocation already covers every decode shape and prefill up to M = 8192.
at::Tensor& workspace_for(const at::Tensor& like, int64_t bytes) {
  static std::mutex mu;
  static std::map<std::pair<int, void*>, at::Tensor> cache;
  static std::vector<at::Tensor> retired;
  std::lock_guard<std::mutex> lock(mu);
  auto key = std::make_pair((int)like.get_device(), current_stream());
  auto it = cache.find(key);
  if (it == cache.end() || it->second.numel() < bytes) {
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing((cudaStream_t)current_stream(), &cap);
    TORCH_CHECK(cap == cudaStreamCaptureStatusNone,
                "awq_b200_linear: the split-K workspace must grow (", bytes, " bytes) while the stream is being "
                "captured; run this shape once eagerly on the capture stream before capturing");
    const int64_t floor_bytes = std::max(awq_b200_workspace_bytes(64, 65536), awq_b200_workspace_bytes(8192, 65536));
    const int64_t alloc = std::max<int64_t>(bytes, floor_bytes);
    if (it != cache.end()) retired.push_back(it->second);
    cache[key] = at::zeros({alloc}, like.options().dtype(at::kByte));
    it = cache.find(key);
  }
  return it->second;
}

at::Tensor awq_dequantize(at::Tensor qweight, at::Tensor scales, at::Tensor qzeros) {
  TORCH_CHECK(qweight.is_cuda() && scales.is_cuda() && qzeros.is_cuda(), "awq_dequantize: tensors must be on CUDA");
  TORCH_CHECK(qweight.dim() == 2 && scales.dim() == 2 && qzeros.dim() == 2, "awq_dequantize: expected 2-D tensors");
  TORCH_CHECK(qweight.scalar_type() == at::kInt && qzeros.scalar_type() == at::kInt,
              "awq_dequantize: qweight and qzeros must be int32");
  TORCH_CHECK(qweight.is_contiguous() && scales.is_contiguous() && qzeros.is_contiguous(),
              "awq_dequantize: tensors must be contiguous");
  const int64_t K = qweight.size(0), NP = qweight.size(1), N = NP * 8;
  TORCH_CHECK(scales.size(1) == N, "awq_dequantize: scales.size(1) must be 8 * qweight.size(1)");
  TORCH_CHECK(qzeros.size(0) == scales.size(0) && qzeros.size(1) == NP, "awq_dequantize: qzeros shape mismatch");
  TORCH_CHECK(scales.size(0) > 0 && K % scales.size(0) == 0, "awq_dequantize: rows not divisible by groups");
  const int dt = dtype_code(scales, "awq_dequantize(scales)");
  const c10::cuda::OptionalCUDAGuard guard(device_of(qweight));
  at::Tensor out = at::empty({K, N}, scales.options());
  check_rc(awq_b200_dequantize(qweight.data_ptr<int>(), scales.data_ptr(), qzeros.data_ptr<int>(), out.data_ptr(), K,
                               N, scales.size(0), dt, current_stream()));
  return out;
}

at::Tensor awq_repack_any(at::Tensor qweight, at::Tensor scales, at::Tensor qzeros, bool gate_up);
at::Tensor awq_repack(at::Tensor qweight, at::Tensor scales, at::Tensor qzeros) {
  return awq_repack_any(qweight, scales, qzeros, false);
}
at::Tensor awq_repack_gate_up(at::Tensor qweight, at::Tensor scales, at::Tensor qzeros) {
  return awq_repack_any(qweight, scales, qzeros, true);
}

at::Tensor awq_repack_any(at::Tensor qweight, at::Tensor scales, at::Tensor qzeros, bool gate_up) {
  TORCH_CHECK(qweight.is_cuda() && scales.is_cuda() && qzeros.is_cuda(), "awq_b200_repack: tensors must be on CUDA");
  TORCH_CHECK(qweight.dim() == 2 && scales.dim() == 2 && qzeros.dim() == 2, "awq_b200_repack: expected 2-D tensors");
  TORCH_CHECK(qweight.scalar_type() == at::kInt && qzeros.scalar_type() == at::kInt,
              "awq_b200_repack: qweight and qzeros must be int32");
  TORCH_CHECK(qweight.is_contiguous() && scales.is_contiguous() && qzeros.is_contiguous(),
              "awq_b200_repack: tensors must be contiguous");
  const int64_t K = qweight.size(0), NP = qweight.size(1), N = NP * 8;
  // group size inferred like the reference op does (awq_kernel.cu:189): K / scales.size(0); 32, 64 or a multiple of 128
  TORCH_CHECK(K % 128 == 0 && scales.size(0) > 0 && K % scales.size(0) == 0, "awq_b200_repack: K must be a multiple of 128 and of the group count");
  const int64_t G = K / scales.size(0);
  const int64_t rec = awq_b200_tile_bytes(G);
  TORCH_CHECK(rec > 0, "awq_b200_repack: group size ", G, " is not supported (32, 64 or a multiple of 128)");
  TORCH_CHECK(!gate_up || G == 128, "awq_b200_repack_gate_up: group size must be 128");
  TORCH_CHECK(scales.size(1) == N && qzeros.size(0) == scales.size(0) && qzeros.size(1) == NP,
              "awq_b200_repack: shape mismatch");
  const int dt = dtype_code(scales, "awq_b200_repack(scales)");
  const c10::cuda::OptionalCUDAGuard guard(device_of(qweight));
  const int64_t T = awq_b200_num_tiles(K, N);
  at::Tensor tiles = at::empty({T, rec / 4}, qweight.options());
  TORCH_CHECK(!gate_up || N % 16 == 0, "awq_b200_repack_gate_up: out_features must be 2 x a multiple of 8");
  if (gate_up)
    check_rc(awq_b200_repack_gate_up(qweight.data_ptr<int>(), scales.data_ptr(), qzeros.data_ptr<int>(), tiles.data_ptr(), K, N,
                                     dt, current_stream()));
  else
    check_rc(awq_b200_repack_g(qweight.data_ptr<int>(), scales.data_ptr(), qzeros.data_ptr<int>(), tiles.data_ptr(), K, N, G, dt,
                               current_stream()));
  return tiles;
}

// returns the group size the records were made for: 128 (8576-byte records; also every multiple of 128), 64 or 32
int64_t check_tiled(const at::Tensor& tiles, int64_t K, int64_t N, const char* op) {
  TORCH_CHECK(tiles.is_cuda(), op, ": tiles must be on CUDA");
  TORCH_CHECK(tiles.scalar_type() == at::kInt && tiles.is_contiguous(), op, ": tiles must be a contiguous int32 tensor");
  const int64_t T = awq_b200_num_tiles(K, N);
  TORCH_CHECK(T >= 0, op, ": in_features must be a multiple of 128");
  for (int64_t G : {128, 64, 32})
    if (tiles.numel() == T * (awq_b200_tile_bytes(G) / 4)) return G;
  TORCH_CHECK(false, op, ": tiles do not match in_features/out_features");
  return 0;
}

at::Tensor awq_dequantize_tiled(at::Tensor tiles, int64_t K, int64_t N, at::ScalarType dtype) {
  const int64_t G = check_tiled(tiles, K, N, "awq_b200_dequantize_tiled");
  TORCH_CHECK(dtype == at::kHalf || dtype == at::kBFloat16, "awq_b200_dequantize_tiled: dtype must be float16/bfloat16");
  const c10::cuda::OptionalCUDAGuard guard(device_of(tiles));
  at::Tensor out = at::empty({K, N}, tiles.options().dtype(dtype));
  check_rc(awq_b200_dequantize_tiled_g(tiles.data_ptr(), out.data_ptr(), K, N, G,
                                       dtype == at::kHalf ? AWQ_B200_F16 : AWQ_B200_BF16, current_stream()));
  return out;
}

at::Tensor awq_linear(at::Tensor x, at::Tensor tiles, c10::optional<at::Tensor> bias, int64_t N) {
  TORCH_CHECK(x.is_cuda(), "awq_b200_linear: x must be on CUDA");
  TORCH_CHECK(x.dim() >= 1, "awq_b200_linear: x must have at least one dimension");
  const int64_t K = x.size(-1);
  const int64_t G = check_tiled(tiles, K, N, "awq_b200_linear");
  const int dt = dtype_code(x, "awq_b200_linear(x)");
  at::Tensor x2 = x.reshape({-1, K}).contiguous();
  const int64_t M = x2.size(0);
  const void* bias_ptr = nullptr;
  at::Tensor b;
  if (bias.has_value() && bias->defined()) {
    b = bias->contiguous();
    TORCH_CHECK(b.is_cuda() && b.numel() == N && b.scalar_type() == x.scalar_type(),
                "awq_b200_linear: bias must be a CUDA tensor of out_features elements in x's dtype");
    bias_ptr = b.data_ptr();
  }
  const c10::cuda::OptionalCUDAGuard guard(device_of(x));
  std::vector<int64_t> out_shape(x.sizes().begin(), x.sizes().end());
  out_shape.back() = N;
  at::Tensor y = at::empty({M, N}, x.options());
  if (M > 0) {
    const int64_t ws_bytes = awq_b200_workspace_bytes(M, N);
    at::Tensor& ws = workspace_for(x, ws_bytes);
    check_rc(awq_b200_linear_g(x2.data_ptr(), tiles.data_ptr(), bias_ptr, y.data_ptr(), M, K, N, G, dt, ws.data_ptr(),
                               ws.numel(), current_stream()));
  }
  return y.reshape(out_shape);
}

at::Tensor awq_linear_silu_mul(at::Tensor x, at::Tensor tiles, int64_t N) {
  TORCH_CHECK(x.is_cuda(), "awq_b200_linear_silu_mul: x must be on CUDA");
  TORCH_CHECK(x.dim() >= 1, "awq_b200_linear_silu_mul: x must have at least one dimension");
  TORCH_CHECK(N % 16 == 0, "awq_b200_linear_silu_mul: out_features must be 2 x a multiple of 8");
  const int64_t K = x.size(-1);
  TORCH_CHECK(check_tiled(tiles, K, N, "awq_b200_linear_silu_mul") == 128, "awq_b200_linear_silu_mul: group size must be 128");
  const int dt = dtype_code(x, "awq_b200_linear_silu_mul(x)");
  at::Tensor x2 = x.reshape({-1, K}).contiguous();
  const int64_t M = x2.size(0);
  const c10::cuda::OptionalCUDAGuard guard(device_of(x));
  std::vector<int64_t> out_shape(x.sizes().begin(), x.sizes().end());
  out_shape.back() = N / 2;
  at::Tensor y = at::empty({M, N / 2}, x.options());
  if (M > 0) {
    at::Tensor& ws = workspace_for(x, awq_b200_workspace_bytes(M, N));
    check_rc(awq_b200_linear_silu_mul(x2.data_ptr(), tiles.data_ptr(), y.data_ptr(), M, K, N, dt, ws.data_ptr(), ws.numel(),
                                      current_stream()));
  }
  return y.reshape(out_shape);
}

// in place like the reference op: residual <- input + residual, input <- rmsnorm(residual) * weight
void fused_add_rmsnorm(at::Tensor input, at::Tensor residual, at::Tensor weight, double eps, bool /*enable_pdl*/) {
  TORCH_CHECK(input.is_cuda() && residual.is_cuda() && weight.is_cuda(), "fused_add_rmsnorm: tensors must be on CUDA");
  TORCH_CHECK(input.dim() == 2 && residual.sizes() == input.sizes() && weight.dim() == 1 && weight.size(0) == input.size(1),
              "fused_add_rmsnorm: expected input/residual [tokens, hidden] and weight [hidden]");
  TORCH_CHECK(input.is_contiguous() && residual.is_contiguous() && weight.is_contiguous(),
              "fused_add_rmsnorm: tensors must be contiguous");
  TORCH_CHECK(residual.scalar_type() == input.scalar_type() && weight.scalar_type() == input.scalar_type(),
              "fused_add_rmsnorm: dtype mismatch");
  const int dt = dtype_code(input, "fused_add_rmsnorm(input)");
  const c10::cuda::OptionalCUDAGuard guard(device_of(input));
  check_rc(awq_b200_residual_rmsnorm(input.data_ptr(), residual.data_ptr(), weight.data_ptr(), input.data_ptr(), input.size(0),
                                     input.size(1), (float)eps, dt, current_stream()));
}

// tiles int32 [E, T, 2144]; expert_ids int32 [pairs]; row_scale float32 [pairs] or None
at::Tensor awq_moe_linear(at::Tensor x, at::Tensor tiles, at::Tensor expert_ids, c10::optional<at::Tensor> row_scale, int64_t x_div,
                          int64_t N, bool silu) {
  TORCH_CHECK(x.is_cuda() && tiles.is_cuda() && expert_ids.is_cuda(), "awq_b200_moe_linear: tensors must be on CUDA");
  TORCH_CHECK(x.dim() == 2 && x.is_contiguous(), "awq_b200_moe_linear: x must be a contiguous [rows, K] tensor");
  TORCH_CHECK(tiles.dim() == 3 && tiles.scalar_type() == at::kInt && tiles.is_contiguous(),
              "awq_b200_moe_linear: tiles must be a contiguous int32 [experts, T, 2144] tensor");
  TORCH_CHECK(expert_ids.scalar_type() == at::kInt && expert_ids.is_contiguous(), "awq_b200_moe_linear: expert_ids must be int32");
  const int64_t K = x.size(1), E = tiles.size(0), pairs = expert_ids.numel();
  const int64_t T = awq_b200_num_tiles(K, N);
  TORCH_CHECK(T >= 0 && tiles.size(1) * tiles.size(2) == T * (AWQ_B200_TILE_BYTES / 4),
              "awq_b200_moe_linear: tiles do not match in_features / out_features");
  TORCH_CHECK(x_div >= 1 && x.size(0) * x_div == pairs, "awq_b200_moe_linear: x must have pairs / x_div rows");
  const float* rs = nullptr;
  at::Tensor r;
  if (row_scale.has_value() && row_scale->defined()) {
    r = row_scale->contiguous();
    TORCH_CHECK(r.is_cuda() && r.scalar_type() == at::kFloat && r.numel() == pairs, "awq_b200_moe_linear: row_scale must be float32 [pairs]");
    rs = r.data_ptr<float>();
  }
  const int dt = dtype_code(x, "awq_b200_moe_linear(x)");
  const c10::cuda::OptionalCUDAGuard guard(device_of(x));
  at::Tensor y = at::empty({pairs, silu ? N / 2 : N}, x.options());
  check_rc(awq_b200_moe_linear(x.data_ptr(), tiles.data_ptr(), expert_ids.data_ptr<int>(), rs, y.data_ptr(), pairs, x_div, K, N, E,
                               silu ? 1 : 0, dt, current_stream()));
  return y;
}

at::Tensor awq_moe_sum(at::Tensor o, int64_t topk) {
  TORCH_CHECK(o.is_cuda() && o.dim() == 2 && o.is_contiguous() && topk >= 1 && o.size(0) % topk == 0,
              "awq_b200_moe_sum: o must be a contiguous CUDA [tokens * topk, hidden] tensor");
  const int dt = dtype_code(o, "awq_b200_moe_sum(o)");
  const c10::cuda::OptionalCUDAGuard guard(device_of(o));
  at::Tensor out = at::empty({o.size(0) / topk, o.size(1)}, o.options());
  check_rc(awq_b200_moe_sum(o.data_ptr(), out.data_ptr(), o.size(0) / topk, topk, o.size(1), dt, current_stream()));
  return out;
}

// reference schema: silu_and_mul(Tensor! out, Tensor input) -> () (sgl-kernel/csrc/common_extension.cc:78-79)
void silu_and_mul(at::Tensor out, at::Tensor input) {
  TORCH_CHECK(input.is_cuda() && out.is_cuda() && input.is_contiguous() && out.is_contiguous(), "silu_and_mul: contiguous CUDA tensors");
  TORCH_CHECK(input.dim() >= 1 && input.size(-1) % 16 == 0 && out.numel() * 2 == input.numel() && out.scalar_type() == input.scalar_type(),
              "silu_and_mul: input [..., 2 d] with d % 8 == 0, out [..., d] of the same dtype");
  const int dt = dtype_code(input, "silu_and_mul(input)");
  const int64_t d = input.size(-1) / 2;
  const c10::cuda::OptionalCUDAGuard guard(device_of(input));
  check_rc(awq_b200_silu_and_mul(input.data_ptr(), out.data_ptr(), input.numel() / (2 * d), d, dt, current_stream()));
}

// reference schema: rmsnorm(Tensor! output, Tensor input, Tensor weight, float eps, bool enable_pdl) -> ()
void rmsnorm(at::Tensor output, at::Tensor input, at::Tensor weight, double eps, bool /*enable_pdl*/) {
  TORCH_CHECK(input.is_cuda() && output.is_cuda() && weight.is_cuda() && input.is_contiguous() && output.is_contiguous() &&
                  weight.is_contiguous(), "rmsnorm: contiguous CUDA tensors");
  TORCH_CHECK(input.dim() == 2 && output.sizes() == input.sizes() && weight.dim() == 1 && weight.size(0) == input.size(1) &&
                  output.scalar_type() == input.scalar_type() && weight.scalar_type() == input.scalar_type(),
              "rmsnorm: expected input / output [tokens, hidden] and weight [hidden] of one dtype");
  const int dt = dtype_code(input, "rmsnorm(input)");
  const c10::cuda::OptionalCUDAGuard guard(device_of(input));
  check_rc(awq_b200_residual_rmsnorm(input.data_ptr(), nullptr, weight.data_ptr(), output.data_ptr(), input.size(0), input.size(1),
                                     (float)eps, dt, current_stream()));
}

}  // namespace

TORCH_LIBRARY_FRAGMENT(sgl_kernel, m) {
  m.def("awq_dequantize(Tensor qweight, Tensor scales, Tensor qzeros) -> Tensor");
  m.impl("awq_dequantize", c10::kCUDA, &awq_dequantize);

  m.def("awq_b200_repack(Tensor qweight, Tensor scales, Tensor qzeros) -> Tensor");
  m.impl("awq_b200_repack", c10::kCUDA, &awq_repack);

  m.def("awq_b200_dequantize_tiled(Tensor tiles, int in_features, int out_features, ScalarType dtype) -> Tensor");
  m.impl("awq_b200_dequantize_tiled", c10::kCUDA, &awq_dequantize_tiled);

  m.def("awq_b200_linear(Tensor x, Tensor tiles, Tensor? bias, int out_features) -> Tensor");
  m.impl("awq_b200_linear", c10::kCUDA, &awq_linear);

  m.def("awq_b200_repack_gate_up(Tensor qweight, Tensor scales, Tensor qzeros) -> Tensor");
  m.impl("awq_b200_repack_gate_up", c10::kCUDA, &awq_repack_gate_up);

  m.def("awq_b200_linear_silu_mul(Tensor x, Tensor tiles, int out_features) -> Tensor");
  m.impl("awq_b200_linear_silu_mul", c10::kCUDA, &awq_linear_silu_mul);

  m.def("awq_b200_moe_linear(Tensor x, Tensor tiles, Tensor expert_ids, Tensor? row_scale, int x_div, int out_features, "
        "bool silu) -> Tensor");
  m.impl("awq_b200_moe_linear", c10::kCUDA, &awq_moe_linear);
  m.def("awq_b200_moe_sum(Tensor o, int topk) -> Tensor");
  m.impl("awq_b200_moe_sum", c10::kCUDA, &awq_moe_sum);

  m.def("silu_and_mul(Tensor! out, Tensor input) -> ()");
  m.impl("silu_and_mul", c10::kCUDA, &silu_and_mul);
  m.def("rmsnorm(Tensor! output, Tensor input, Tensor weight, float eps, bool enable_pdl) -> ()");
  m.impl("rmsnorm", c10::kCUDA, &rmsnorm);

  m.def("fused_add_rmsnorm(Tensor! input, Tensor! residual, Tensor weight, float eps, bool enable_pdl) -> ()");
  m.impl("fused_add_rmsnorm", c10::kCUDA, &fused_add_rmsnorm);
}
