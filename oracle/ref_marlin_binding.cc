// Test infrastructure (oracle/): registers the REFERENCE's own Marlin W4A16 kernels -- compiled unmodified from
// /root/reference/sgl-kernel/csrc/gemm/marlin/{gptq_marlin.cu, awq_marlin_repack.cu} by oracle/marlin_ref.py --
// under a separate op namespace (ref_marlin::*), so that bench.py can time "baseline C" (what an
// AWQ checkpoint resolves to by default in the reference, awq.py:267-289) beside our kernels on the same GPU.
// Only tests/ and bench.py load the resulting oracle/_ref/ref_marlin.so.
#include <torch/all.h>
#include <torch/library.h>

#include <optional>

// defined in the reference sources (sgl-kernel/include/sgl_kernel_ops.h:264-297)
torch::Tensor gptq_marlin_gemm(torch::Tensor& a, std::optional<torch::Tensor> c_or_none, torch::Tensor& b_q_weight,
                               torch::Tensor& b_scales, std::optional<torch::Tensor> const& global_scale_or_none,
                               std::optional<torch::Tensor> const& b_zeros_or_none,
                               std::optional<torch::Tensor> const& g_idx_or_none,
                               std::optional<torch::Tensor> const& perm_or_none, torch::Tensor& workspace,
                               int64_t const& b_q_type_id, int64_t size_m, int64_t size_n, int64_t size_k, bool is_k_full,
                               bool use_atomic_add, bool use_fp32_reduce, bool is_zp_float);
torch::Tensor awq_marlin_repack(torch::Tensor& b_q_weight, int64_t size_k, int64_t size_n, int64_t num_bits);

static torch::Tensor marlin_gemm(torch::Tensor a, torch::Tensor b_q_weight, torch::Tensor b_scales, torch::Tensor b_zeros,
                                 torch::Tensor g_idx, torch::Tensor perm, torch::Tensor workspace, int64_t b_q_type_id,
                                 int64_t size_m, int64_t size_n, int64_t size_k, bool use_atomic_add, bool use_fp32_reduce) {
  std::optional<torch::Tensor> none;
  std::optional<torch::Tensor> zeros(b_zeros), g(g_idx), p(perm);
  return gptq_marlin_gemm(a, none, b_q_weight, b_scales, none, zeros, g, p, workspace, b_q_type_id, size_m, size_n, size_k,
                          /*is_k_full=*/true, use_atomic_add, use_fp32_reduce, /*is_zp_float=*/false);
}

static torch::Tensor marlin_repack(torch::Tensor b_q_weight, int64_t size_k, int64_t size_n, int64_t num_bits) {
  return awq_marlin_repack(b_q_weight, size_k, size_n, num_bits);
}

TORCH_LIBRARY(ref_marlin, m) {
  m.def("gptq_marlin_gemm(Tensor a, Tensor b_q_weight, Tensor b_scales, Tensor b_zeros, Tensor g_idx, Tensor perm, "
        "Tensor workspace, int b_q_type_id, int size_m, int size_n, int size_k, bool use_atomic_add, bool use_fp32_reduce) "
        "-> Tensor");
  m.impl("gptq_marlin_gemm", c10::kCUDA, &marlin_gemm);
  m.def("awq_marlin_repack(Tensor b_q_weight, int size_k, int size_n, int num_bits) -> Tensor");
  m.impl("awq_marlin_repack", c10::kCUDA, &marlin_repack);
}
