// Test infrastructure (oracle/): registers the REFERENCE's own CUDA dequant kernel -- compiled
// unmodified from /root/reference/sgl-kernel/csrc/gemm/awq_kernel.cu by oracle/build_ref.py --
// under a separate op namespace, ref_sgl_kernel::awq_dequantize, so that the GPU tests and the
// bench can run it beside ours (bit-exact oracle on the device; timed "reference on this GPU"
// baseline = this op + torch.matmul, i.e. AWQLinearMethod.apply of awq.py:434-451).
// Only tests/, __graft_entry__.smoke() and bench.py load the resulting oracle/_ref/ref_awq.so.
#include <torch/all.h>
#include <torch/library.h>

// defined in the reference source (sgl-kernel/include/sgl_kernel_ops.h:201)
torch::Tensor awq_dequantize(torch::Tensor qweight, torch::Tensor scales, torch::Tensor qzeros);

TORCH_LIBRARY(ref_sgl_kernel, m) {
  m.def("awq_dequantize(Tensor qweight, Tensor scales, Tensor qzeros) -> Tensor");
  m.impl("awq_dequantize", c10::kCUDA, &awq_dequantize);
}
