"""RMSNorm with the residual / all-reduce fusions that sit right after the AWQ row-parallel linears.

Mirror of the slice of python/sglang/srt/layers/layernorm.py the decoder layers use around `o_proj` and
`down_proj` (reference llama.py:253-275): `forward(x, residual)` = fused_add_rmsnorm (layernorm.py:104-126),
`forward_with_allreduce_fusion(x, residual)` (layernorm.py:255-285) = all-reduce + residual add + RMSNorm, which
the reference delegates to flashinfer's trtllm_allreduce_fusion (flashinfer_comm_fusion.py:126-200) and this
package runs as ONE peer-memory kernel (csrc/awq_allreduce.cu).  CUDA only, no fallback.
"""
from __future__ import annotations

from typing import Optional, Tuple, Union

import torch

from ..distributed import get_tp_group


class RMSNorm(torch.nn.Module):
    def __init__(self, hidden_size: int, eps: float = 1e-6, params_dtype: Optional[torch.dtype] = None) -> None:
        super().__init__()
        self.weight = torch.nn.Parameter(torch.ones(hidden_size, dtype=params_dtype or torch.get_default_dtype()),
                                         requires_grad=False)
        self.variance_epsilon = eps
        self.hidden_size = hidden_size

    def forward_native(self, x: torch.Tensor, residual: Optional[torch.Tensor] = None):
        """the arithmetic (reference layernorm.py:178-222); used by tests as the torch restatement"""
        orig_dtype = x.dtype
        x = x.to(torch.float32)
        if residual is not None:
            x = x + residual.to(torch.float32)
            residual = x.to(orig_dtype)
        variance = x.pow(2).mean(dim=-1, keepdim=True)
        x = (x * torch.rsqrt(variance + self.variance_epsilon) * self.weight).to(orig_dtype)
        return x if residual is None else (x, residual)

    def forward(self, x: torch.Tensor, residual: Optional[torch.Tensor] = None
                ) -> Union[torch.Tensor, Tuple[torch.Tensor, torch.Tensor]]:
        from sglang_awq_b200 import sgl_kernel

        if residual is not None:  # in place, like the reference: x <- norm, residual <- x + residual
            sgl_kernel.fused_add_rmsnorm(x.view(-1, self.hidden_size), residual.view(-1, self.hidden_size),
                                         self.weight.data, self.variance_epsilon)
            return x, residual
        return sgl_kernel.rmsnorm(x.reshape(-1, self.hidden_size).contiguous(), self.weight.data,
                                  self.variance_epsilon).view(x.shape)

    def forward_with_allreduce_fusion(self, x: torch.Tensor, residual: Optional[torch.Tensor] = None):
        """x = the partial sums of a row-parallel linear called with skip_all_reduce=True"""
        tp = get_tp_group()
        if residual is not None and tp.world_size > 1:
            return tp.all_reduce_residual_rmsnorm(x, residual, self.weight.data, self.variance_epsilon)
        if tp.world_size > 1:
            x = tp.all_reduce(x)
        return self.forward(x, residual)
