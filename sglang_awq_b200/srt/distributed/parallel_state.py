"""Tensor-parallel process group: one process per GPU, torch.distributed underneath.

Mirrors the slice of the reference's `parallel_state` that the AWQ linear path touches
(python/sglang/srt/distributed/parallel_state.py:544-656 `GroupCoordinator.all_reduce`,
:1484-1695 group setup): a single TP group whose `all_reduce` closes RowParallelLinear.  The
reference dispatches between symmetric-memory NCCL, a P2P custom all-reduce, mscclpp, quick
all-reduce and plain NCCL; here decode-sized messages go through the one-shot peer-memory
kernel (peer_all_reduce.py -> csrc/awq_allreduce.cu, NVLink 5 / NVSwitch P2P) and everything
else through NCCL (backend "nccl"; NVLS in-switch reduction when NCCL selects it), both on the
current CUDA stream; gloo on CPU for the host-logic tests.  Out of scope: PP / DP-attention / EP groups,
shm broadcast, message queues.
"""
from __future__ import annotations

import os
from typing import List, Optional

import torch
import torch.distributed as dist


class GroupCoordinator:
    """A process group plus the collectives the linear layers call."""

    def __init__(self, ranks: List[int], rank: int, backend: Optional[str] = None):
        self.ranks = list(ranks)
        self.world_size = len(self.ranks)
        self.rank = rank
        self.rank_in_group = self.ranks.index(rank) if rank in self.ranks else -1
        self.device_group = None
        self.peer_ar = None
        if self.world_size > 1:
            self.device_group = dist.new_group(self.ranks, backend=backend)
            if self.rank_in_group >= 0 and torch.cuda.is_available() and dist.get_backend(self.device_group) == "nccl":
                from .peer_all_reduce import try_create

                self.peer_ar = try_create(self.device_group, self.rank_in_group, self.world_size,
                                          torch.device("cuda", torch.cuda.current_device()))

    # reference: parallel_state.py:544-623 (in-place on the current stream for CUDA tensors)
    def all_reduce(self, input_: torch.Tensor) -> torch.Tensor:
        if self.world_size == 1:
            return input_
        # reference order (parallel_state.py:583-623): P2P custom all-reduce when the message
        # qualifies, else the NCCL all-reduce
        if self.peer_ar is not None and self.peer_ar.should_use(input_):
            return self.peer_ar.all_reduce(input_)
        dist.all_reduce(input_, group=self.device_group)
        return input_

    def all_reduce_residual_rmsnorm(self, input_: torch.Tensor, residual: torch.Tensor, weight: torch.Tensor,
                                    eps: float = 1e-6):
        """sum over the group of `input_` (a row-parallel linear's partial sums, skip_all_reduce=True), then
        residual <- sum + residual, norm <- RMSNorm(residual) * weight.  Returns (norm, residual) with `norm`
        written over `input_` and `residual` updated in place.  ONE kernel for decode-sized messages (reference:
        flashinfer_allreduce_residual_rmsnorm, python/sglang/srt/layers/flashinfer_comm_fusion.py:126-200)."""
        from sglang_awq_b200 import _C

        x2, r2 = input_.view(-1, input_.shape[-1]), residual.view(-1, residual.shape[-1])
        # the gate uses only properties that are equal on every rank (as in all_reduce); rank-local details
        # (a misaligned or strided view) are staged through aligned temporaries instead of changing the path
        if self.world_size > 1 and self.peer_ar is not None and self.peer_ar.should_use(input_):
            def staged(t):
                return t if t.is_contiguous() and t.data_ptr() % 16 == 0 else t.contiguous().clone()
            xs, rs = staged(x2), staged(r2)
            self.peer_ar.all_reduce_residual_rmsnorm(xs, rs, weight, eps)
            if xs is not x2:
                x2.copy_(xs)
            if rs is not r2:
                r2.copy_(rs)
            return input_, residual
        self.all_reduce(input_)
        _C.residual_rmsnorm(x2, r2, weight, eps)
        return input_, residual

    def all_gather(self, input_: torch.Tensor, dim: int = -1) -> torch.Tensor:
        if self.world_size == 1:
            return input_
        if dim < 0:
            dim += input_.dim()
        parts = [torch.empty_like(input_) for _ in range(self.world_size)]
        dist.all_gather(parts, input_.contiguous(), group=self.device_group)
        return torch.cat(parts, dim=dim)

    def barrier(self) -> None:
        if self.world_size > 1:
            dist.barrier(group=self.device_group)


_TP: Optional[GroupCoordinator] = None


def init_distributed_environment(world_size: int = -1, rank: int = -1, distributed_init_method: str = "env://",
                                 local_rank: int = -1, backend: str = "nccl") -> None:
    """reference: parallel_state.py init_distributed_environment.  Rendezvous on 127.0.0.1."""
    if dist.is_initialized():
        return
    if distributed_init_method == "env://":
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
    kwargs = {}
    if backend == "nccl" and local_rank >= 0:
        torch.cuda.set_device(local_rank)
        kwargs["device_id"] = torch.device("cuda", local_rank)
    dist.init_process_group(backend=backend, init_method=distributed_init_method, world_size=world_size, rank=rank,
                            **kwargs)


def initialize_model_parallel(tensor_model_parallel_size: int = 1, backend: Optional[str] = None) -> None:
    """One TP group spanning `tensor_model_parallel_size` consecutive ranks."""
    global _TP
    if not dist.is_initialized():
        assert tensor_model_parallel_size == 1, "init_distributed_environment() first"
        _TP = GroupCoordinator([0], 0)
        return
    world, rank = dist.get_world_size(), dist.get_rank()
    assert world % tensor_model_parallel_size == 0
    mine = None
    for g in range(world // tensor_model_parallel_size):
        ranks = list(range(g * tensor_model_parallel_size, (g + 1) * tensor_model_parallel_size))
        coord = GroupCoordinator(ranks, rank, backend=backend)  # new_group must be called by every rank
        if rank in ranks:
            mine = coord
    _TP = mine


def model_parallel_is_initialized() -> bool:
    return _TP is not None


def get_tp_group() -> GroupCoordinator:
    global _TP
    if _TP is None:  # single-process use: TP=1 without any rendezvous
        _TP = GroupCoordinator([0], 0)
    return _TP


def get_tensor_model_parallel_world_size() -> int:
    return get_tp_group().world_size


def get_tensor_model_parallel_rank() -> int:
    return get_tp_group().rank_in_group


def destroy_model_parallel() -> None:
    """Release the TP group's peer mapping.  Teardown order for a process that captured CUDA
    graphs over the collectives: (1) delete the graphs, (2) synchronize, (3) this function,
    (4) destroy_distributed_environment()."""
    global _TP
    if _TP is not None and _TP.peer_ar is not None:
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        _TP.peer_ar.close()
        _TP.peer_ar = None
    _TP = None


def destroy_distributed_environment() -> None:
    if dist.is_initialized():
        dist.destroy_process_group()
