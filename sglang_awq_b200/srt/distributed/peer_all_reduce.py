"""One-shot all-reduce over NVLink peer memory for decode-sized messages.

Host side of `awq_b200_allreduce` (sglang_awq_b200/csrc/awq_allreduce.cu).  Mirrors the role of the
reference's `CustomAllreduce` (python/sglang/srt/distributed/device_communicators/custom_all_reduce.py:
44-50 size gate, :237-249 buffer registration, :315-336 `custom_all_reduce`) inside
`GroupCoordinator.all_reduce` (parallel_state.py:583-623): messages that fit the staging slot go
through the P2P kernel, larger ones through NCCL.

Plumbing only: the staging buffer is a PyTorch symmetric-memory allocation, whose rendezvous maps
every rank's buffer into every process and yields the device array of base pointers the kernel
takes.  No compute happens here.
"""
from __future__ import annotations

import logging
import os
from typing import Optional

import torch
import torch.distributed as dist

logger = logging.getLogger(__name__)

MAX_MESSAGE_BYTES = 1 << 20  # M = 64 tokens x hidden 8192 x fp16 (reference gate: 8 MB, custom_all_reduce.py:44-50)
SUPPORTED_WORLD_SIZES = (2, 4, 8)


class PeerAllReduce:
    """Owns the staging buffer of one rank and the peer mapping of the whole TP group."""

    def __init__(self, group: dist.ProcessGroup, rank_in_group: int, world_size: int, device: torch.device,
                 max_message_bytes: int = MAX_MESSAGE_BYTES):
        import torch.distributed._symmetric_memory as symm_mem

        from sglang_awq_b200 import _C

        self._C = _C
        self.rank = rank_in_group
        self.world_size = world_size
        self.device = device
        self.max_message_bytes = max_message_bytes
        self.buffer_bytes = int(_C.lib().awq_b200_allreduce_buffer_bytes(max_message_bytes))
        self.buffer = symm_mem.empty(self.buffer_bytes, dtype=torch.uint8, device=device)
        self.buffer.zero_()  # epochs and flags start at 0 on every rank ...
        self.handle = symm_mem.rendezvous(self.buffer, group)
        torch.cuda.synchronize(device)
        dist.barrier(group=group)  # ... before any peer may signal
        assert self.handle.world_size == world_size and self.handle.rank == rank_in_group
        self.peer_ptrs_dev = int(self.handle.buffer_ptrs_dev)

    def should_use(self, t: torch.Tensor) -> bool:
        """Gate on properties that are identical on every rank of a collective (dtype, element
        count, size): a per-rank property such as alignment or contiguity would let the ranks of
        one call split between this kernel and NCCL, and both sides would wait forever."""
        return (self.buffer is not None and t.is_cuda and t.dtype in (torch.float16, torch.bfloat16)
                and t.numel() % 8 == 0 and 0 < t.numel() * t.element_size() <= self.max_message_bytes)

    def all_reduce(self, t: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """sum over the group, in place unless `out` is given; on the current stream."""
        out = t if out is None else out
        C = self._C
        src = t
        if not t.is_contiguous() or t.data_ptr() % 16 != 0:
            src = t.contiguous()          # rank-local detail: stage through an aligned temporary
            if src.data_ptr() % 16 != 0:  # a contiguous view at an odd offset
                src = src.clone(memory_format=torch.contiguous_format)
        dst = out
        if not out.is_contiguous() or out.data_ptr() % 16 != 0:
            dst = src if src is not t else torch.empty_like(src)
        with torch.cuda.device(t.device):
            C._check(C.lib().awq_b200_allreduce(src.data_ptr(), dst.data_ptr(), t.numel(), C._dt(t), self.peer_ptrs_dev,
                                                self.buffer_bytes, self.rank, self.world_size, C._stream()))
        if dst is not out:
            out.copy_(dst.view_as(out) if dst.shape != out.shape else dst)
        return out

    def linear_all_reduce(self, x: torch.Tensor, tiles: torch.Tensor, bias: Optional[torch.Tensor], out_features: int) -> torch.Tensor:
        """RowParallelLinear.forward in one call (awq_b200_linear_allreduce): y = sum over ranks of x_rank @ W_rank
        (+ bias on the rank that holds it).  The decode kernels that serve the per-rank shards push their partial
        sums to the peers from their own epilogue; the collective that follows only polls and sums.  Same bits as
        linear -> all_reduce.  x [M, K] contiguous; the caller checked should_use() on the [M, out_features] result."""
        C = self._C
        M, K = x.shape
        y = torch.empty((M, out_features), dtype=x.dtype, device=x.device)
        with torch.cuda.device(x.device):
            ws = C.workspace(x.device, max_m=max(16, M))
            C._check(C.lib().awq_b200_linear_allreduce(
                x.data_ptr(), tiles.data_ptr(), None if bias is None else bias.data_ptr(), y.data_ptr(), M, K, out_features,
                C._dt(x), ws.data_ptr(), ws.numel(), self.peer_ptrs_dev, self.buffer_bytes, self.rank, self.world_size, C._stream()))
        return y

    def all_reduce_residual_rmsnorm(self, x: torch.Tensor, residual: torch.Tensor, weight: torch.Tensor,
                                    eps: float) -> torch.Tensor:
        """x [rows, hidden] partial sums -> x <- RMSNorm(sum + residual) * weight, residual <- sum + residual;
        contiguous, 16-byte aligned (the caller checks), on the current stream."""
        C = self._C
        rows, hidden = x.shape
        with torch.cuda.device(x.device):
            C._check(C.lib().awq_b200_allreduce_residual_rmsnorm(
                x.data_ptr(), residual.data_ptr(), weight.data_ptr(), x.data_ptr(), rows, hidden, float(eps), C._dt(x),
                self.peer_ptrs_dev, self.buffer_bytes, self.rank, self.world_size, C._stream()))
        return x

    def error_flag(self) -> int:
        """non-zero once a call gave up waiting for a peer (~35 s).  The kernel traps in that case,
        so the failure also surfaces as a CUDA error on the next synchronising call."""
        return int(self.buffer[8:12].view(torch.int32).item())

    def close(self) -> None:
        """Drop the peer mapping.  Call after every CUDA graph that captured `all_reduce` has been
        destroyed and the device is idle, and before `destroy_process_group` (the symmetric-memory
        handle keeps the NCCL communicator's peers mapped)."""
        self.handle = None
        self.buffer = None
        self.peer_ptrs_dev = 0


def try_create(group, rank_in_group: int, world_size: int, device: torch.device) -> Optional[PeerAllReduce]:
    """None when the P2P path is unavailable (world size, no peer access, rendezvous refused): the
    caller then stays on NCCL, which is the reference's own fallback (parallel_state.py:612-623)."""
    if os.environ.get("AWQ_B200_DISABLE_PEER_ALLREDUCE", "0") == "1":
        return None
    if world_size not in SUPPORTED_WORLD_SIZES or device.type != "cuda":
        return None
    ok = 1
    ar = None
    try:
        ar = PeerAllReduce(group, rank_in_group, world_size, device)
    except Exception as e:  # noqa: BLE001 -- any failure here means "no peer mapping on this box"
        logger.warning("peer all-reduce unavailable (%s: %s); row-parallel layers use NCCL", type(e).__name__, e)
        ok = 0
    # all ranks must agree, or some would wait in the kernel for peers that went to NCCL
    flag = torch.tensor([ok], dtype=torch.int32, device=device)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
    return ar if int(flag.item()) == 1 else None
