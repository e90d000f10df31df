"""-m gpu (needs >= 2 GPUs, skipped otherwise): the one-shot peer-memory all-reduce that closes
RowParallelLinear (reference: tensor_model_parallel_all_reduce, communication_op.py:11-13 ->
custom_all_reduce.cuh:168-273) against the fp32 rank-order sum -- bit-exact, identical on every
rank, over call sequences and CUDA-graph replays.  Runs scripts/ar_check.py under torchrun."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_peer_all_reduce_two_ranks():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29541", os.path.join(ROOT, "scripts", "ar_check.py")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert p.returncode == 0, p.stderr[-2000:]
    line = [l for l in p.stdout.splitlines() if l.startswith("{")][-1]
    res = json.loads(line)
    assert res["peer_path"], "symmetric-memory rendezvous failed: the ranks fell back to NCCL"
    assert res["ok"] and not res["bad"] and res["error_flag"] == 0, res


def test_all_reduce_argument_checks(cabi):
    """single GPU: the C ABI rejects bad arguments before touching peer memory"""
    L = cabi.lib()
    assert L.awq_b200_allreduce_buffer_bytes(1 << 20) >= (1 << 21) + 8192
    x = torch.zeros(64, dtype=torch.float16, device="cuda")
    ptrs = torch.zeros(8, dtype=torch.int64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    nbytes = L.awq_b200_allreduce_buffer_bytes(1 << 20)
    assert L.awq_b200_allreduce(x.data_ptr(), x.data_ptr(), 60, 0, ptrs.data_ptr(), nbytes, 0, 2, st) != 0   # numel % 8
    assert L.awq_b200_allreduce(x.data_ptr(), x.data_ptr(), 64, 0, ptrs.data_ptr(), nbytes, 0, 3, st) != 0   # world 3
    assert L.awq_b200_allreduce(x.data_ptr(), x.data_ptr(), 64, 0, ptrs.data_ptr(), nbytes, 2, 2, st) != 0   # rank >= world
    assert L.awq_b200_allreduce(x.data_ptr(), x.data_ptr(), 64, 7, ptrs.data_ptr(), nbytes, 0, 2, st) != 0   # dtype
    assert L.awq_b200_allreduce(x.data_ptr(), x.data_ptr(), 64, 0, ptrs.data_ptr(), 4096, 0, 2, st) != 0     # buffer too small
    assert L.awq_b200_allreduce(x.data_ptr(), x.data_ptr(), 0, 0, ptrs.data_ptr(), nbytes, 0, 2, st) == 0    # empty message
