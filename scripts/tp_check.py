"""Multi-GPU parity check of the TP path (run under torchrun on N GPUs of one node):
every rank loads ITS shard of the same seeded AWQ checkpoint through the layer API, runs
qkv -> o (all-reduce) and gate_up -> down (all-reduce) on real NCCL, and rank 0 compares with the
unsharded layers.  Prints one JSON line."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from layer_util import awq_cfg, build_layers, make_checkpoint  # noqa: E402
from oracle import awq_oracle as oracle  # noqa: E402
from sglang_awq_b200.srt import distributed as D  # noqa: E402


def main():
    world, rank, local = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    D.init_distributed_environment(world_size=world, rank=rank, local_rank=local, backend="nccl")
    D.initialize_model_parallel(world)
    dev = torch.device("cuda", local)
    dims = dict(hidden=1024, heads=8, kv_heads=8, head_dim=128, inter=2048)
    ck = make_checkpoint(oracle, **dims)

    def finish(ls):
        for l in ls.values():
            l.to(dev)
            l.quant_method.process_weights_after_loading(l)
        return ls

    mine = finish(build_layers(ck, awq_cfg(), None, None, **dims))
    g = torch.Generator(device=dev).manual_seed(3)
    res = {}
    for M in (1, 16, 200):
        x = torch.randn(M, 1024, generator=g, device=dev).half()
        with torch.no_grad():
            qkv, _ = mine["qkv"](x)                       # column parallel: [M, (8+8+8)/world * 128]
            attn_in = qkv[:, :mine["qkv"].q_proj_shard_size].contiguous()  # this rank's q heads stand in for attention
            o, _ = mine["o"](attn_in)                      # row parallel + NCCL all-reduce
            gu, _ = mine["gate_up"](o)
            inter_r = 2048 // world
            h = (gu[:, :inter_r] * 0.01).contiguous()
            dn, _ = mine["down"](h)                        # row parallel + NCCL all-reduce
        if rank == 0:
            one = finish(build_layers(ck, awq_cfg(), 0, 1, **dims))
            with torch.no_grad():
                qkv1, _ = one["qkv"](x)
                o1, _ = one["o"](qkv1[:, :1024].contiguous())
                gu1, _ = one["gate_up"](o1)
                dn1, _ = one["down"]((gu1[:, :2048] * 0.01).contiguous())
            fro = lambda a, b: float(((a.float() - b.float()).norm() / b.float().norm()).item())  # noqa: E731
            res[f"M{M}"] = {"o": fro(o, o1), "down": fro(dn, dn1)}
    dist.barrier()
    if rank == 0:
        ok = all(v["o"] < 2e-3 and v["down"] < 3e-3 for v in res.values())
        print(json.dumps({"tp": world, "ok": ok, "rel_fro": res}))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
