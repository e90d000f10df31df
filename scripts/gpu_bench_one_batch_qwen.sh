#!/bin/bash
# config 4: Qwen2.5-32B-AWQ linear stack, bench_one_batch protocol (--load-format dummy), TP = 1 / 2 / 4
mkdir -p gpurun_out; rm -f gpurun_out/r2x_bob.jsonl
for tp in 1 2 4; do
  for fuse in "" "--fuse-silu-mul"; do
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $tp --master-addr 127.0.0.1 --master-port 2955$tp \
      -m sglang_awq_b200.bench_one_batch --model qwen2.5-32b --batch-size 1 16 --input-len 1024 --output-len 16 $fuse \
      --result-filename gpurun_out/r2x_bob.jsonl > gpurun_out/r2x_bob_tp$tp.log 2>&1; echo "tp=$tp fuse=[$fuse] rc=$?"
  done
done
cat gpurun_out/r2x_bob.jsonl | cut -c1-400
