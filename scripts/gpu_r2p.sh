#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_layers.py tests/test_gpu_allreduce.py -x -q > gpurun_out/r2p_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2p_pytest.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r2p_bench2.json 2> gpurun_out/r2p_bench2.err; echo "bench2 rc=$?"; tail -5 gpurun_out/r2p_bench2.err; cut -c1-2500 gpurun_out/r2p_bench2.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 -m sglang_awq_b200.bench_one_batch --model qwen2.5-32b --tp-size 2 --batch-size 1 --input-len 512 --output-len 16 --num-layers 8 > gpurun_out/r2p_b1b.log 2>&1; echo "b1b rc=$?"; tail -12 gpurun_out/r2p_b1b.log
