#!/bin/bash
# round-2 call A: parity suite + new bench defaults
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r2a_smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2a_pytest.log
tail -5 gpurun_out/r2a_pytest.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r2a_bench70b.json 2> gpurun_out/r2a_bench70b.err; echo "bench70b rc=$?"
timeout 600 python bench.py --workload llama3-8b --steps 20 --warmup 3 --no-extras > gpurun_out/r2a_bench8b.json 2> gpurun_out/r2a_bench8b.err; echo "bench8b rc=$?"
timeout 600 python -m sglang_awq_b200.bench_one_batch --model qwen2.5-32b --batch-size 1 --input-len 1024 --output-len 16 --result-filename gpurun_out/r2a_bob_qwen_tp1.jsonl > gpurun_out/r2a_bob_qwen_tp1.log 2>&1; echo "bob rc=$?"
tail -3 gpurun_out/r2a_bob_qwen_tp1.log
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2a_bench_ref.json 2> gpurun_out/r2a_bench_ref.err; echo "ref rc=$?"
head -c 1500 gpurun_out/r2a_bench70b.json; echo; tail -5 gpurun_out/r2a_bench70b.err
