#!/bin/bash
# round-2 call U (N GPUs): all-reduce parity + latency, then the TP bench
N=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 scripts/ar_check.py > gpurun_out/r2_ar_tp$N.json 2> gpurun_out/r2_ar_tp$N.err; echo "ar_check rc=$?"
tail -c 3000 gpurun_out/r2_ar_tp$N.json; tail -5 gpurun_out/r2_ar_tp$N.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r2_bench_tp$N.json 2> gpurun_out/r2_bench_tp$N.err; echo "bench rc=$?"
head -c 1500 gpurun_out/r2_bench_tp$N.json; tail -3 gpurun_out/r2_bench_tp$N.err
