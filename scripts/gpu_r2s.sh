#!/bin/bash
# round-2 call S: full parity suite, default bench line, ncu launch list + --set full captures of the dominant kernel
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r2s_smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2s_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2s_pytest.log
tail -4 gpurun_out/r2s_pytest.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r2s_bench70b.json 2> gpurun_out/r2s_bench70b.err; echo "bench70b rc=$?"
head -c 1200 gpurun_out/r2s_bench70b.json; echo; tail -3 gpurun_out/r2s_bench70b.err
for M in 1 4 16; do
  python scripts/ncu_target.py gemv $M 70b > gpurun_out/r2s_plain_$M.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:gemv -s 2 -c 1 -o gpurun_out/r2s_gemv_70b_gate_up_m$M -f python scripts/ncu_target.py gemv $M 70b > gpurun_out/r2s_ncu_$M.log 2>&1
  echo "ncu M=$M rc=$?"
done
timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2s_launches.csv python bench.py --steps 1 --warmup 3 --no-extras > gpurun_out/r2s_ncu_bench.log 2>&1; echo "launch list rc=$?"
wc -l gpurun_out/r2s_launches.csv
