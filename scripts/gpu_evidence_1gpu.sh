#!/bin/bash
# round-2 final evidence, one GPU: smoke, default bench line, ncu --set full of the routed dominant kernels, launch list
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r2_smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r2_bench70b.json 2> gpurun_out/r2_bench70b.err; echo "bench70b rc=$?"
head -c 900 gpurun_out/r2_bench70b.json; echo
for M in 1 4 16; do
  python scripts/ncu_target.py gemv $M 70b > gpurun_out/r2_plain_$M.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:gemv -s 2 -c 1 -o gpurun_out/r2_gemv_70b_gate_up_m$M -f python scripts/ncu_target.py gemv $M 70b > gpurun_out/r2_ncu_$M.log 2>&1
  echo "ncu M=$M rc=$?"
done
timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_launches.csv python bench.py --steps 1 --warmup 3 --no-extras > gpurun_out/r2_ncu_bench.log 2>&1; echo "launch list rc=$?"
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_ref.json 2> gpurun_out/r2_bench_ref.err; echo "ref rc=$?"; head -c 400 gpurun_out/r2_bench_ref.json
