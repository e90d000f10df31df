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
  # the reports are 15-40 MB each with --import-source on (gpurun brings back 64 MB at most): summarise here, drop them
  python scripts/summarize_profiles.py gpurun_out/r2_gemv_70b_gate_up_m$M.ncu-rep "Llama-3-70B gate_up_proj (K=8192, N=57344) decode, M=$M, routed kernel" gpurun_out/r02_gemv_70b_gate_up_m$M.txt > /dev/null 2>&1
  python scripts/make_dram_traffic.py gpurun_out/r2_gemv_70b_gate_up_m$M.ncu-rep:8192:57344:$M:profiles/r02_gemv_70b_gate_up_m$M.txt > /dev/null 2>&1
  rm -f gpurun_out/r2_gemv_70b_gate_up_m$M.ncu-rep
done
cp profiles/dram_traffic.json gpurun_out/dram_traffic.json
timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_launches.csv python bench.py --steps 1 --warmup 3 --no-extras > gpurun_out/r2_ncu_bench.log 2>&1; echo "launch list rc=$?"
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_ref.json 2> gpurun_out/r2_bench_ref.err; echo "ref rc=$?"; head -c 400 gpurun_out/r2_bench_ref.json
