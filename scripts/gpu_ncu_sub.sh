#!/bin/bash
# ncu --set full of gemv_sub_kernel on Llama-3-70B gate_up at group 64 / 32, M = 1 (and G = 32, M = 4: the 16-column variant)
mkdir -p gpurun_out
for cfg in 64:1 32:1 32:4; do
  G=${cfg%%:*}; M=${cfg##*:}
  PROBE_GROUP=$G python scripts/ncu_target.py linear_g $M 70b > gpurun_out/r2_plain_sub_$G_$M.log 2>&1 && \
  PROBE_GROUP=$G timeout 300 ncu --set full --clock-control none --import-source on -k regex:gemv_sub -s 2 -c 1 -o gpurun_out/r2_gemv_sub_g${G}_m$M -f python scripts/ncu_target.py linear_g $M 70b > gpurun_out/r2_ncu_sub_${G}_$M.log 2>&1
  echo "ncu G=$G M=$M rc=$?"
  python scripts/summarize_profiles.py gpurun_out/r2_gemv_sub_g${G}_m$M.ncu-rep "Llama-3-70B gate_up_proj (K=8192, N=57344) decode at AWQ group size $G, M=$M: gemv_sub_kernel" gpurun_out/r02_gemv_sub_70b_gate_up_g${G}_m$M.txt > /dev/null 2>&1
  rm -f gpurun_out/r2_gemv_sub_g${G}_m$M.ncu-rep
  grep -E "dram__throughput.avg.pct|gpu__time_duration.sum|smsp__issue_active.avg.pct|dram__bytes_read.sum " gpurun_out/r02_gemv_sub_70b_gate_up_g${G}_m$M.txt | head -6
done
