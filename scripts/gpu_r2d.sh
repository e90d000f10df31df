#!/bin/bash
mkdir -p gpurun_out
export PROBE_IMPL=5 PROBE_VARIANT=${TCV:-0}
python scripts/ncu_target.py gemv 16 70b > gpurun_out/r2l_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:gemv_tc -s 2 -c 1 -o gpurun_out/r2l_tc_m16_70b -f python scripts/ncu_target.py gemv 16 70b > gpurun_out/r2l_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/r2l_ncu.log
