#!/bin/bash
mkdir -p gpurun_out
export PROBE_GEMM_BN=64 PROBE_GEMM_ALIGN=1
python scripts/ncu_target.py gemm 4096 o > gpurun_out/r2r_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:gemm_kernel -s 2 -c 1 -o gpurun_out/r2r_gemm_bn64_o -f python scripts/ncu_target.py gemm 4096 o > gpurun_out/r2r_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/r2r_ncu.log
