#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_linear.py tests/test_gpu_stack.py -x -q > gpurun_out/r2q_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2q_pytest.log
GEMM_MS=${GEMM_MS:-128,512,2048,4096} timeout 1500 python scripts/gemm_matrix.py ${GEMM_GROUPS:-8b 70b} > gpurun_out/r2q_gemm_matrix.txt 2> gpurun_out/r2q_gemm_matrix.err; echo "rc=$?"; tail -5 gpurun_out/r2q_gemm_matrix.err
cat gpurun_out/r2q_gemm_matrix.txt | cut -c1-220
