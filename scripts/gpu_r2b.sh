#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_decode_kernels.py -x -q > gpurun_out/r2b_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r2b_pytest.log
timeout 1500 python scripts/kernel_matrix.py > gpurun_out/r2b_matrix.txt 2> gpurun_out/r2b_matrix.err; echo "matrix rc=$?"; tail -3 gpurun_out/r2b_matrix.err
head -40 gpurun_out/r2b_matrix.txt
