#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_decode_kernels.py tests/test_gpu_dequant.py -x -q > gpurun_out/r2i_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2i_pytest.log
MATRIX_TP=1 MATRIX_MS=1,8,16,32 timeout 900 python scripts/kernel_matrix.py 70b 8b > gpurun_out/r2i_matrix.txt 2> gpurun_out/r2i_matrix.err; echo "matrix rc=$?"; tail -3 gpurun_out/r2i_matrix.err
cat gpurun_out/r2i_matrix.txt | cut -c1-160
