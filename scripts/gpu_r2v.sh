#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_decode_kernels.py -x -q 2>&1 | tail -2
MATRIX_MS=4,16 MATRIX_TP=1,8 MATRIX_KERNELS="auto:0:-1,tc:5:0" timeout 900 python scripts/kernel_matrix.py 70b 8b > gpurun_out/r2v_matrix.txt 2>&1; cat gpurun_out/r2v_matrix.txt | cut -c1-120
