#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_decode_kernels.py -x -q -k "tc or tensor" > gpurun_out/r2k_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2k_pytest.log
MATRIX_KERNELS="g3d3:5:0,g3d2:5:1,g2d4:5:2,persist:3:-1,legacy:2:-1" MATRIX_TP=1 MATRIX_MS=1,16,32 timeout 900 python scripts/kernel_matrix.py 70b 8b > gpurun_out/r2k_matrix.txt 2> gpurun_out/r2k_matrix.err; echo "matrix rc=$?"; tail -3 gpurun_out/r2k_matrix.err
cat gpurun_out/r2k_matrix.txt | cut -c1-160
