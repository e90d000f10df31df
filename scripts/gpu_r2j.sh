#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_decode_kernels.py -x -q -k "tc or tensor or tcgen05" > gpurun_out/r2k_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2k_pytest.log
MATRIX_KERNELS="g3d2b2:5:0,g3d2b1:5:1,g3d3b2:5:2,g4d2b1:5:3,persist:3:-1,legacy:2:-1" MATRIX_TP=1 MATRIX_MS=1,16,32 timeout 900 python scripts/kernel_matrix.py 70b 8b > gpurun_out/r2k_matrix.txt 2> gpurun_out/r2k_matrix.err; echo "matrix rc=$?"; tail -3 gpurun_out/r2k_matrix.err
cat gpurun_out/r2k_matrix.txt | cut -c1-160
