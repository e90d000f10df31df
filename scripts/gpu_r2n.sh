#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_decode_kernels.py tests/test_gpu_dequant.py -x -q > gpurun_out/r2n_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2n_pytest.log
MATRIX_KERNELS="direct:1:-1,persist:3:-1,cluster:4:-1,tc:5:0,legacy:2:-1" timeout 1500 python scripts/kernel_matrix.py 8b 70b qwen l2 > gpurun_out/r2n_matrix.txt 2> gpurun_out/r2n_matrix.err; echo "matrix rc=$?"; tail -3 gpurun_out/r2n_matrix.err
cp gpurun_out/kernel_matrix.json gpurun_out/r2n_kernel_matrix.json
tail -5 gpurun_out/r2n_matrix.txt | cut -c1-180
