#!/bin/bash
mkdir -p gpurun_out
MATRIX_KERNELS="base:5:0,nol2:5:32,b1:5:1,b1nol2:5:33,persist:3:-1" MATRIX_TP=1 MATRIX_MS=1,16,32 timeout 900 python scripts/kernel_matrix.py 70b 8b > gpurun_out/r2m_matrix.txt 2> gpurun_out/r2m_matrix.err; echo "matrix rc=$?"; tail -3 gpurun_out/r2m_matrix.err
cat gpurun_out/r2m_matrix.txt | cut -c1-180
