#!/bin/bash
mkdir -p gpurun_out
for v in new head new head; do
  if [ $v = head ]; then export AWQ_B200_LIB=$PWD/sglang_awq_b200/lib/libawq_b200_head.so; else unset AWQ_B200_LIB; fi
  echo "== $v"
  MATRIX_MS=16 MATRIX_TP=1,8 MATRIX_KERNELS="tc:5:0" timeout 900 python scripts/kernel_matrix.py 70b 2>&1 | grep -v "^#\|^shape" | cut -c1-75
done
