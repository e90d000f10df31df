#!/bin/bash
for v in new head new head; do
  if [ $v = head ]; then export AWQ_B200_LIB=$PWD/sglang_awq_b200/lib/libawq_b200_head.so; else unset AWQ_B200_LIB; fi
  echo "== $v"
  MATRIX_MS=1,4,8 MATRIX_TP=1,8 MATRIX_KERNELS="persist:3:-1,legacy:2:-1" timeout 900 python scripts/kernel_matrix.py 70b 8b 2>&1 | grep -v "^#\|^shape" | cut -c1-62
done
