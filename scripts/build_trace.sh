#!/bin/bash
# dev build with per-CTA phase timestamps compiled in (never shipped, never loaded by the package)
set -e
cd "$(dirname "$0")/../sglang_awq_b200"
mkdir -p build/trace lib
for f in api awq_dequant awq_repack awq_gemv awq_gemv_direct awq_gemv_cluster awq_gemv_persist awq_gemv_tc awq_gemm_tcgen05 awq_allreduce; do
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC --expt-relaxed-constexpr \
       -DAWQ_TRACE -I ../include -c csrc/$f.cu -o build/trace/$f.o &
done
wait
nvcc -shared -o lib/libawq_b200_trace.so build/trace/*.o -gencode arch=compute_100a,code=sm_100a -lcudart
echo built lib/libawq_b200_trace.so
