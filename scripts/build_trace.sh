#!/bin/bash
# development aid: libawq_b200_trace.so = the library with the tcgen05 decode kernel built -DAWQ_TC_TRACE
# (use with AWQ_B200_LIB=sglang_awq_b200/lib/libawq_b200_trace.so python scripts/trace_tc.py ...)
set -e
cd "$(dirname "$0")/.."
python -c "from sglang_awq_b200 import build as b; b.build_c_abi()"
B=sglang_awq_b200/build
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC --expt-relaxed-constexpr -DAWQ_TC_TRACE \
  -I include -c sglang_awq_b200/csrc/awq_gemv_tc.cu -o $B/awq_gemv_tc_trace.o
OBJS=$(ls $B/*.o | grep -v "awq_gemv_tc.o" | grep -v "_trace.o")
nvcc -shared -o sglang_awq_b200/lib/libawq_b200_trace.so $OBJS $B/awq_gemv_tc_trace.o -gencode arch=compute_100a,code=sm_100a -lcudart
echo built sglang_awq_b200/lib/libawq_b200_trace.so
