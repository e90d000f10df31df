#!/bin/bash
# awq_b200_gemv_static: alone in a fresh process, then after every other GPU test (the round-end order, pytest -x)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q --durations=5 > gpurun_out/r2z_alone.log 2>&1; echo "alone rc=$?"; tail -12 gpurun_out/r2z_alone.log
timeout 1300 python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/r2z_all.log 2>&1; echo "all rc=$?"; tail -14 gpurun_out/r2z_all.log
