#!/bin/bash
# round-2 call T: new epilogue tests first, then the whole suite
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_epilogues.py -x -q > gpurun_out/r2t_epi.log 2>&1; echo "epilogues rc=$?"; tail -15 gpurun_out/r2t_epi.log
timeout 1200 python -m pytest tests -m gpu -q --deselect tests/test_gpu_epilogues.py > gpurun_out/r2t_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/r2t_pytest.log
