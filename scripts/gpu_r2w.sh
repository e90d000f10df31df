#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_reference_kernel.py -x -q -k marlin 2>&1 | tail -8
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r2w_bench70b.json 2> gpurun_out/r2w_bench70b.err; echo "bench rc=$?"; tail -3 gpurun_out/r2w_bench70b.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/r2w_bench70b.json"))
print(d["value"], d["roofline"]["frac"], d["decode_tokens_per_s"])
for k, v in d["extras"].get("reference_marlin_same_gpu", {}).items():
    print(k, v)
PY
