#!/bin/bash
# Llama-3-70B-AWQ linear stack, decode batch 1..64, at N GPUs of this box (BASELINE.json configs[4]).
# usage: scripts/scale70b.sh N   -> gpurun_out/bench70b_tp${N}.json
N=${1:-1}
mkdir -p gpurun_out
if [ "$N" = "1" ]; then
  timeout 600 python bench.py --workload llama3-70b --ms 1,2,4,8,16,32,64 --steps 5 --warmup 3 > gpurun_out/bench70b_tp1.json 2> gpurun_out/bench70b_tp1.err
else
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 \
    bench.py --gpus $N --workload llama3-70b --ms 1,2,4,8,16,32,64 --steps 5 --warmup 3 > gpurun_out/bench70b_tp$N.json 2> gpurun_out/bench70b_tp$N.err
fi
echo rc=$?
python - <<PY
import json
l=[x for x in open("gpurun_out/bench70b_tp$N.json") if x.startswith("{")]
if l:
    d=json.loads(l[-1]); print($N, round(d["value"]), round(d["ms_per_step"],3), json.dumps(d["decode_tokens_per_s_linears_only"]))
else:
    print(open("gpurun_out/bench70b_tp$N.err").read()[-1500:])
PY
