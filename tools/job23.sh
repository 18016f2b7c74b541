#!/bin/bash
# scaling check: the driver's own launch line at N GPUs
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name,clocks.sm,power.draw --format=csv | head -10
nproc; free -g | head -2
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu > gpurun_out/job23_n$N.json 2> gpurun_out/job23_n$N.err
echo "rc=$?"; tail -3 gpurun_out/job23_n$N.err
python - <<PY
import json
l=json.loads(open('gpurun_out/job23_n$N.json').read().strip().splitlines()[-1])
print("N", l["n_gpus"], "value", round(l["value"],2), "per gpu", round(l["value"]/l["n_gpus"],2), "e2e", round(l["e2e"]["value"],2), "sc8", round(l["e2e_sc8"]["value"],2), "ms", round(l["ms_per_step"],2), l["clocks"], l["stage_ms_per_step"])
PY
