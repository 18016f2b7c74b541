#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_detect.py -x -q -m gpu 2>&1 | tail -3
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/job18.json 2> gpurun_out/job18.err
echo "bench rc=$?"; grep detect_batch gpurun_out/job18.err | tail -3
python - <<'PY'
import json
l=json.loads(open('gpurun_out/job18.json').read().strip().splitlines()[-1])
print(l["value"], l["e2e"]["value"], l["e2e_sc8"]["value"], l["ms_per_step"])
print(json.dumps(l["signal_stage"], indent=1))
PY
