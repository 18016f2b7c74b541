#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_signal.py -x -q -m gpu 2>&1 | tail -3
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/job11_bench.json 2> gpurun_out/job11_bench.err
echo "bench rc=$?"
tail -c 600 gpurun_out/job11_bench.err
python - <<'PY'
import json
l=json.loads(open('gpurun_out/job11_bench.json').read().strip().splitlines()[-1])
print(l["value"], l["e2e"]["value"], l["ms_per_step"])
print(json.dumps(l["signal_stage"], indent=1))
PY
