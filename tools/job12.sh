#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu 2>&1 | tail -4
for d in 2 3 4; do
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu --e2e-depth $d > gpurun_out/job12_d$d.json 2> gpurun_out/job12_d$d.err
echo "depth $d rc=$?"; tail -c 300 gpurun_out/job12_d$d.err
python - <<PY
import json
l=json.loads(open('gpurun_out/job12_d$d.json').read().strip().splitlines()[-1])
print("value", l["value"], "e2e sc16", l["e2e"]["value"], l["e2e"]["ms_per_step"], "fc32", l["e2e_fc32"]["value"], "sc8", l["e2e_sc8"]["value"], l["e2e_sc8"]["ms_per_step"])
PY
done
