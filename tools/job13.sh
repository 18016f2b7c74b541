#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -X faulthandler -m pytest tests -x -q -m gpu > gpurun_out/job13.log 2>&1
echo "rc=$?"
grep -v "^  File\|^$" gpurun_out/job13.log | head -40
grep -n "File \"/root/repo\|File \"/.*tests/" gpurun_out/job13.log | head -20
true
python - <<PY
import json
l=json.loads(open('gpurun_out/job13_b.json').read().strip().splitlines()[-1])
print("value", l["value"], "e2e sc16", l["e2e"]["value"], l["e2e"]["ms_per_step"], "fc32", l["e2e_fc32"]["value"], "sc8", l["e2e_sc8"]["value"], l["e2e_sc8"]["ms_per_step"])
PY
