#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out/job24.txt; : > $O
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -2
B="timeout 300 python bench.py --steps 10 --warmup 3 --resident-only"
run() { echo "## $*" >> $O; ( "$@" ) 2>> gpurun_out/job24.err | tail -1 >> $O; }
run $B
run $B
python - <<'PY'
import json
for l in open('gpurun_out/job24.txt'):
    if l.startswith('#'): continue
    try:
        d=json.loads(l); print(round(d["value"],2), round(d["ms_per_step"],2), d["stage_ms_per_step"])
    except Exception as e: print("ERR", l[:200])
PY
