#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out/job22.txt; : > $O
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -2
B="timeout 300 python bench.py --steps 5 --warmup 3 --resident-only"
run() { echo "## $*" >> $O; ( "$@" ) 2>> gpurun_out/job22.err | tail -1 >> $O; }
run $B
CMD="python bench.py --no-cpu --steps 2 --warmup 3 --packets 14208 --e2e-packets 14208"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"demod|post_kernel" -s 4 -c 2 -o gpurun_out/prof_demod2 $CMD > gpurun_out/job22_ncu.log 2>&1
tail -2 gpurun_out/job22_ncu.log
python - <<'PY'
import json
for l in open('gpurun_out/job22.txt'):
    if l.startswith('#'): print(l.strip()[-70:]); continue
    try:
        d=json.loads(l); print(round(d["value"],2), d["stage_ms_per_step"])
    except Exception as e: print("ERR", l[:200])
PY
