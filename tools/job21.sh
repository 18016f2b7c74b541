#!/bin/bash
# demod / post changes: parity, then per-kernel times of the product build and the demod variants
mkdir -p gpurun_out
O=gpurun_out/job21.txt; : > $O
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
B="timeout 300 python bench.py --steps 5 --warmup 3 --resident-only"
run() { echo "## $*" >> $O; ( "$@" ) 2>> gpurun_out/job21.err | tail -1 >> $O; }
run $B
for v in d_nobf d_mb7 d_h0 d_h1; do
  run env W80211_LIB=fork-gr-ieee80211_b200/variants/libw80211rx_$v.so $B
  W80211_LIB=fork-gr-ieee80211_b200/variants/libw80211rx_$v.so timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -1
done
python - <<'PY'
import json
for l in open('gpurun_out/job21.txt'):
    if l.startswith('#'): print(l.strip()[-70:]); continue
    try:
        d=json.loads(l); print(round(d["value"],2), d["stage_ms_per_step"])
    except Exception as e: print("ERR", l[:200])
PY
