#!/bin/bash
mkdir -p gpurun_out
B="timeout 300 python bench.py --steps 4 --warmup 3 --resident-only"
for v in "" abs pad4 abspad4; do
  if [ -z "$v" ]; then L=""; else L="fork-gr-ieee80211_b200/variants/libw80211rx_$v.so"; fi
  W80211_LIB=$L $B 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('variant=$v', round(d['value'],2), d['stage_ms_per_step']['demod'], d['fcs_ok_per_step'])"
  W80211_LIB=$L timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -1
done
