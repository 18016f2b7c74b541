#!/bin/bash
mkdir -p gpurun_out
B="timeout 300 python bench.py --steps 4 --warmup 3 --resident-only"
for v in "" spc16 spc24 spc32 spc48 spc64; do
  if [ -z "$v" ]; then L=""; else L="fork-gr-ieee80211_b200/variants/libw80211rx_$v.so"; fi
  W80211_LIB=$L $B 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('variant=$v', round(d['value'],2), d['stage_ms_per_step']['demod'], d['fcs_ok_per_step'])"
done
