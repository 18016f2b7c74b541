#!/bin/bash
mkdir -p gpurun_out
for i in 1 2; do
W80211_DEMOD_AFTER_TAIL=1 W80211_HDR_PRIO=0 W80211_TRACE=1 timeout 300 python bench.py --steps 10 --warmup 3 --resident-only 2> gpurun_out/tl0.err | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('AFTER_TAIL', round(d['value'],2), round(d['ms_per_step'],2))"
done
grep timeline gpurun_out/tl0.err | sed -n 10,12p
