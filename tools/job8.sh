mkdir -p gpurun_out
python -m pytest tests -q -m gpu -x 2>&1 | tail -3
for v in base m6 m8; do
  lib=$PWD/fork-gr-ieee80211_b200/variants/libw80211rx_$v.so
  [ "$v" = base ] && lib=$PWD/fork-gr-ieee80211_b200/libw80211rx.so
  W80211_LIB=$lib python bench.py --steps 6 --warmup 3 --no-cpu 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$v', round(d['value'],2), {k:round(v,2) for k,v in d['stage_ms_per_step'].items()})"
done
