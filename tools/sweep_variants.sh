# tuning helper: bench each kernel variant built by csrc/build_variant.sh (never used by tests or the default bench)
mkdir -p gpurun_out
for v in base "$@"; do
  lib=$PWD/fork-gr-ieee80211_b200/variants/libw80211rx_$v.so
  [ "$v" = base ] && lib=$PWD/fork-gr-ieee80211_b200/libw80211rx.so
  echo "== variant $v"
  W80211_LIB=$lib python bench.py --steps 6 --warmup 3 --no-cpu 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(round(d['value'],2), round(d['e2e']['value'],2), {k:round(v,2) for k,v in d['stage_ms_per_step'].items()})"
done
