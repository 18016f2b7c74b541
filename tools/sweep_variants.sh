python -m pytest tests -q -m gpu -x 2>&1 | tail -4
for v in a b c d; do
  echo "== variant $v"
  W80211_LIB=$PWD/fork-gr-ieee80211_b200/variants/libw80211rx_$v.so python bench.py --packets 65536 --steps 3 --warmup 2 --no-cpu 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(round(d['value'],2), round(d['e2e']['value'],2), {k:round(v,2) for k,v in d['stage_ms_per_step'].items()})"
done
