mkdir -p gpurun_out
run() { v=$1; p=$2
  lib=$PWD/fork-gr-ieee80211_b200/variants/libw80211rx_$v.so
  [ "$v" = base ] && lib=$PWD/fork-gr-ieee80211_b200/libw80211rx.so
  W80211_LIB=$lib python bench.py --steps 6 --warmup 3 --no-cpu --packets $p 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); s=d['stage_ms_per_step']; print('$v', $p, 'value', round(d['value'],2), 'vit ms', round(s['viterbi'],2), 'pkts/ms', round($p/s['viterbi'],1))"
}
run base 75776
run c 75776
run h 66304
run f 56832
run f 113664
run i 47360
run i 94720
run g 37888
run g 75776
