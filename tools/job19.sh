#!/bin/bash
# overlap sweep: persistent capped Viterbi grid on a shared stream vs one stream per slot
mkdir -p gpurun_out
O=gpurun_out/job19.txt; : > $O
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -3
run() { echo "## $*" >> $O; ( "$@" ) 2>> gpurun_out/job19.err | tail -1 >> $O; }
B="timeout 300 python bench.py --steps 8 --warmup 3 --resident-only"
run $B --one-stream --packets 113664 --slots 2
run $B --packets 113664 --slots 2
run $B --packets 75776 --slots 3
run $B --packets 37888 --slots 4
run env W80211_VIT_PRIO=0 $B --packets 75776 --slots 3
run env W80211_LIB=fork-gr-ieee80211_b200/variants/libw80211rx_r128.so $B --packets 75776 --slots 3
run env W80211_LIB=fork-gr-ieee80211_b200/variants/libw80211rx_r128.so $B --packets 113664 --slots 3 --vit-bps 12
run env W80211_LIB=fork-gr-ieee80211_b200/variants/libw80211rx_r128.so $B --packets 37888 --slots 4
run $B --packets 113664 --slots 3 --vit-bps 12
cat $O
