#!/bin/bash
# usage: cmp_variants.sh "<configs>" lib...   -- resident bench of each config with the product library and each variant
cfgs=$1; shift
for c in $cfgs; do
  for l in product "$@"; do
    if [ $l = product ]; then o=$(python bench.py --config $c --resident-only --steps 6); else o=$(python bench.py --config $c --resident-only --steps 6 --lib $l); fi
    echo "$c $(basename $l) $(echo "$o" | python -c 'import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print("gbit/s %.2f fcs %d viterbi_ms %.3f demod_ms %.3f tb_ms %.3f total %.3f" % (d["value"], d["fcs_ok_per_step"], d["stage_ms_per_step"]["viterbi"], d["stage_ms_per_step"]["demod"], d["stage_ms_per_step"]["traceback"], d["ms_per_step"]))')"
  done
done
