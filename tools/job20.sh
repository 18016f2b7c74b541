#!/bin/bash
# validation of the restored tree + source-level ncu capture of demod / header / post at full packet length
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
timeout 600 python bench.py > gpurun_out/job20_bench.json 2> gpurun_out/job20_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/job20_ref.json 2> gpurun_out/job20_ref.err; echo "ref rc=$?"
CMD="python bench.py --no-cpu --steps 2 --warmup 3 --packets 14208 --e2e-packets 14208"
timeout 600 $CMD > gpurun_out/job20_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"demod|hdr_kernel|post_kernel" -s 6 -c 3 -o gpurun_out/prof_demod_full $CMD > gpurun_out/job20_ncu.log 2>&1
tail -2 gpurun_out/job20_ncu.log
python - <<'PY'
import json
l=json.loads(open('gpurun_out/job20_bench.json').read().strip().splitlines()[-1])
print(l["value"], l["e2e"]["value"], l["e2e_sc8"]["value"], l["ms_per_step"], l["stage_ms_per_step"])
print(open('gpurun_out/job20_ref.json').read()[:600])
PY
