mkdir -p gpurun_out
./tools/divcheck | tee gpurun_out/divcheck.txt
python -m pytest tests -q -m gpu -x 2>&1 | tail -3
python bench.py --no-cpu --steps 8 > gpurun_out/bench_1gpu.json 2> gpurun_out/bench_1gpu.err; tail -2 gpurun_out/bench_1gpu.err; python -c "
import json; d=json.load(open('gpurun_out/bench_1gpu.json')); print(d['value'], d['e2e']['value'], d['ms_per_step'], d['stage_ms_per_step'])"
