set -x
mkdir -p gpurun_out
python -m pytest tests -q -m gpu -x 2>&1 | tail -3
python bench.py --no-cpu --steps 10 > gpurun_out/bench_1gpu.json 2> gpurun_out/bench_1gpu.err; tail -2 gpurun_out/bench_1gpu.err; python -c "
import json; d=json.load(open('gpurun_out/bench_1gpu.json')); print(d['value'], d['e2e']['value'], d['ms_per_step'], d['stage_ms_per_step'])"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_2gpu.json 2> gpurun_out/bench_2gpu.err; tail -3 gpurun_out/bench_2gpu.err; python -c "
import json; d=json.loads(open('gpurun_out/bench_2gpu.json').read().strip().splitlines()[-1]); print(d['n_gpus'], d['value'], d['e2e']['value'], d['ms_per_step'])"
