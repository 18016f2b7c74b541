set -x
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,power.limit --format=csv
mkdir -p gpurun_out; ./tools/pipebench > gpurun_out/pipebench.txt 2>&1; cat gpurun_out/pipebench.txt
python -m pytest tests -q -m gpu -x 2>&1 | tail -4
python bench.py --packets 65536 --steps 5 --warmup 3 --no-cpu > gpurun_out/bench_65536.json 2> gpurun_out/bench_65536.err; tail -c 3000 gpurun_out/bench_65536.json; tail -5 gpurun_out/bench_65536.err
