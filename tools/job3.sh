set -x
mkdir -p gpurun_out
nproc; free -g | head -2
( time python bench.py ) > gpurun_out/bench_r1_default.json 2> gpurun_out/bench_r1_default.err; tail -3 gpurun_out/bench_r1_default.err; cat gpurun_out/bench_r1_default.json
( time python bench.py --impl reference --steps 2 --warmup 1 ) > gpurun_out/bench_r1_reference.json 2> gpurun_out/bench_r1_reference.err; tail -3 gpurun_out/bench_r1_reference.err; cat gpurun_out/bench_r1_reference.json
nvidia-smi --query-gpu=memory.used,memory.total --format=csv
