#!/bin/bash
# the driver's launch line for N GPUs of one box; appends the JSON line to gpurun_out/scale_N.json
N=${1:-2}
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/scale_$N.json 2> gpurun_out/scale_$N.err
tail -c 1500 gpurun_out/scale_$N.json
