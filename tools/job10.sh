#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gr_blocks.py tests/test_gpu_signal.py -x -q -m gpu > gpurun_out/job10.log 2>&1
echo "rc=$?" >> gpurun_out/job10.log
tail -40 gpurun_out/job10.log
