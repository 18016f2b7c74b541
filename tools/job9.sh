#!/bin/bash
# signal stage: new tests first, then the whole GPU suite
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_signal.py -x -q -m gpu > gpurun_out/job9_signal.log 2>&1
echo "signal rc=$?" >> gpurun_out/job9_signal.log
tail -30 gpurun_out/job9_signal.log
timeout 900 python -m pytest tests -q -m gpu > gpurun_out/job9_all.log 2>&1
echo "all rc=$?" >> gpurun_out/job9_all.log
tail -5 gpurun_out/job9_all.log
