#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -X faulthandler -m pytest tests/test_gpu_more.py -x -q -m gpu -k garbage > gpurun_out/job16.log 2>&1
echo "rc=$?"; tail -30 gpurun_out/job16.log
