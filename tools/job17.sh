#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -X faulthandler -m pytest tests/test_gpu_detect.py -x -q -m gpu > gpurun_out/job17.log 2>&1
echo "rc=$?"; tail -40 gpurun_out/job17.log
