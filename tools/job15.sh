#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -X faulthandler -m pytest tests/test_gpu_mu.py -x -q -m gpu > gpurun_out/job15.log 2>&1
echo "mu rc=$?"; tail -40 gpurun_out/job15.log
timeout 900 python -X faulthandler -m pytest tests -q -m gpu > gpurun_out/job15_all.log 2>&1
echo "all rc=$?"; tail -5 gpurun_out/job15_all.log
