#!/bin/bash
# one GPU lease: parity tests, then whatever measurement the round needs (arguments are run as a shell command afterwards)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv,noheader
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log

if [ -n "$1" ]; then bash -c "$1"; fi
