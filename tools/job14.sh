#!/bin/bash
mkdir -p gpurun_out
W80211_BENCH_TRACE=1 timeout 900 python bench.py --steps 6 --warmup 3 --no-cpu --e2e-packets 56832 > gpurun_out/job14.json 2> gpurun_out/job14.err
grep -n "trace" gpurun_out/job14.err | awk 'NR>60' | head -120
