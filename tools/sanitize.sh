#!/bin/bash
# compute-sanitizer over a parity subset (every kernel of the path runs: header, demod 1 and 2 antennas, both Viterbi
# layouts, traceback, post, signal stage, detection, the no-allocation entry point, the block path with LLR leases).
# Logs -> gpurun_out/sanitizer_{memcheck,racecheck,initcheck}.log; copy them to profiles/ (tools/summarize_profiles.py does).
mkdir -p gpurun_out
SUB_MEM="tests/test_gpu_parity.py tests/test_gpu_mu.py tests/test_gpu_signal.py tests/test_gpu_detect.py tests/test_zz_no_alloc.py tests/test_gr_blocks.py tests/test_gpu_more.py"
KMEM="not full_size and not garbage_sig and not c2_c3_c4 and not plain_c_port and not many_subframes and not zero_length"
for tool in memcheck racecheck; do
  if [ $tool = racecheck ]; then
    SUB="tests/test_gpu_parity.py tests/test_gpu_mu.py tests/test_zz_no_alloc.py tests/test_gpu_more.py"
    K="(siso_all_mcs or mimo_2x2 or mu or no_alloc or ragged) and not c2_c3_c4"
  else
    SUB="$SUB_MEM"; K="$KMEM"
  fi
  echo "== compute-sanitizer --tool $tool: pytest $SUB -k '$K'" > gpurun_out/sanitizer_$tool.log
  timeout 1500 compute-sanitizer --tool $tool --target-processes all --error-exitcode 77 \
      python -m pytest $SUB -q -x -m gpu -k "$K" -p no:cacheprovider >> gpurun_out/sanitizer_$tool.log 2>&1
  echo "== exit code $?" >> gpurun_out/sanitizer_$tool.log
  tail -4 gpurun_out/sanitizer_$tool.log
done
