mkdir -p gpurun_out
python -m pytest tests -q -m gpu -x -k "configs_c2 or sc16 or drop_gates or mimo_flat or fixtures or decode_block" > gpurun_out/plain_san.log 2>&1 && \
compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests -q -m gpu -x -k "configs_c2 or sc16 or drop_gates or mimo_flat or fixtures or decode_block" > gpurun_out/memcheck.log 2>&1
echo "memcheck rc=$?"; tail -15 gpurun_out/memcheck.log
