# Round profile: bench line at the default size, ncu launch list (full size, with DRAM bytes), ncu --set full of the
# three big kernels on a short-packet variant (same occupancy and loops, 10x less memory for ncu to save/restore).
mkdir -p gpurun_out
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err || { tail -20 gpurun_out/bench_default.err; exit 1; }
cat gpurun_out/bench_default.json
CMD_FULL="python bench.py --no-cpu --steps 2 --warmup 3"
$CMD_FULL > gpurun_out/plain_full.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_full.csv $CMD_FULL > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
CMD_SHORT="python bench.py --no-cpu --steps 2 --warmup 3 --mpdu-len 400 --packets 113664"
$CMD_SHORT > gpurun_out/plain_short.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"viterbi|demod|traceback" -s 9 -c 3 -o gpurun_out/prof_round $CMD_SHORT > gpurun_out/ncu_short.log 2>&1
tail -2 gpurun_out/ncu_short.log
