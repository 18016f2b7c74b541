# Round profile: bench line at the default size, ncu launch list (full size, with DRAM bytes), ncu --set full of the
# three big kernels on a short-packet variant (same occupancy and loops, 10x less memory for ncu to save/restore), and of
# the front-end kernels (presiso, run edges, sync events, L-SIG, CFO copy).
mkdir -p gpurun_out
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err || { tail -20 gpurun_out/bench_default.err; exit 1; }
cat gpurun_out/bench_default.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err
cat gpurun_out/bench_reference.json
CMD_FULL="python bench.py --no-cpu --steps 2 --warmup 3"
$CMD_FULL > gpurun_out/plain_full.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_full.csv $CMD_FULL > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
CMD_SHORT="python bench.py --no-cpu --steps 2 --warmup 3 --mpdu-len 400 --packets 113664"
$CMD_SHORT > gpurun_out/plain_short.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"viterbi|demod|traceback" -s 9 -c 3 -f -o gpurun_out/prof_round $CMD_SHORT > gpurun_out/ncu_short.log 2>&1
tail -2 gpurun_out/ncu_short.log
ncu --set full --clock-control none --import-source on -k regex:"presiso|cfo_copy|lsig|sync_event|run_edges" -c 6 -f -o gpurun_out/prof_front $CMD_SHORT > gpurun_out/ncu_front.log 2>&1
tail -2 gpurun_out/ncu_front.log
CMD_C5="python bench.py --resident-only --config C5 --packets 14208 --steps 2 --warmup 1"
$CMD_C5 > gpurun_out/plain_c5.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:demod -s 2 -c 1 -f -o gpurun_out/prof_demod2 $CMD_C5 > gpurun_out/ncu_c5.log 2>&1
tail -2 gpurun_out/ncu_c5.log
