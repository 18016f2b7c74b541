mkdir -p gpurun_out
CMD_SHORT="python bench.py --no-cpu --steps 2 --warmup 3 --mpdu-len 400 --packets 75776"
$CMD_SHORT > gpurun_out/plain_short.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"demod" -s 3 -c 1 -o gpurun_out/prof_demod $CMD_SHORT > gpurun_out/ncu_demod.log 2>&1
tail -3 gpurun_out/ncu_demod.log
