# ncu evidence for the bench command (run under gpurun; one ncu "use" per call)
mkdir -p gpurun_out
CMD="python bench.py --steps 40 --warmup 3 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_k1.csv $CMD > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:impala_loss -s 10 -c 3 -o gpurun_out/prof_k1 $CMD > gpurun_out/ncu2.log 2>&1
CMDP="python bench.py --workload ppo --steps 8 --warmup 3 --no-cpu-baseline --no-e2e"
$CMDP > gpurun_out/plain_ppo.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"ppo_loss|scan_" -s 6 -c 4 -o gpurun_out/prof_ppo $CMDP > gpurun_out/ncu3.log 2>&1
CMDR="python bench.py --workload replay --steps 8 --warmup 3 --no-cpu-baseline --no-e2e"
$CMDR > gpurun_out/plain_replay.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"row_copy|sample_uniform" -s 6 -c 4 -o gpurun_out/prof_replay $CMDR > gpurun_out/ncu4.log 2>&1
tail -n 3 gpurun_out/ncu1.log gpurun_out/ncu2.log gpurun_out/ncu3.log gpurun_out/ncu4.log
