mkdir -p gpurun_out
timeout 300 python bench.py --workload ppo --no-cpu-baseline > gpurun_out/bench_ppo.log 2>&1
timeout 300 python bench.py --workload replay --no-cpu-baseline > gpurun_out/bench_replay.log 2>&1
timeout 300 python bench.py --layout time_major --no-cpu-baseline > gpurun_out/bench_tm.log 2>&1
CMD="python bench.py --steps 40 --warmup 3 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_k1.csv $CMD > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:impala_loss -s 10 -c 3 -o gpurun_out/prof_k1 $CMD > gpurun_out/ncu2.log 2>&1
tail -2 gpurun_out/bench_ppo.log gpurun_out/bench_replay.log gpurun_out/bench_tm.log; tail -3 gpurun_out/ncu1.log gpurun_out/ncu2.log
