# round-2 evidence refresh with the final kernels (run under gpurun, 1 GPU): ncu captures + launch list + traces
mkdir -p gpurun_out
bash scripts/gpu_profile_k1.sh > gpurun_out/profile_k1.log 2>&1; tail -2 gpurun_out/profile_k1.log
CMD="python bench.py --workload ppo --steps 6 --warmup 3 --repeats 2 --no-cpu-baseline --no-e2e --others="
$CMD > gpurun_out/plain_ppo.log 2>&1 && timeout 300 ncu --set full --clock-control none --import-source on -k regex:"ppo_loss|scan_" -s 8 -c 4 -o gpurun_out/prof_ppo -f $CMD > gpurun_out/ncu3.log 2>&1; tail -1 gpurun_out/ncu3.log
CMDR="python bench.py --workload replay --steps 8 --warmup 3 --repeats 2 --no-cpu-baseline --no-e2e --others="
$CMDR > gpurun_out/plain_replay.log 2>&1 && timeout 300 ncu --set full --clock-control none -k regex:"fields_copy|row_copy|sample_uniform" -s 6 -c 4 -o gpurun_out/prof_replay -f $CMDR > gpurun_out/ncu4.log 2>&1; tail -1 gpurun_out/ncu4.log
CMDB="python bench.py --steps 4 --warmup 3 --repeats 2 --no-cpu-baseline --no-e2e --others="
$CMDB > gpurun_out/plain_bench.log 2>&1 && timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_bench.csv $CMDB > gpurun_out/ncu5.log 2>&1; grep -c impala_loss gpurun_out/launches_bench.csv
timeout 200 python scripts/k1_trace.py > gpurun_out/k1_trace.txt 2>&1; grep -c . gpurun_out/k1_trace.txt
timeout 300 python scripts/k1_sweep.py "" B200RL_K1_TILE=1 B200RL_K1_TILE=2 B200RL_K1_GENERIC=1 B200RL_K1_PREFETCH=0 B200RL_K1_DEBUG=1 B200RL_K1_DEBUG=2 B200RL_K1_DEBUG=3 B200RL_K1_G=5 B200RL_PDL=0 > gpurun_out/k1_sweep.txt 2>&1; head -3 gpurun_out/k1_sweep.txt
timeout 200 python scripts/parity_errors.py > gpurun_out/parity_errors.txt 2>&1; tail -1 gpurun_out/parity_errors.txt
