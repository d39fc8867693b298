# round-2 evidence (run under gpurun, 1 GPU): CPU arms of every config, K1 phase traces and knob sweep, ncu captures
mkdir -p gpurun_out
lscpu | grep -E "Model name|^CPU\(s\)|Thread|Socket" > gpurun_out/lscpu.txt; cat gpurun_out/lscpu.txt
for wl in impala_loss ppo replay learner; do
  timeout 600 python bench.py --impl reference --workload $wl --steps 3 --warmup 1 2>/dev/null | tail -1 > gpurun_out/ref_$wl.json
  python -c "import json; d=json.load(open('gpurun_out/ref_$wl.json')); print('$wl', '%.4g' % d['value'], d['unit'], 'ms/step %.2f' % d['ms_per_step'], d['cpu_baseline']['sample'])"
done
timeout 300 python scripts/k1_trace.py > gpurun_out/k1_trace.txt 2>&1; grep -c . gpurun_out/k1_trace.txt
timeout 600 python scripts/k1_sweep.py "" B200RL_K1_TILE=1 B200RL_K1_TILE=2 B200RL_K1_GENERIC=1 B200RL_K1_PREFETCH=0 B200RL_K1_DEBUG=1 B200RL_K1_DEBUG=2 B200RL_K1_DEBUG=3 B200RL_K1_G=5 B200RL_PDL=0 > gpurun_out/k1_sweep.txt 2>&1; cat gpurun_out/k1_sweep.txt
bash scripts/gpu_profile_k1.sh > gpurun_out/profile_k1.log 2>&1; tail -4 gpurun_out/profile_k1.log
CMD="python bench.py --workload ppo --steps 6 --warmup 3 --repeats 2 --no-cpu-baseline --no-e2e --others="
$CMD > gpurun_out/plain_ppo.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"ppo_loss|scan_" -s 8 -c 4 -o gpurun_out/prof_ppo $CMD > gpurun_out/ncu3.log 2>&1; tail -2 gpurun_out/ncu3.log
CMDR="python bench.py --workload replay --steps 8 --warmup 3 --repeats 2 --no-cpu-baseline --no-e2e --others="
$CMDR > gpurun_out/plain_replay.log 2>&1 && ncu --set full --clock-control none -k regex:"row_copy|sample_uniform" -s 6 -c 4 -o gpurun_out/prof_replay $CMDR > gpurun_out/ncu4.log 2>&1; tail -2 gpurun_out/ncu4.log
CMDB="python bench.py --steps 4 --warmup 3 --repeats 2 --no-cpu-baseline --no-e2e --others="
$CMDB > gpurun_out/plain_bench.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_bench.csv $CMDB > gpurun_out/ncu5.log 2>&1; grep -c impala_loss gpurun_out/launches_bench.csv
