# K1 tuning sweep: CTA tile size x PDL mode (run under gpurun)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
B="python bench.py --no-cpu-baseline --no-e2e --steps 4000 --warmup 20"
for S in 80 140 160 280 320 560; do
  for mode in "1 independent" "1 serial" "0 serial"; do
    set -- $mode
    echo -n "SAMPLES=$S PDL=$1 pipe=$2 : "
    B200RL_K1_SAMPLES=$S B200RL_PDL=$1 timeout 120 $B --pipeline $2 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('%.3f us  frac %.3f'%(d['ms_per_step']*1e3, d['roofline']['frac']))
    elif 'rror' in l: print(l.strip()[:200])
"
  done
done | tee gpurun_out/sweep_k1.txt
