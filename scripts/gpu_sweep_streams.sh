mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_impala.py -m gpu -x -q 2>&1 | grep -v "^$" | head -60 > gpurun_out/pytest_k1.log; tail -5 gpurun_out/pytest_k1.log
timeout 200 python scripts/k1_trace.py 2>&1 | tee gpurun_out/k1_trace.txt
B="python bench.py --no-cpu-baseline --no-e2e --steps 4000 --warmup 20"
for st in 1 2 3 4 6 8; do for pipe in serial independent; do
  echo -n "streams=$st pipe=$pipe : "
  timeout 120 $B --pipeline $pipe --streams $st 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('%.3f us  frac %.3f'%(d['ms_per_step']*1e3, d['roofline']['frac']))
    elif 'rror' in l: print(l.strip()[:200])
"
done; done | tee gpurun_out/sweep_streams.txt
