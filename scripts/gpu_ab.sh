# A/B helper: VAR "v1 v2 ..." [extra bench args]  (run under gpurun)
VAR=$1; VALS=$2; shift 2
for v in $VALS; do
  echo -n "$VAR=$v $@ : "
  env $VAR=$v timeout 120 python bench.py --no-cpu-baseline --no-e2e --steps 8000 --warmup 20 "$@" 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('%.3f us  frac %.3f'%(d['ms_per_step']*1e3, d['roofline']['frac']))
    elif 'rror' in l: print(l.strip()[:200])
"
done
