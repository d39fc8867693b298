mkdir -p gpurun_out
B="python bench.py --no-cpu-baseline --no-e2e --steps 8000 --warmup 20"
for layout in batch_major time_major; do for S in 80 120 160 200 240 320 480 640; do
  echo -n "layout=$layout SAMPLES=$S : "
  B200RL_K1_SAMPLES=$S timeout 120 $B --layout $layout 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('%.3f us  frac %.3f'%(d['ms_per_step']*1e3, d['roofline']['frac']))
    elif 'rror' in l: print(l.strip()[:200])
"
done; done | tee gpurun_out/sweep_tiles.txt
