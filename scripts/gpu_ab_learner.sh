# A/B of one environment switch on the learner workload (run under gpurun): VAR "v1 v2 ..."
VAR=$1; VALS=$2
for v in $VALS; do
  echo -n "$VAR=$v: "
  env $VAR=$v timeout 300 python bench.py --workload learner --steps 20 --warmup 5 --no-e2e --no-cpu-baseline --others none 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print(d['ms_per_step'], 'ms/step', d['value'], d['unit'])"
done
