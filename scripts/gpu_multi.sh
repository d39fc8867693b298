# multi-GPU check (run under gpurun --gpus N): bench workloads under torchrun, as the driver launches them
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
for wl in impala_loss ppo learner; do
  timeout 600 $TR bench.py --gpus $N --workload $wl --no-e2e > gpurun_out/bench_${wl}_n$N.log 2>&1
  grep '^{' gpurun_out/bench_${wl}_n$N.log | tail -1 | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('$wl N=$N', '%.4g %s'%(d['value'], d['unit']), 'ms/step %.4f'%d['ms_per_step'], 'frac', d['roofline'] and round(d['roofline']['frac'],3))
" || tail -5 gpurun_out/bench_${wl}_n$N.log
done
timeout 300 python bench.py --workload learner > gpurun_out/bench_learner_n1.log 2>&1; tail -1 gpurun_out/bench_learner_n1.log | cut -c1-400
