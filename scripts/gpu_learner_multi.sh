# learner workload (BASELINE config 5) under torchrun on N GPUs, as the driver launches bench.py (run under gpurun --gpus N)
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus $N --workload learner --no-e2e --no-cpu-baseline > gpurun_out/bench_learner_n$N.log 2>&1
grep '^{' gpurun_out/bench_learner_n$N.log | tail -1 > gpurun_out/bench_learner_n$N.json
python -c "
import json; d=json.load(open('gpurun_out/bench_learner_n$N.json')); print('learner N=$N', d['value'], d['unit'], 'ms/step', d['ms_per_step'], d['config'].get('grad_exchange'))" || tail -5 gpurun_out/bench_learner_n$N.log
