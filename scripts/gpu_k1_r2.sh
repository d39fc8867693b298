mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_replay_learner.py tests/test_gpu_reference_golden.py -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
for f in 1 0; do
echo "replay fused=$f: $(B200RL_BENCH_REPLAY_FUSED=$f timeout 600 python bench.py --workload replay --steps 100 --warmup 3 --repeats 10 --no-cpu-baseline --no-e2e --others= 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('ms/step %.5f frac %.3f' % (d['ms_per_step'], d['roofline']['frac']))")"
done
echo "learner: $(timeout 600 python bench.py --workload learner --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --others= 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('ms/step %.4f' % d['ms_per_step'], d['config']['breakdown_ms'], d['config']['launches_per_step'])")"
