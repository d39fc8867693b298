mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_replay_learner.py -m gpu -q -x -k "sample_gather or replay or ring or gather" > gpurun_out/pytest_gpu.log 2>&1; tail -2 gpurun_out/pytest_gpu.log
for f in 1 0; do
echo "replay fused=$f: $(B200RL_BENCH_REPLAY_FUSED=$f timeout 600 python bench.py --workload replay --steps 100 --warmup 3 --repeats 10 --no-cpu-baseline --no-e2e --others= 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('ms/step %.5f frac %.3f' % (d['ms_per_step'], d['roofline']['frac']))")"
done
