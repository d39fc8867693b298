mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ppo_returns.py tests/test_gpu_reference_golden.py -q -x > gpurun_out/pytest_gpu.log 2>&1; tail -1 gpurun_out/pytest_gpu.log
for B in 8192 4096 2048 1024; do
echo "ppo B=$B: $(timeout 600 python bench.py --workload ppo --B $B --steps 50 --warmup 3 --repeats 10 --no-cpu-baseline --no-e2e --others '' 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('ms/step %.5f frac %.3f' % (d['ms_per_step'], d['roofline']['frac']))")"
done
