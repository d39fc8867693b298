mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_impala.py tests/test_gpu_ppo_returns.py tests/test_gpu_reference_golden.py -q -x > gpurun_out/pytest_gpu.log 2>&1; tail -8 gpurun_out/pytest_gpu.log
timeout 600 python scripts/k1_sweep.py "" B200RL_K1_DEBUG=1 B200RL_K1_DEBUG=2 B200RL_K1_DEBUG=3 B200RL_K1_TILE=1 > gpurun_out/k1_sweep.txt 2>&1; cat gpurun_out/k1_sweep.txt
