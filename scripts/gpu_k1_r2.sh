mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_ppo_returns.py -m gpu -q -x -k "ragged_large" -o faulthandler_timeout=90 > gpurun_out/pytest_gpu.log 2>&1; tail -2 gpurun_out/pytest_gpu.log; grep -n "^E " gpurun_out/pytest_gpu.log | head -5
