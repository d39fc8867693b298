mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_replay_learner.py tests/test_gpu_optim.py -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log; grep -n "^E " gpurun_out/pytest_gpu.log | head -20
