timeout 600 python -m pytest tests/test_gpu_ppo_returns.py -m gpu -x -q 2>&1 | tail -4
for nw in 8 4 2 1; do B200RL_K2_WARPS=$nw bash scripts/gpu_ab.sh B200RL_K2_SMEM "1" --workload ppo --steps 1500; done
bash scripts/gpu_ab.sh B200RL_K2_SMEM "0" --workload ppo --steps 1500
