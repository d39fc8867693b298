# round 2: persistent K1 -- parity, sweep of the launch knobs, phase trace (run under gpurun)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_impala.py tests/test_gpu_reference_golden.py -x -q > gpurun_out/pytest_k1.log 2>&1; tail -3 gpurun_out/pytest_k1.log
timeout 600 python scripts/k1_sweep.py "" B200RL_K1_DEBUG=1 B200RL_K1_DEBUG=2 B200RL_K1_DEBUG=3 B200RL_K1_DEBUG=3,B200RL_K1_GPC=1 B200RL_K1_DEBUG=3,B200RL_K1_GPC=5 > gpurun_out/k1_sweep.txt 2>&1; cat gpurun_out/k1_sweep.txt
timeout 300 python scripts/k1_trace.py serial > gpurun_out/k1_trace.txt 2>&1; cat gpurun_out/k1_trace.txt
B200RL_K1_DEBUG=3 timeout 300 python scripts/k1_trace.py serial > gpurun_out/k1_trace_d3.txt 2>&1; cat gpurun_out/k1_trace_d3.txt
