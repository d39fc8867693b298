# compute-sanitizer over the small-shape GPU tests (ONE tool per gpurun call: pass memcheck|racecheck|initcheck|synccheck)
TOOL=${1:-memcheck}
mkdir -p gpurun_out
timeout 1500 compute-sanitizer --tool $TOOL --error-exitcode 99 python -m pytest tests/test_gpu_impala.py tests/test_gpu_ppo_returns.py tests/test_gpu_optim.py -m gpu -x -q \
  -k "not full_size and not 630_352 and not 1_300_001 and not 630352 and not 1300001" > gpurun_out/sanitize_$TOOL.log 2>&1
echo "rc=$?" >> gpurun_out/sanitize_$TOOL.log
tail -15 gpurun_out/sanitize_$TOOL.log
