# long random-shape cross-check of the two-implementation paths (run under gpurun): 10x the draws of the committed test
mkdir -p gpurun_out
B200RL_FUZZ_SCALE=${1:-10} timeout 900 python -m pytest tests/test_gpu_fuzz.py -m gpu -q -o faulthandler_timeout=200 -p no:cacheprovider > gpurun_out/fuzz.log 2>&1
echo "rc=$?" >> gpurun_out/fuzz.log; tail -15 gpurun_out/fuzz.log
timeout 300 python -m pytest tests -m gpu -q -x -o faulthandler_timeout=120 -p no:cacheprovider 2>&1 | tail -2
