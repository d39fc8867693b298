# end-of-round soak (run under gpurun): 10x drawn shapes, the GPU suite twice more, K1 both-kernel stress at five shapes
mkdir -p gpurun_out
B200RL_FUZZ_SCALE=10 timeout 900 python -m pytest tests/test_gpu_fuzz.py -m gpu -q -o faulthandler_timeout=200 -p no:cacheprovider 2>&1 | tail -2
for i in 1 2; do timeout 300 python -m pytest tests -m gpu -q -x -o faulthandler_timeout=120 -p no:cacheprovider 2>&1 | tail -1; done
for shape in "4096 20 15" "256 20 15" "8192 20 6" "1366 20 15" "100 64 18" "30000 20 15"; do
  echo "stress $shape: $(timeout 120 python scripts/k1_stress.py $shape 60 2>&1 | awk '/grads equal/{n++; if ($5=="True") t++} /FAIL/{print} END{print t"/"n" iterations bit-equal"}')"
done
