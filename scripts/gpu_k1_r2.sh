mkdir -p gpurun_out
for i in 1 2 3; do timeout 250 python -m pytest tests -m gpu -q -x -o faulthandler_timeout=120 -p no:cacheprovider 2>&1 | tail -1; done
for shape in "4096 20 15" "256 20 15" "8192 20 6" "1366 20 15" "100 64 18"; do
  echo "stress $shape: $(timeout 120 python scripts/k1_stress.py $shape 60 2>&1 | awk '/grads equal/{n++; if ($5=="True") t++} /FAIL/{print} END{print t"/"n" iterations bit-equal"}')"
done
