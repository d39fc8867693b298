# full GPU check (run under gpurun): every -m gpu test, smoke, the default bench exactly as the driver runs it
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_default.log 2>&1; tail -1 gpurun_out/bench_default.log | python scripts/bench_digest.py
