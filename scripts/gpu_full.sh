# full GPU check: tests, smoke, all bench workloads (run under gpurun)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
timeout 400 python bench.py > gpurun_out/bench_impala.log 2>&1; tail -1 gpurun_out/bench_impala.log
timeout 300 python bench.py --workload ppo --no-cpu-baseline > gpurun_out/bench_ppo.log 2>&1; tail -1 gpurun_out/bench_ppo.log
timeout 300 python bench.py --workload replay --no-cpu-baseline > gpurun_out/bench_replay.log 2>&1; tail -1 gpurun_out/bench_replay.log
timeout 300 python bench.py --layout time_major --no-cpu-baseline --no-e2e > gpurun_out/bench_tm.log 2>&1; tail -1 gpurun_out/bench_tm.log
timeout 300 python bench.py --streams 1 --pipeline serial --no-cpu-baseline --no-e2e > gpurun_out/bench_serial.log 2>&1; tail -1 gpurun_out/bench_serial.log
