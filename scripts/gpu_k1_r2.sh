mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_impala.py tests/test_zz_no_alloc.py -m gpu -q -x -o faulthandler_timeout=90 > gpurun_out/pytest_gpu.log 2>&1; tail -2 gpurun_out/pytest_gpu.log
