mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "alone: $(tail -1 gpurun_out/pytest_noalloc_alone.log)"
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "full: $(tail -1 gpurun_out/pytest_gpu.log)"; grep -c "test_zz_no_alloc" gpurun_out/pytest_gpu.log
timeout 600 python -m pytest tests -m gpu -q -x --co 2>/dev/null | tail -3
