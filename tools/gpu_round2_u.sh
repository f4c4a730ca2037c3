mkdir -p gpurun_out
timeout -k 5 300 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/pytest_no_alloc_alone.log 2>&1; echo "alone exit=$?"; tail -15 gpurun_out/pytest_no_alloc_alone.log
timeout -k 5 1500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu_full.log 2>&1; echo "full exit=$?"; tail -6 gpurun_out/pytest_gpu_full.log
