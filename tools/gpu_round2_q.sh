mkdir -p gpurun_out
timeout -k 5 1500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu_full.log 2>&1; echo "pytest exit=$?"; tail -5 gpurun_out/pytest_gpu_full.log
timeout -k 5 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit=$?"; tail -n 3 gpurun_out/smoke.log
