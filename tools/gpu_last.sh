#!/bin/bash
# last check of the round on the final build: gpu tests + smoke + bench + train bench (trimmed quick pass)
mkdir -p gpurun_out
timeout -k 5 300 python -u -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.txt 2>&1; echo "gpu tests exit=$?"; tail -n 3 gpurun_out/pytest_gpu.txt
timeout -k 5 120 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit=$?"; tail -n 1 gpurun_out/smoke.log
timeout -k 5 200 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit=$?"; cut -c1-200 gpurun_out/bench.json
timeout -k 5 120 python tools/bench_train.py --steps 5 --graph > gpurun_out/bench_train.log 2>&1; echo "bench_train exit=$?"; tail -n 1 gpurun_out/bench_train.log | cut -c1-200
