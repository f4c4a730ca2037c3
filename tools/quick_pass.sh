#!/bin/bash
# short round-end check on one B200 (run under gpurun): gpu tests + smoke + bench + reference arm + the two micro-benches whose
# numbers the docs quote.  tools/final_pass.sh is the long version (ncu launch lists and full captures).
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
timeout -k 5 900 python -u -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.txt 2>&1; echo "gpu tests exit=$?"; tail -n 3 gpurun_out/pytest_gpu.txt
timeout -k 5 300 python tools/bench_train.py --steps 5 --graph > gpurun_out/bench_train.log 2>&1; echo "bench_train exit=$?"; tail -n 2 gpurun_out/bench_train.log | cut -c1-400
timeout -k 5 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit=$?"; tail -n 2 gpurun_out/smoke.log
timeout -k 5 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit=$?"; tail -n 3 gpurun_out/bench.err; cut -c1-300 gpurun_out/bench.json
timeout -k 5 400 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "reference arm exit=$?"; cut -c1-300 gpurun_out/bench_reference.json
timeout -k 5 300 python tools/bench_gemm.py > gpurun_out/bench_gemm.log 2>&1; echo "bench_gemm exit=$?"
timeout -k 5 300 python tools/bench_attn_shapes.py > gpurun_out/attn_shapes.txt 2>&1; echo "attn shapes exit=$?"
timeout -k 5 200 python tools/bench_freq.py > gpurun_out/bench_freq.txt 2>&1; echo "bench_freq exit=$?"; cut -c1-200 gpurun_out/bench_freq.txt | head -6
