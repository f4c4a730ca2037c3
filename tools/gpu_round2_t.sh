mkdir -p gpurun_out
timeout -k 5 1500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu_full.log 2>&1; echo "pytest exit=$?"; tail -6 gpurun_out/pytest_gpu_full.log
timeout -k 5 300 python tools/bench_gemm.py > gpurun_out/bench_gemm.txt 2>&1; echo "bench_gemm exit=$?"; tail -30 gpurun_out/bench_gemm.txt
timeout -k 5 300 python tools/bench_train.py --steps 5 > gpurun_out/bench_train.json 2> gpurun_out/bench_train.err; echo "train exit=$?"; cut -c1-500 gpurun_out/bench_train.json
timeout -k 5 600 python bench.py --steps 3 --warmup 3 --no-train > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit=$?"; cut -c1-1600 gpurun_out/bench.json
