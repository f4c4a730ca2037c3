mkdir -p gpurun_out
timeout -k 5 600 python -m pytest tests/test_gpu_train.py -x -q -m gpu -k "graphed or fused" -s > gpurun_out/pytest_graph.log 2>&1; echo "pytest exit=$?"; tail -25 gpurun_out/pytest_graph.log
timeout -k 5 300 python tools/bench_train.py --steps 5 --graph > gpurun_out/bench_train_graph.json 2> gpurun_out/bench_train_graph.err; echo "graph exit=$?"; cut -c1-500 gpurun_out/bench_train_graph.json; tail -5 gpurun_out/bench_train_graph.err
timeout -k 5 300 python tools/bench_train.py --steps 5 > gpurun_out/bench_train.json 2> gpurun_out/bench_train.err; echo "eager exit=$?"; cut -c1-500 gpurun_out/bench_train.json
