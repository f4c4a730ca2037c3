mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
timeout -k 5 600 python -u -m pytest tests/test_gpu_train.py -q -m gpu -s -x > gpurun_out/train_tests.log 2>&1; echo "train tests exit=$?"
tail -n 40 gpurun_out/train_tests.log
timeout -k 5 300 python -u -m pytest tests/test_gpu_parity.py -q -m gpu -k "not sample20" > gpurun_out/parity.log 2>&1; echo "parity exit=$?"; tail -n 5 gpurun_out/parity.log
timeout -k 5 400 python -u -m pytest tests/test_gpu_parity.py -q -m gpu -s -k "sample20" > gpurun_out/sample20.log 2>&1; echo "sample20 exit=$?"; tail -n 25 gpurun_out/sample20.log
timeout -k 5 300 python tools/bench_train.py --config small --clips 2 --steps 3 > gpurun_out/bench_train_small.json 2> gpurun_out/bench_train_small.err; echo "bench small exit=$?"; cat gpurun_out/bench_train_small.json; tail -n 5 gpurun_out/bench_train_small.err
timeout -k 5 400 python tools/bench_train.py --clips 4 --steps 3 > gpurun_out/bench_train.json 2> gpurun_out/bench_train.err; echo "bench exit=$?"; cat gpurun_out/bench_train.json; tail -n 5 gpurun_out/bench_train.err
