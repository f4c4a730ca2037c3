mkdir -p gpurun_out
timeout -k 5 900 python -u -m pytest tests/test_gpu_train.py -q -m gpu -s > gpurun_out/train_tests.log 2>&1; echo "train tests exit=$?"
grep -v "^$" gpurun_out/train_tests.log | tail -n 40
timeout -k 5 300 python -u -m pytest tests/test_gpu_parity.py -q -m gpu -k "not sample20" > gpurun_out/parity.log 2>&1; echo "parity exit=$?"; tail -n 5 gpurun_out/parity.log
timeout -k 5 400 python tools/bench_train.py --clips 4 --steps 3 > gpurun_out/bench_train.json 2> gpurun_out/bench_train.err; echo "bench train exit=$?"; cat gpurun_out/bench_train.json; tail -n 5 gpurun_out/bench_train.err
