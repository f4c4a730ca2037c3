mkdir -p gpurun_out
timeout -k 5 600 python -u -m pytest tests/test_gpu_train.py -q -m gpu > gpurun_out/train_tests.log 2>&1; echo "train tests exit=$?"; tail -n 3 gpurun_out/train_tests.log
timeout -k 5 300 python -u -m pytest tests/test_gpu_parity.py -q -m gpu -k "not sample20" > gpurun_out/parity.log 2>&1; echo "parity exit=$?"; tail -n 3 gpurun_out/parity.log
timeout -k 5 400 python tools/bench_train.py --clips 4 --steps 3 > gpurun_out/bench_train.json 2> gpurun_out/bench_train.err; echo "bench train exit=$?"; cat gpurun_out/bench_train.json; tail -n 3 gpurun_out/bench_train.err
CMD="python tools/bench_train.py --clips 4 --steps 1 --warmup 1"
timeout -k 5 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/train_launches.csv $CMD > gpurun_out/ncu_train.log 2>&1
echo "train launch list exit=$?"; python tools/launch_summary.py gpurun_out/train_launches.csv | head -45
