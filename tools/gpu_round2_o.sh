mkdir -p gpurun_out
timeout -k 5 300 python -u -m pytest tests/test_gpu_train.py -q -m gpu -x -k "linear_tn or odd_token" > gpurun_out/tn_tests.log 2>&1; echo "tn tests exit=$?"; tail -n 30 gpurun_out/tn_tests.log
timeout -k 5 600 python -u -m pytest tests/test_gpu_train.py -q -m gpu > gpurun_out/train_tests.log 2>&1; echo "train tests exit=$?"; tail -n 4 gpurun_out/train_tests.log
timeout -k 5 400 python tools/bench_train.py --clips 4 --steps 3 > gpurun_out/bench_train.json 2> gpurun_out/bench_train.err; echo "bench train exit=$?"; cat gpurun_out/bench_train.json; tail -n 3 gpurun_out/bench_train.err
HV_WGRAD_TN=0 timeout -k 5 400 python tools/bench_train.py --clips 4 --steps 3 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('HV_WGRAD_TN=0 ms_per_step', d['ms_per_step'])"
