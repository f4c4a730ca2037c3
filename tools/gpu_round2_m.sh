mkdir -p gpurun_out
timeout -k 5 300 python -u -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "attention_backward or block_backward" > gpurun_out/attn_bwd_tests.log 2>&1; echo "attn bwd tests exit=$?"; tail -n 3 gpurun_out/attn_bwd_tests.log
timeout -k 5 120 python tools/bench_attn_shapes.py > gpurun_out/attn_shapes.txt 2>&1; echo "attn shapes exit=$?"; cat gpurun_out/attn_shapes.txt
timeout -k 5 600 python -u -m pytest tests/test_gpu_train.py -q -m gpu > gpurun_out/train_tests.log 2>&1; echo "train tests exit=$?"; tail -n 4 gpurun_out/train_tests.log
timeout -k 5 400 python tools/bench_train.py --clips 4 --steps 3 > gpurun_out/bench_train.json 2> gpurun_out/bench_train.err; echo "bench train exit=$?"; cat gpurun_out/bench_train.json; tail -n 3 gpurun_out/bench_train.err
