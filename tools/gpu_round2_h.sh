mkdir -p gpurun_out
timeout -k 5 300 python -u -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "attention_backward or block_backward" > gpurun_out/attn_bwd_tests.log 2>&1; echo "attn bwd tests exit=$?"; tail -n 15 gpurun_out/attn_bwd_tests.log
timeout -k 5 120 python tools/bench_attn_shapes.py > gpurun_out/attn_shapes.txt 2>&1; echo "attn shapes exit=$?"; cat gpurun_out/attn_shapes.txt
