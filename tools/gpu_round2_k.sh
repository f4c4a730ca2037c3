mkdir -p gpurun_out
timeout -k 5 300 python -u -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "attention_backward or block_backward" > gpurun_out/attn_bwd_tests.log 2>&1; echo "attn bwd tests exit=$?"; tail -n 5 gpurun_out/attn_bwd_tests.log
timeout -k 5 120 python tools/bench_attn_shapes.py > gpurun_out/attn_shapes.txt 2>&1; echo "attn shapes exit=$?"; cat gpurun_out/attn_shapes.txt
timeout -k 5 120 python tools/attn_bwd_phases.py > gpurun_out/attn_bwd_phases.txt 2>&1; head -32 gpurun_out/attn_bwd_phases.txt
