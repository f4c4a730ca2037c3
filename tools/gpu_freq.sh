#!/bin/bash
# freq_split3d tensor-core kernel: parity tests + A/B micro-bench + one ncu capture (run under gpurun)
mkdir -p gpurun_out
timeout -k 5 300 python -u -m pytest tests/test_gpu_parity.py -x -q -k "freq or dwt or idwt or pyramid or haar or wavelet" > gpurun_out/pytest_freq.txt 2>&1; echo "freq tests exit=$?"; tail -n 15 gpurun_out/pytest_freq.txt
timeout -k 5 200 python tools/bench_freq.py > gpurun_out/bench_freq.txt 2>&1; echo "bench_freq exit=$?"; cat gpurun_out/bench_freq.txt | cut -c1-300
bash tools/ncu_one.sh freq freq_split3d_mma
