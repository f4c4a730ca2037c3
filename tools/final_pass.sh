#!/bin/bash
# round-end measurement pass on one B200 (run under gpurun): gpu tests + smoke + bench, reference arm, launch list,
# full ncu captures of the dominant kernels, per-kernel micro-benches.  Everything lands in gpurun_out/.
mkdir -p gpurun_out
bash scripts_gpu_tests.sh bench || exit $?
timeout -k 5 400 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "reference arm exit=$?"; cat gpurun_out/bench_reference.json
timeout -k 5 300 python tools/bench_gemm.py > gpurun_out/bench_gemm.log 2>&1; echo "bench_gemm exit=$?"
timeout -k 5 300 python tools/bench_rows.py > gpurun_out/bench_rows.log 2>&1; echo "bench_rows exit=$?"
timeout -k 5 300 python tools/bench_optim.py > gpurun_out/bench_optim.log 2>&1; echo "bench_optim exit=$?"
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --ncu-range"
timeout -k 5 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout -k 5 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1
echo "launch list exit=$?"
bash tools/ncu_one.sh gemm_ff1 gemm_tc_kernel
bash tools/ncu_one.sh gemm_out gemm_tc_kernel
bash tools/ncu_one.sh attn attn_tc_kernel
bash tools/ncu_one.sh ln ln_modulate
bash tools/ncu_one.sh dwt 'pyramid|dwt2'
ls -la gpurun_out | head -80
