#!/bin/bash
# launch list of one timed bench step + full captures of the dominant kernels (run under gpurun)
set -x
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --ncu-range"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1
echo "launch list exit=$?"
CMD2="python tools/bench_gemm.py"
REPS=1 $CMD2 > gpurun_out/plain2.log 2>&1 &&
REPS=1 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 8 -c 12 -o gpurun_out/prof_gemm -f $CMD2 > gpurun_out/ncu_gemm.log 2>&1
echo "gemm capture exit=$?"
CMD3="python tools/bench_rows.py"
REPS=1 $CMD3 > gpurun_out/plain3.log 2>&1 &&
REPS=1 ncu --set full --clock-control none --import-source on -k regex:'attn_tc_kernel|attn_small_kernel|ln_modulate|dwt2|idwt2|freq_split' -s 20 -c 16 -o gpurun_out/prof_rows -f $CMD3 > gpurun_out/ncu_rows.log 2>&1
echo "rows capture exit=$?"
ls -la gpurun_out
