#!/bin/bash
# round-end measurement pass on one B200 (run under gpurun): gpu tests + smoke + bench, reference arm, micro-benches, ncu launch
# lists of the inference step and of the train step, full ncu captures of the dominant kernels.  Everything lands in gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
nproc >> gpurun_out/gpu.txt; lscpu | grep -E "Model name|^CPU\(s\)" >> gpurun_out/gpu.txt
timeout -k 5 900 python -u -m pytest tests -q -m gpu > gpurun_out/pytest_gpu.txt 2>&1; echo "gpu tests exit=$?"; tail -n 3 gpurun_out/pytest_gpu.txt
timeout -k 5 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit=$?"; tail -n 4 gpurun_out/smoke.log
timeout -k 5 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit=$?"; tail -n 3 gpurun_out/bench.err; cat gpurun_out/bench.json
timeout -k 5 400 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "reference arm exit=$?"; cat gpurun_out/bench_reference.json
timeout -k 5 300 python tools/bench_gemm.py > gpurun_out/bench_gemm.log 2>&1; echo "bench_gemm exit=$?"
timeout -k 5 300 python tools/bench_rows.py > gpurun_out/bench_rows.log 2>&1; echo "bench_rows exit=$?"
timeout -k 5 300 python tools/bench_attn_shapes.py > gpurun_out/attn_shapes.txt 2>&1; echo "attn shapes exit=$?"
timeout -k 5 300 python tools/attn_bwd_phases.py > gpurun_out/attn_bwd_phases.txt 2>&1
timeout -k 5 400 python tools/bench_train.py --clips 4 --steps 3 --profile > gpurun_out/bench_train_profile.json 2>/dev/null; echo "bench train exit=$?"
timeout -k 5 900 python tests/tools/library_bar.py > gpurun_out/library_bar.json 2> gpurun_out/library_bar.err; echo "library bar exit=$?"
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-train --no-secondary --no-graph --ncu-range"
timeout -k 5 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout -k 5 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1
echo "launch list exit=$?"
CMD="python tools/bench_train.py --clips 4 --steps 1 --warmup 1"
timeout -k 5 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/train_launches.csv $CMD > gpurun_out/ncu_train.log 2>&1
echo "train launch list exit=$?"
bash tools/ncu_one.sh gemm_ff1 gemm_tc_kernel
bash tools/ncu_one.sh gemm_out gemm_tc_kernel
bash tools/ncu_one.sh wgrad gemm_tc_kernel
bash tools/ncu_one.sh attn attn_tc2_kernel
bash tools/ncu_one.sh attn_bwd_dit attn_bwd_dq_tc
cp gpurun_out/prof_attn_bwd_dit.ncu-rep gpurun_out/prof_attn_bwd_dq.ncu-rep 2>/dev/null
bash tools/ncu_one.sh attn_bwd_dit attn_bwd_dkv_tc
mv gpurun_out/prof_attn_bwd_dit.ncu-rep gpurun_out/prof_attn_bwd_dkv.ncu-rep 2>/dev/null
bash tools/ncu_one.sh ln ln_modulate
bash tools/ncu_one.sh dwt 'pyramid|dwt2'
bash tools/ncu_one.sh freq freq_split3d_mma
ls -la gpurun_out | head -80
