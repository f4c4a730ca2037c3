#!/bin/bash
# runs the gpu tier in groups, each under its own timeout, so that one hung kernel cannot eat the whole call;
# the first group that times out stops the run (a hang would otherwise repeat in every later group)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
nproc >> gpurun_out/gpu.txt; lscpu | grep -E "Model name|^CPU\(s\)" >> gpurun_out/gpu.txt
run() { name=$1; shift; timeout -k 5 "$T" python -u -m pytest tests/test_gpu_parity.py -q -m gpu -k "$@" > gpurun_out/test_$name.log 2>&1; rc=$?; echo "$name exit=$rc" | tee -a gpurun_out/summary.txt; tail -n 3 gpurun_out/test_$name.log; if [ $rc == 124 ]; then echo "TIMEOUT in $name: stopping"; exit 124; fi; }
: > gpurun_out/summary.txt
T=120 run signal "dwt or freq or pyramid"
T=120 run simt "linear_f32 or rope or patch_embed or unpatchify"
T=120 run attn "attention"
T=120 run gemm "linear_bf16 or qknorm or gated"
T=120 run blocks "blocks or adaln"
T=180 run small "small_model"
T=180 run variant "object_only or temporal_motion_encoder or imgspatial or sample_cross or amdmodel"
T=120 run optim "adamw or clip"
T=180 run backward "transpose or linear_backward or gelu_and_layernorm or attention_backward or headnorm64 or block_backward"
T=300 run default "default_model or natively"
if [ "$1" == "bench" ]; then
  timeout -k 5 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit=$?" | tee -a gpurun_out/summary.txt; tail -n 4 gpurun_out/smoke.log
  timeout -k 5 600 python bench.py --steps 2 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit=$?" | tee -a gpurun_out/summary.txt; tail -n 5 gpurun_out/bench.err; cat gpurun_out/bench.json
fi
