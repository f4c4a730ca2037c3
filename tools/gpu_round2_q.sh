mkdir -p gpurun_out
timeout -k 5 1500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu_full.log 2>&1; echo "pytest exit=$?"; tail -4 gpurun_out/pytest_gpu_full.log
timeout -k 5 300 python tools/bench_gemm.py > gpurun_out/bench_gemm.txt 2>&1; echo "bench_gemm exit=$?"; head -12 gpurun_out/bench_gemm.txt
timeout -k 5 600 python bench.py --steps 3 --warmup 3 --no-secondary --no-cpu-baseline > gpurun_out/bench_epi.json 2>/dev/null; echo "bench exit=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/bench_epi.json")); print(round(d["value"],3), "clips/s", round(d["ms_per_step"],1), "ms roof", d["roofline"]["frac"], d["clocks"]["sm_mhz"], "train", d["train"]["ms_per_step"], d["train"]["ms_per_step_eager"])
PY
