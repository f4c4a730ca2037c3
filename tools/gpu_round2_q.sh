mkdir -p gpurun_out
timeout -k 5 1500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu_full.log 2>&1; echo "pytest exit=$?"; tail -4 gpurun_out/pytest_gpu_full.log
for P in 0 1 0 1; do
HV_PDL=$P timeout -k 5 600 python bench.py --steps 3 --warmup 3 --no-secondary --no-cpu-baseline > gpurun_out/bench_pdl$P.json 2>gpurun_out/bench_pdl$P.err; echo "bench pdl=$P exit=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/bench_pdl$P.json")); print("PDL=$P", round(d["value"],3), "clips/s", round(d["ms_per_step"],1), "ms  clk", d["clocks"]["sm_mhz"], " train", d["train"]["ms_per_step"], "eager", d["train"]["ms_per_step_eager"])
PY
done
