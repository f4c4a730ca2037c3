mkdir -p gpurun_out
timeout -k 5 600 python -m pytest tests/test_gpu_train.py -x -q -m gpu -k "graphed" > gpurun_out/pytest_graph.log 2>&1; echo "pytest exit=$?"; tail -12 gpurun_out/pytest_graph.log
for C in 1 2 4 8; do
for G in "" "--graph"; do
timeout -k 5 600 python bench.py --steps 3 --warmup 3 --clips-per-gpu $C --no-train --no-secondary --no-cpu-baseline $G > gpurun_out/sweep_c${C}${G}.json 2> gpurun_out/sweep.err; echo "clips=$C $G exit=$?"
python -c "import json; d=json.load(open('gpurun_out/sweep_c${C}${G}.json')); print('clips/gpu=$C $G:', round(d['value'],3), 'clips/s', round(d['ms_per_step'],1), 'ms  e2e', round(d['e2e']['value'],3), 'launches', d['gpu_launches'])"
done; done
