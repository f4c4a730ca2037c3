mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout -k 5 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench n=$N exit=$?"; tail -n 4 gpurun_out/bench_n$N.err | cut -c1-400
python - <<PY
import json
d=json.load(open('gpurun_out/bench_n$N.json'))
print('value', round(d['value'],3), 'ms', round(d['ms_per_step'],1), 'e2e', round(d['e2e']['value'],3), 'launches', d['gpu_launches'], 'roof', d['roofline']['frac'], d['clocks'])
print(json.dumps(d['train']))
print(json.dumps(d['secondary'])[:600])
PY
