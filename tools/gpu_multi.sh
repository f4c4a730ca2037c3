mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
echo "GPUs: $N"
timeout -k 5 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/ddp_check.py > gpurun_out/ddp_check_n$N.json 2> gpurun_out/ddp_check_n$N.err; echo "ddp check exit=$?"; cat gpurun_out/ddp_check_n$N.json
timeout -k 5 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 tools/ddp_graph_check.py > gpurun_out/ddp_graph_check_n$N.json 2> gpurun_out/ddp_graph_check_n$N.err; echo "ddp graph check exit=$?"; cat gpurun_out/ddp_graph_check_n$N.json
timeout -k 5 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench n=$N exit=$?"; tail -n 3 gpurun_out/bench_n$N.err; cat gpurun_out/bench_n$N.json
# BASELINE configs[2]: 32 clips sharded over the GPUs, 10 / 25 / 50 Euler steps
C=$((32 / N))
for S in 10 25 50; do
  timeout -k 5 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --steps 2 --warmup 3 --sample-steps $S --clips-per-gpu $C --no-train --no-secondary --no-cpu-baseline > gpurun_out/sweep_n${N}_s$S.json 2> gpurun_out/sweep_n${N}_s$S.err; echo "sweep n=$N steps=$S exit=$?"
  python -c "import json; d=json.load(open('gpurun_out/sweep_n${N}_s$S.json')); print('config3 N=$N clips/gpu=$C steps=$S:', round(d['value'],3), 'clips/s', round(d['ms_per_step'],1), 'ms')"
done
