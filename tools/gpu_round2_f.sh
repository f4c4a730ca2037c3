mkdir -p gpurun_out
N=${NGPU:-2}
timeout -k 5 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/ddp_check.py > gpurun_out/ddp_check_n$N.json 2> gpurun_out/ddp_check_n$N.err; echo "ddp check exit=$?"; cat gpurun_out/ddp_check_n$N.json; tail -n 5 gpurun_out/ddp_check_n$N.err
timeout -k 5 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 2 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench n=$N exit=$?"; tail -n 5 gpurun_out/bench_n$N.err; cat gpurun_out/bench_n$N.json
