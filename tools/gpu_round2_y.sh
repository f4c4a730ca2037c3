mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout -k 5 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 tools/ddp_graph_check.py > gpurun_out/ddp_graph_check_n$N.json 2> gpurun_out/ddp_graph_check_n$N.err; echo "exit=$?"; cat gpurun_out/ddp_graph_check_n$N.json; tail -15 gpurun_out/ddp_graph_check_n$N.err | cut -c1-300
