mkdir -p gpurun_out
timeout -k 5 900 python -m pytest tests -x -q -m gpu -k "attention or block or train or golden" > gpurun_out/pytest_attn.log 2>&1; echo "pytest exit=$?"; tail -3 gpurun_out/pytest_attn.log
timeout -k 5 300 python tools/bench_attn_shapes.py > gpurun_out/attn_shapes.txt 2>&1; echo "shapes exit=$?"; cat gpurun_out/attn_shapes.txt
timeout -k 5 300 python tools/bench_train.py --steps 5 --graph > gpurun_out/bench_train_graph.json 2> gpurun_out/bench_train_graph.err; echo "train exit=$?"; cut -c1-400 gpurun_out/bench_train_graph.json
bash tools/ncu_one.sh attn attn_tc2_kernel
