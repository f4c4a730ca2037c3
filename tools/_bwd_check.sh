mkdir -p gpurun_out
timeout -k 5 900 python -m pytest tests -x -q -m gpu -k "attention or block or train or golden" > gpurun_out/pytest_attn.log 2>&1; echo "pytest exit=$?"; tail -3 gpurun_out/pytest_attn.log
timeout -k 5 300 python tools/bench_attn_shapes.py > gpurun_out/attn_shapes.txt 2>&1; echo "shapes exit=$?"; sed -n 3p gpurun_out/attn_shapes.txt; sed -n 7p gpurun_out/attn_shapes.txt
HV_ATTN_BWD_PROF=1 timeout -k 5 300 python tools/attn_bwd_phases.py > gpurun_out/attn_bwd_phases.txt 2>&1; head -19 gpurun_out/attn_bwd_phases.txt
timeout -k 5 300 python tools/bench_train.py --steps 5 --graph > gpurun_out/bench_train_graph.json 2> gpurun_out/bench_train_graph.err; echo "train exit=$?"; cut -c1-260 gpurun_out/bench_train_graph.json
