mkdir -p gpurun_out
timeout -k 5 300 python tools/bench_attn_shapes.py > gpurun_out/attn_shapes.txt 2>&1; echo "attn shapes exit=$?"; cat gpurun_out/attn_shapes.txt
CMD="python tools/bench_train.py --clips 4 --steps 1 --warmup 1"
timeout -k 5 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/train_launches.csv $CMD > gpurun_out/ncu_train.log 2>&1
echo "train launch list exit=$?"; tail -3 gpurun_out/ncu_train.log; wc -l gpurun_out/train_launches.csv
