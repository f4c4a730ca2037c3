mkdir -p gpurun_out
for v in 0 1 2 3; do echo "HV_ATTN_BWD_PIPE=$v"; HV_ATTN_BWD_PIPE=$v timeout -k 5 120 python tools/bench_attn_shapes.py 2>&1 | grep "frames  64 L 27[24]\|frames 128 L 274"; done > gpurun_out/attn_variants.txt 2>&1
cat gpurun_out/attn_variants.txt
CMD="python tools/bench_attn_shapes.py"
timeout -k 5 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/attn_launches.csv -k regex:'attn_bwd|attn_small_bwd' $CMD > /dev/null 2>&1
python tools/launch_summary.py gpurun_out/attn_launches.csv | head -12
