mkdir -p gpurun_out
for L in 256 274; do L=$L timeout 120 python tools/attn_timeline.py > gpurun_out/attn_tc2_timeline_L$L.txt 2>&1; sed -n 1,1p gpurun_out/attn_tc2_timeline_L$L.txt; sed -n 6,9p gpurun_out/attn_tc2_timeline_L$L.txt; tail -1 gpurun_out/attn_tc2_timeline_L$L.txt; done
timeout 300 python tools/attn_tc2_check.py > gpurun_out/attn_tc2_check.txt 2>&1; grep "worst\|FAIL\|frames 128" gpurun_out/attn_tc2_check.txt
HV_ATTN_TC2=0 timeout 300 python tools/attn_tc2_check.py 2>&1 | grep "frames 128" | sed 's/^/v1 /'
