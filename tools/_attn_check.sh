mkdir -p gpurun_out
timeout 300 python tools/attn_tc2_check.py > gpurun_out/attn_tc2_check.txt 2>&1; grep "worst\|FAIL\|frames 128" gpurun_out/attn_tc2_check.txt | sed 's/^/v2 /'
HV_ATTN_EXACT=1 timeout 300 python tools/attn_tc2_check.py 2>&1 | grep "worst\|FAIL" | sed 's/^/v2 exact /'
HV_ATTN_TC2=0 timeout 300 python tools/attn_tc2_check.py 2>&1 | grep "frames 128" | sed 's/^/v1 /'
for L in 256 274; do L=$L timeout 120 python tools/attn_timeline.py > gpurun_out/attn_tc2_timeline_L$L.txt 2>&1; sed -n 1p gpurun_out/attn_tc2_timeline_L$L.txt | cut -c1-170; sed -n 6,10p gpurun_out/attn_tc2_timeline_L$L.txt | cut -c1-170; tail -1 gpurun_out/attn_tc2_timeline_L$L.txt; done
