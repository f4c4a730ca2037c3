mkdir -p gpurun_out
HV_ATTN_TC2=1 timeout -k 5 300 python tools/attn_tc2_check.py > gpurun_out/attn_tc2_check.txt 2>&1; echo "v2 exit=$?"; cat gpurun_out/attn_tc2_check.txt
HV_ATTN_TC2=1 HV_ATTN_EXACT=1 timeout -k 5 300 python tools/attn_tc2_check.py > gpurun_out/attn_tc2_check_exact.txt 2>&1; echo "v2 exact exit=$?"; tail -12 gpurun_out/attn_tc2_check_exact.txt
timeout -k 5 300 python tools/attn_tc2_check.py > gpurun_out/attn_tc2_check_v1.txt 2>&1; echo "v1 exit=$?"; tail -11 gpurun_out/attn_tc2_check_v1.txt
bash tools/probes/run_probes.sh
