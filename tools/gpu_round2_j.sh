mkdir -p gpurun_out
timeout -k 5 120 python tools/attn_bwd_phases.py > gpurun_out/attn_bwd_phases.txt 2>&1; cat gpurun_out/attn_bwd_phases.txt
