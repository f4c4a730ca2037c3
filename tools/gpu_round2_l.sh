mkdir -p gpurun_out
cat > /tmp/one_bwd.py <<'PY'
import os, sys
sys.path.insert(0, os.getcwd())
import torch
from hivae_b200 import engine, backward as B
dev = torch.device("cuda:0")
frames, L, H, D = 64, 274, 16, 1024
qkv = torch.randn(frames * L, 3 * D, device=dev).bfloat16()
att, dout = (torch.randn(frames * L, D, device=dev).bfloat16() for _ in range(2))
g = engine.dense_geom(frames, L)
for _ in range(3):
    B.attention_backward(qkv[:, :D], qkv[:, D:2 * D], qkv[:, 2 * D:], att, dout, g.seqs, H, L, L, g.qmap, g.kvmap, 0.125)
torch.cuda.synchronize()
PY
python /tmp/one_bwd.py && timeout 600 ncu --set full --clock-control none --import-source on -k regex:'attn_bwd_d' -s 2 -c 2 -o gpurun_out/prof_attn_bwd -f python /tmp/one_bwd.py > gpurun_out/ncu_attn_bwd.log 2>&1
echo "ncu exit=$?"; ls -la gpurun_out/*.ncu-rep
