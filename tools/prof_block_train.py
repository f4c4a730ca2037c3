"""One forward + backward of a BasicTransformerBlock at the object-encoder shape (128 x 264 tokens, D = 512, bf16) for the
ncu launch list: `ncu --metrics gpu__time_duration.sum --clock-control none --csv python tools/prof_block_train.py`."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hivae_b200 import backward as Bk
from hivae_b200.modules import BasicTransformerBlock
dev = torch.device("cuda:0")
torch.manual_seed(0)
blk = BasicTransformerBlock(512, 8, 64, attention_bias=True).to(dev)
blk.hv_compute_dtype = torch.bfloat16
x = torch.randn(128, 264, 512, device=dev, requires_grad=True)
dy = torch.randn(128, 264, 512, device=dev)
for _ in range(int(os.environ.get("REPS", "1"))):
    Bk.basic_block_apply(blk, x).backward(dy)
torch.cuda.synchronize()
print("done")
