"""Per-phase cycle breakdown of the softmax warps of attn_tc_kernel (HV_ATTN_PROF=1), DiT spatial shape."""
import os, sys
os.environ["HV_ATTN_PROF"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes as C
import numpy as np
import torch
from hivae_b200 import _lib, engine, ops
dev = torch.device("cuda:0")
L, H, D, seqs = int(os.environ.get("L", 274)), 16, 1024, 128
qkv = (torch.randn(seqs * L, 3 * D, device=dev) * 0.5).to(torch.bfloat16)
att = torch.empty(seqs * L, D, device=dev, dtype=torch.bfloat16)
g = engine.dense_geom(seqs, L)
run = lambda: ops.attention(qkv[:, :D], qkv[:, D:2 * D], qkv[:, 2 * D:], att, g.seqs, H, 64, L, L, g.qmap, g.kvmap, 0.125)
buf = np.zeros(148 * 16, dtype=np.uint64)
run(); torch.cuda.synchronize()
_lib.load().hv_attn_prof_read(buf.ctypes.data_as(C.c_void_p))  # discard the warm-up
run(); torch.cuda.synchronize()
assert _lib.load().hv_attn_prof_read(buf.ctypes.data_as(C.c_void_p)) == 1
c = buf.reshape(148, 2, 8).astype(np.float64)
names = ["wait S", "max pass", "max exchange", "O epilogue", "exp pass", "publish P + sums"]
for h in range(2):
    tiles = c[:, h, 6].sum() + c[:, h, 7].sum()
    print("key half %d: %d full + %d tail tiles over 148 CTAs; cycles per tile (all tiles):" % (h, c[:, h, 6].sum(), c[:, h, 7].sum()))
    tot = 0.0
    for i, n in enumerate(names):
        v = c[:, h, i].sum() / tiles
        tot += v
        print("   %-18s %8.0f" % (n, v))
    print("   %-18s %8.0f" % ("sum", tot))
