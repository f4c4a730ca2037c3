"""Timeline of CTA 0 of attn_tc2_kernel (HV_ATTN_TIMELINE=1): clock64 stamps per tile, relative to the first stamp.
   L=256 python tools/attn_timeline.py
events per tile: key half h (warp quarter 0): S ready, exp done, arrived on p_full; key half 0 also: o_full seen, epilogue done
(both for the tile itself, executed during the next tile); bounds done (norm warp); MMA warp: top of iteration, p_full seen,
P V issued, S(T+1).H1 issued."""
import os, sys
os.environ["HV_ATTN_TIMELINE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes as C
import numpy as np
import torch
from hivae_b200 import _lib, engine, ops
dev = torch.device("cuda:0")
L, H, D, seqs = int(os.environ.get("L", 256)), 16, 1024, 128
qkv = (torch.randn(seqs * L, 3 * D, device=dev) * 0.5).to(torch.bfloat16)
att = torch.empty(seqs * L, D, device=dev, dtype=torch.bfloat16)
g = engine.dense_geom(seqs, L)
run = lambda: ops.attention(qkv[:, :D], qkv[:, D:2 * D], qkv[:, 2 * D:], att, g.seqs, H, 64, L, L, g.qmap, g.kvmap, 0.125)
buf = np.zeros(1024, dtype=np.uint64)
run(); torch.cuda.synchronize()
_lib.load().hv_attn_timeline_read(buf.ctypes.data_as(C.c_void_p))
run(); torch.cuda.synchronize()
assert _lib.load().hv_attn_timeline_read(buf.ctypes.data_as(C.c_void_p)) == 1
t = buf.reshape(64, 16).astype(np.int64)
t0 = t[t > 0].min()
names = {9: "n k_full", 10: "n K done", 5: "n q_full", 11: "bounds", 15: "mma top", 0: "h0 S rdy", 1: "h0 exp", 2: "h0 arrive", 6: "h1 S rdy", 7: "h1 exp", 8: "h1 arrive", 12: "mma p_full",
         13: "pv issued", 14: "s.h1 iss", 3: "h0 o_full", 4: "h0 epi"}
order = [9, 10, 5, 11, 15, 0, 1, 2, 6, 7, 8, 12, 13, 14, 3, 4]
print("L = %d; " % L + " | ".join(names[i] for i in order))
for T in range(10, 22):
    print("tile %2d: " % T + " ".join("%8d" % (t[T, i] - t0 if t[T, i] else -1) for i in order))
per = (t[21, 8] - t[11, 8]) / 10.0
print("cycles per tile (h1 arrive to h1 arrive, tiles 11..21): %.0f" % per)
