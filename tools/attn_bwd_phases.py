"""HV_ATTN_BWD_PROF=1 python tools/attn_bwd_phases.py — cycles per tile and phase of the tcgen05 attention-backward kernels"""
import ctypes, os, sys
os.environ["HV_ATTN_BWD_PROF"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hivae_b200 import _lib, engine, backward as B
dev = torch.device("cuda:0")
names = {0: ["-", "wait S", "pass 1 (max)", "max exchange", "pass 2 (exp -> regs)", "sum exchange + stats", "wait dP", "pass 3 (dS -> smem)", "-", "deferred dQ read-back"],
         1: ["-", "wait S^T", "-", "stage P^T", "pass 2 (exp -> regs)", "stage dS^T", "wait dP^T", "pass 3 (dS^T in regs)", "wait dV", "deferred dV / dK read-back"]}
for frames, L in ((64, 274), (128, 264)):
    H, D = 16, 1024
    qkv = torch.randn(frames * L, 3 * D, device=dev).bfloat16()
    att, dout = (torch.randn(frames * L, D, device=dev).bfloat16() for _ in range(2))
    g = engine.dense_geom(frames, L)
    buf = (ctypes.c_ulonglong * (2 * 148 * 16))()
    B.attention_backward(qkv[:, :D], qkv[:, D:2 * D], qkv[:, 2 * D:], att, dout, g.seqs, H, L, L, g.qmap, g.kvmap, 0.125)
    _lib.load().hv_attn_bwd_prof_read(buf)
    for _ in range(3):
        B.attention_backward(qkv[:, :D], qkv[:, D:2 * D], qkv[:, 2 * D:], att, dout, g.seqs, H, L, L, g.qmap, g.kvmap, 0.125)
    assert _lib.load().hv_attn_bwd_prof_read(buf)
    for k, kn in ((0, "dq"), (1, "dkv")):
        tot = [sum(buf[(k * 148 + c) * 16 + p] for c in range(148)) for p in range(16)]
        tiles = tot[15]
        print("frames %d L %d, %s kernel: %d tiles, %.0f clk per tile" % (frames, L, kn, tiles, sum(tot[:10]) / tiles))
        for p in range(10):
            if names[k][p] != "-":
                print("   %-32s %8.0f" % (names[k][p], tot[p] / tiles))
