"""attention forward / backward at the training shapes (64 frames, 16 heads): which L is slow?"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hivae_b200 import ops, engine, backward as B
dev = torch.device("cuda:0")
def t(fn, reps=10):
    for _ in range(3): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
H, D = 16, 1024
for frames in (64, 128):
    for L in (264, 272, 274, 288):
        qkv = torch.randn(frames * L, 3 * D, device=dev).bfloat16()
        att = torch.empty(frames * L, D, device=dev, dtype=torch.bfloat16)
        dout = torch.randn(frames * L, D, device=dev).bfloat16()
        g = engine.dense_geom(frames, L)
        q, k, v = qkv[:, :D], qkv[:, D:2 * D], qkv[:, 2 * D:]
        f = t(lambda: ops.attention(q, k, v, att, g.seqs, H, 64, L, L, g.qmap, g.kvmap, 0.125))
        b = t(lambda: B.attention_backward(q, k, v, att, dout, g.seqs, H, L, L, g.qmap, g.kvmap, 0.125))
        fl = 4.0 * frames * H * L * L * 64
        print("frames %3d L %3d: fwd %.3f ms (%.0f TFLOP/s)  bwd %.3f ms (%.0f TFLOP/s algorithmic)" % (frames, L, f, fl / f / 1e9, b, 2.5 * fl / b / 1e9), flush=True)
n, S, T = 4, 256, 16
qkv = torch.randn(n * T * S, 3 * D, device=dev).bfloat16()
att = torch.empty(n * T * S, D, device=dev, dtype=torch.bfloat16)
dout = torch.randn(n * T * S, D, device=dev).bfloat16()
g = engine.strided_geom(n, S, T)
q, k, v = qkv[:, :D], qkv[:, D:2 * D], qkv[:, 2 * D:]
f = t(lambda: ops.attention(q, k, v, att, g.seqs, H, 64, T, T, g.qmap, g.kvmap, 0.125))
b = t(lambda: B.attention_backward(q, k, v, att, dout, g.seqs, H, T, T, g.qmap, g.kvmap, 0.125))
print("temporal %d seqs x 16 tokens: fwd %.3f ms, bwd %.3f ms (bytes fwd %.0f MB)" % (g.seqs, f, b, n * T * S * D * 2 * 4 / 1e6))
