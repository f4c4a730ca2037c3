"""attention forward (attn_tc2.cu): accuracy against fp64 over the shapes of the path + the bound / exact switch, and timing.
   python tools/attn_tc2_check.py            (HV_ATTN_TC2=0 -> first design, HV_ATTN_EXACT=1 -> exact maximum pass)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hivae_b200 import ops, engine
dev = torch.device("cuda:0")


def ref(q, k, v, B, H, Lq, Lkv, scale):
    q, k, v = (x.double().view(B, -1, H, 64).transpose(1, 2) for x in (q, k, v))
    s = torch.einsum("bhqd,bhkd->bhqk", q, k) * scale
    return torch.einsum("bhqk,bhkd->bhqd", torch.softmax(s, -1), v).transpose(1, 2).reshape(B * Lq, H * 64)


def t(fn, reps=20):
    for _ in range(3): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


print("variant: HV_ATTN_TC2=%s HV_ATTN_EXACT=%s" % (os.environ.get("HV_ATTN_TC2", "1"), os.environ.get("HV_ATTN_EXACT", "0")))
worst = 0.0
torch.manual_seed(0)
for (B, H, Lq, Lkv) in [(3, 2, 274, 274), (2, 8, 264, 264), (2, 4, 272, 272), (1, 16, 64, 128), (2, 2, 100, 49), (5, 3, 288, 288),
                        (2, 2, 129, 200), (3, 1, 160, 48), (2, 2, 256, 256), (1, 2, 20, 274), (300, 16, 274, 274), (7, 2, 33, 64)]:
    for qscale in (1.0, 12.0):   # 12: |q| |k| scale log2e ~ 12*8*8*0.125*1.44 = 138 > 56 -> the kernel itself switches to the exact path
        D = H * 64
        q = (torch.randn(B * Lq, D, device=dev) * qscale).bfloat16()
        k = torch.randn(B * Lkv, D, device=dev).bfloat16()
        v = torch.randn(B * Lkv, D, device=dev).bfloat16()
        out = torch.empty(B * Lq, D, device=dev, dtype=torch.bfloat16)
        ops.attention(q, k, v, out, B, H, 64, Lq, Lkv, ops.seq_map(1, Lq, 0, 1), ops.seq_map(1, Lkv, 0, 1), 0.125)
        torch.cuda.synchronize()
        r = ref(q, k, v, B, H, Lq, Lkv, 0.125)
        err = float((out.double() - r).abs().max() / r.abs().max())
        worst = max(worst, err)
        print("B %3d H %2d Lq %3d Lkv %3d qscale %4.1f: rel err %.2e %s" % (B, H, Lq, Lkv, qscale, err, "" if err < 1e-2 else "  <-- FAIL"), flush=True)
print("worst rel err %.2e" % worst)
H, D = 16, 1024
for frames in (64, 128):
    for L in (256, 264, 272, 274, 288):
        qkv = torch.randn(frames * L, 3 * D, device=dev).bfloat16()
        att = torch.empty(frames * L, D, device=dev, dtype=torch.bfloat16)
        g = engine.dense_geom(frames, L)
        q, k, v = qkv[:, :D], qkv[:, D:2 * D], qkv[:, 2 * D:]
        f = t(lambda: ops.attention(q, k, v, att, g.seqs, H, 64, L, L, g.qmap, g.kvmap, 0.125))
        fl = 4.0 * frames * H * L * L * 64
        print("frames %3d L %3d: fwd %.3f ms (%.0f TFLOP/s)" % (frames, L, f, fl / f / 1e9), flush=True)
