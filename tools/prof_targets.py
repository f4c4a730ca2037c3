"""One or two launches of a chosen kernel at the DiT's shapes, for `ncu -k regex:<kernel> --set full` captures.
usage: python tools/prof_targets.py attn|ln|small|gemm_out|gemm_ff1|gemm_qkv|dwt|freq|attn_bwd"""
import sys, os, math
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import hivae_b200 as hv
from hivae_b200 import ops, engine
torch.set_grad_enabled(False)
dev = torch.device("cuda:0")
what = sys.argv[1]
frames, L, D, H = 128, 274, 1024, 16
M = frames * L
reps = 2
if what == "attn":
    qkv = torch.randn(M, 3 * D, device=dev).bfloat16()
    att = torch.empty(M, D, device=dev, dtype=torch.bfloat16)
    g = engine.dense_geom(frames, L)
    for _ in range(reps):
        ops.attention(qkv[:, :D], qkv[:, D:2 * D], qkv[:, 2 * D:], att, g.seqs, H, 64, L, L, g.qmap, g.kvmap, 0.125)
elif what == "small":
    n, S, T = 8, 256, 16
    qkv = torch.randn(n * T * S, 3 * D, device=dev).bfloat16()
    att = torch.empty(n * T * S, D, device=dev, dtype=torch.bfloat16)
    g = engine.strided_geom(n, S, T)
    for _ in range(reps):
        ops.attention(qkv[:, :D], qkv[:, D:2 * D], qkv[:, 2 * D:], att, g.seqs, H, 64, T, T, g.qmap, g.kvmap, 0.125)
elif what == "ln":
    r = torch.randn(frames, L, D, device=dev)
    table = torch.randn(frames, 6 * D, device=dev)
    w, b = torch.ones(D, device=dev), torch.zeros(D, device=dev)
    out = torch.empty(M, D, device=dev, dtype=torch.bfloat16)
    rows = ops.Rows(r, 0, L, 18, 256)
    for _ in range(reps):
        ops.ln_modulate(rows, w, b, 1e-5, out, ops.Ada(table), 0, D, 3 * D, 4 * D)
elif what.startswith("gemm"):
    g = torch.Generator().manual_seed(0)
    mk = lambda r_, c: (torch.randn(r_, c, generator=g) / math.sqrt(c)).bfloat16().to(dev)
    N, K = {"gemm_out": (D, D), "gemm_ff2": (D, 4 * D), "gemm_ff1": (4 * D, D), "gemm_qkv": (3 * D, D)}[what]
    A, W, bias = mk(M, K), mk(N, K), torch.randn(N, device=dev)
    res = torch.randn(frames, L, D, device=dev)
    table = torch.randn(frames, 6 * D, device=dev)
    rows = ops.Rows(res, 0, L, 18, 256)
    outb = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    ln = torch.ones(64, device=dev), torch.zeros(64, device=dev)
    for _ in range(reps):
        if N == D:
            ops.linear(A, W, ops.epi_residual(rows, bias, ops.Ada(table, 1, 1), 2 * D, 5 * D))
        elif what == "gemm_ff1":
            ops.linear(A, W, ops.epi_store(outb, bias, gelu=True))
        else:
            ops.linear(A, W, ops.epi_qknorm(outb, bias, ln[0], ln[1], ln[0], ln[1], D, D))
elif what == "freq":
    from hivae_b200.frequency_utils import freq_split_gauss
    x = torch.randn(128, 4, 32, 32, 32, device=dev)
    for _ in range(reps):
        freq_split_gauss(x, 0.6, 0.6)
elif what == "dwt":
    x = torch.randn(33, 3, 512, 512, device=dev)
    for _ in range(reps):
        bands, ll = hv.dwt_pyramid(x, 3)
        hv.iwt_pyramid(bands, ll)
elif what == "attn_bwd_dit":
    from hivae_b200 import backward as Bk
    frames, L, D, H = 64, 274, 1024, 16  # DiT object block, 4 clips (train step)
    M = frames * L
    qkv = torch.randn(M, 3 * D, device=dev).bfloat16()
    o, do = torch.randn(M, D, device=dev).bfloat16(), torch.randn(M, D, device=dev).bfloat16()
    g = engine.dense_geom(frames, L)
    for _ in range(reps):
        Bk.attention_backward(qkv[:, :D], qkv[:, D:2 * D], qkv[:, 2 * D:], o, do, g.seqs, H, L, L, g.qmap, g.kvmap, 0.125)
elif what == "wgrad":
    M = 17536  # 4 clips x 16 frames x 274 tokens
    dy, x = torch.randn(M, 4 * D, device=dev).bfloat16(), torch.randn(M, D, device=dev).bfloat16()
    dw = torch.zeros(4 * D, D, device=dev)
    for _ in range(reps):
        e = ops.epi_residual(ops.Rows(dw.view(1, 4 * D, D), 0, 4 * D))
        e.k_split = 1
        ops.linear_tn(dy, x, e)
elif what == "attn_bwd":
    from hivae_b200 import backward as Bk
    frames, L, D, H = 128, 264, 512, 8  # object motion encoder, 4 clips
    M = frames * L
    qkv = torch.randn(M, 3 * D, device=dev).bfloat16()
    o, do = torch.randn(M, D, device=dev).bfloat16(), torch.randn(M, D, device=dev).bfloat16()
    g = engine.dense_geom(frames, L)
    for _ in range(reps):
        Bk.attention_backward(qkv[:, :D], qkv[:, D:2 * D], qkv[:, 2 * D:], o, do, g.seqs, H, L, L, g.qmap, g.kvmap, 0.125)
torch.cuda.synchronize()
print("done", what)
