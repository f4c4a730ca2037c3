"""Forward + backward of one BasicTransformerBlock (the object motion encoder's block, D=512, 8 heads, 264 tokens per
frame, 32 frames per clip -> 128 sequences for the 4 clips/GPU of scripts/amd_new.sh:40) through
backward.basic_block_apply.  CUDA events, 10 reps after warm-up.  Algorithmic FLOPs: forward 24 D^2 M + 4 L^2 D per
sequence, backward 2x the linears + 2.5x the attention (recompute included)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hivae_b200 import backward as Bk, ops
from hivae_b200.modules import BasicTransformerBlock
dev = torch.device("cuda:0")
torch.manual_seed(0)
B, L, D, H = 128, 264, 512, 8
blk = BasicTransformerBlock(D, H, 64, attention_bias=True).to(dev)
blk.hv_compute_dtype = torch.bfloat16
x = torch.randn(B, L, D, device=dev, requires_grad=True)
dy = torch.randn(B, L, D, device=dev)


def timeit(fn, reps=10):
    fn(); fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def fwd_bwd():
    for p in blk.parameters(): p.grad = None
    x.grad = None
    with torch.enable_grad():
        Bk.basic_block_apply(blk, x).backward(dy)


M = B * L
lin, att = 24.0 * D * D * M, 4.0 * L * L * D * B
with torch.no_grad():
    ms_inf = timeit(lambda: blk(x))
n0 = ops.launch_count()
ms = timeit(fwd_bwd)
launches = (ops.launch_count() - n0) // 12
print("BasicTransformerBlock D=%d heads=%d, %d sequences x %d tokens (M=%d), bf16 compute" % (D, H, B, L, M))
print("inference forward            %8.3f ms  %7.1f TFLOP/s" % (ms_inf, (lin + att) / ms_inf / 1e9))
print("training forward + backward  %8.3f ms  %7.1f TFLOP/s algorithmic (3x forward), %d kernel launches" % (ms, 3.0 * (lin + att) / ms / 1e9, launches))

# ---- component breakdown of the backward at the same shape
ctx, bp, _ = blk._prepared(x)
with torch.no_grad():
    ms_ft = timeit(lambda: Bk.block_forward_train(ctx, bp, x))
    out, s = Bk.block_forward_train(ctx, bp, x)
    print("training forward (saves activations, 2 extra GEMMs) %8.3f ms" % ms_ft)
    d16 = dy.reshape(M, D).to(torch.bfloat16)
    dF = torch.randn(M, 4 * D, device=dev).to(torch.bfloat16)
    d3 = torch.randn(M, 3 * D, device=dev).to(torch.bfloat16)
    g = s["geom"]
    for name, fn in (
        ("linear_backward ff2 (N=512, K=2048)", lambda: Bk.linear_backward(d16, s["h"], bp.w2)),
        ("gelu_tanh_backward", lambda: Bk.gelu_tanh_backward(s["hpre"], dF)),
        ("linear_backward ff1 (N=2048, K=512)", lambda: Bk.linear_backward(dF, s["xn2"], bp.w1)),
        ("layernorm_backward (accumulate)", lambda: Bk.layernorm_backward(s["x1"].view(M, D), d16, bp.n2w, bp.eps, dx=out.view(M, D), accumulate=True)),
        ("linear_backward out (N=512, K=512)", lambda: Bk.linear_backward(d16, s["att"], bp.wo)),
        ("attention_backward", lambda: Bk.attention_backward(s["q"], s["k"], s["v"], s["att"], d16, g.seqs, H, L, L, g.qmap, g.kvmap, s["scale"])),
        ("headnorm64_backward", lambda: Bk.headnorm64_backward(s["q_pre"], d16, bp.qn_w, H)),
        ("linear_backward qkv (N=1536, K=512)", lambda: Bk.linear_backward(d3, s["xn1"], bp.wqkv)),
        ("transpose2d (M x 2048)", lambda: Bk.transpose2d(dF)),
        ("colsum (M x 2048)", lambda: Bk.colsum(dF)),
    ):
        print("  %-40s %8.3f ms" % (name, timeit(fn)), flush=True)
