"""micro-benchmark of the bandwidth-bound kernels (BASELINE.json config 4: Haar / frequency sweep) plus
LayerNorm-modulate and the attention kernels at the DiT's shapes; GB/s (algorithmic bytes) per launch."""
import sys, os, math, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import hivae_b200 as hv
from hivae_b200 import ops, engine
from hivae_b200.frequency_utils import freq_split_gauss
torch.set_grad_enabled(False)
dev = torch.device("cuda:0")
reps = int(os.environ.get("REPS", 20))
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
def timeit(fn, flush_l2=True):
    for _ in range(3): fn()
    tot = 0.0
    for _ in range(reps):
        if flush_l2: flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / reps
res = {}
def report(name, ms, nbytes=None, flops=None):
    s = "%-44s %8.3f ms" % (name, ms)
    if nbytes: s += " %8.1f GB/s" % (nbytes / ms / 1e6); res[name] = round(nbytes / ms / 1e6, 1)
    if flops: s += " %8.1f TFLOP/s" % (flops / ms / 1e9); res[name + "/tflops"] = round(flops / ms / 1e9, 1)
    print(s, flush=True)
def timeit_rotating(make_fn, nbytes_one, reps=5):
    """kernel-level time of a short kernel: `copies` independent input sets totalling > 2x L2 are processed back to
    back inside ONE CUDA graph (no host gaps, every call sees cold data); time per call = graph time / copies."""
    copies = max(4, int(2 * 126e6 / max(nbytes_one, 1)) + 1)
    copies = min(copies, 64)
    fns = [make_fn(i) for i in range(copies)]
    for f in fns: f()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for f in fns: f()
    g.replay(); torch.cuda.synchronize()
    tot = 0.0
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / reps / copies
# ---- Haar pyramids (fp32 + bf16), 2-3 levels (BASELINE.json configs[3]); algorithmic bytes = SURVEY 8d formula
for (T, H, L) in ((17, 256, 2), (33, 256, 3), (17, 384, 2), (33, 384, 3), (17, 512, 2), (33, 512, 3)):
    for dt in (torch.float32, torch.bfloat16):
        es = 4 if dt == torch.float32 else 2
        numel = T * 3 * H * H
        geo = sum(4.0 ** -k for k in range(L))
        xs = {}
        def mk_fwd(i):
            xs[i] = torch.randn(T, 3, H, H, device=dev).to(dt)
            return lambda: hv.dwt_pyramid(xs[i], L)
        report("dwt_pyramid %s T=%d %dx%d L=%d" % (str(dt)[6:], T, H, H, L), timeit_rotating(mk_fwd, numel * es), 2 * numel * es * geo)
        pyr = {}
        def mk_inv(i):
            pyr[i] = hv.dwt_pyramid(xs[i % len(xs)], L)
            return lambda: hv.iwt_pyramid(*pyr[i])
        report("iwt_pyramid %s T=%d %dx%d L=%d" % (str(dt)[6:], T, H, H, L), timeit_rotating(mk_inv, numel * es), numel * geo * (es + 4))
        del xs, pyr
for n in (8, 32):
    x = torch.randn(n, 4, 32, 32, 32, device=dev)
    report("freq_split3d n=%d" % n, timeit(lambda: freq_split_gauss(x, 0.6, 0.6)), 3 * x.numel() * 4)
# ---- LayerNorm + adaLN modulate at the DiT shape (fp32 residual -> bf16 activations)
frames, L, D = 128, 274, 1024
r = torch.randn(frames, L, D, device=dev)
table = torch.randn(frames, 6 * D, device=dev)
w, b = torch.ones(D, device=dev), torch.zeros(D, device=dev)
out = torch.empty(frames * L, D, device=dev, dtype=torch.bfloat16)
rows = ops.Rows(r, 0, L, 18, 256)
report("ln_modulate joint 35072x1024 fp32->bf16", timeit(lambda: ops.ln_modulate(rows, w, b, 1e-5, out, ops.Ada(table), 0, D, 3 * D, 4 * D)),
       frames * L * D * 6)
report("ln_modulate plain 35072x1024 fp32->bf16", timeit(lambda: ops.ln_modulate(rows, w, b, 1e-5, out)), frames * L * D * 6)
# ---- attention kernels
H = 16
qkv = torch.randn(frames * L, 3 * D, device=dev).bfloat16()
att = torch.empty(frames * L, D, device=dev, dtype=torch.bfloat16)
g = engine.dense_geom(frames, L)
report("attention spatial 128x16 heads L=274", timeit(lambda: ops.attention(qkv[:, :D], qkv[:, D:2 * D], qkv[:, 2 * D:], att, g.seqs, H, 64, L, L, g.qmap, g.kvmap, 0.125)),
       frames * L * D * 2 * 4, 4.0 * frames * H * L * L * 64)
n, S, T = 8, 256, 16
qkv2 = torch.randn(n * T * S, 3 * D, device=dev).bfloat16()
att2 = torch.empty(n * T * S, D, device=dev, dtype=torch.bfloat16)
g2 = engine.strided_geom(n, S, T)
report("attention temporal 2048x16 heads L=16", timeit(lambda: ops.attention(qkv2[:, :D], qkv2[:, D:2 * D], qkv2[:, 2 * D:], att2, g2.seqs, H, 64, T, T, g2.qmap, g2.kvmap, 0.125)),
       n * T * S * D * 2 * 4, 4.0 * g2.seqs * H * T * T * 64)
print(json.dumps(res))
