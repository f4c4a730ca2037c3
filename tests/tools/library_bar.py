"""LIBRARY BAR (test infrastructure, not product): the reference's real execution path on the same B200 — its modules in
torch eager on `cuda` under torch.autocast(bf16) (cuBLASLt GEMMs, F.scaled_dot_product_attention, ATen LayerNorm; call sites
model/modules.py:699-719, 1043-1063) — timed next to hivae_b200 on BASELINE configs[1] (default 695 M model, batch 8,
encode + 20 Euler steps), plus the per-op library times at the DiT shapes.

The reference itself cannot be imported on the GPU box (no /root/reference there, diffusers / timm absent), so its torch
path is run through the oracle restatement (oracle/amd_oracle.py, checked against the unmodified reference in
tests/golden/make_golden*.py) with the attention routed through F.scaled_dot_product_attention exactly like diffusers'
AttnProcessor2_0.  Lives under tests/ because only tests/ may import oracle/.

    python tests/tools/library_bar.py [--clips 8] [--steps 20] > profiles/r02_library_bar.json
"""
import argparse
import json
import math
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import amd_oracle as O  # noqa: E402
import hivae_b200 as hv  # noqa: E402
from hivae_b200 import configs, ops, synth  # noqa: E402


def sdpa_attention(sd, p, x, enc, heads):
    """diffusers Attention + AttnProcessor2_0: same arithmetic as O.attention, through F.scaled_dot_product_attention"""
    B = x.shape[0]
    enc = x if enc is None else enc
    q, k, v = O._lin(sd, p + ".to_q", x), O._lin(sd, p + ".to_k", enc), O._lin(sd, p + ".to_v", enc)
    hd = q.shape[-1] // heads
    q, k, v = (z.view(B, -1, heads, hd).transpose(1, 2) for z in (q, k, v))
    q = O._ln(sd, p + ".norm_q", q, 1e-6)
    k = O._ln(sd, p + ".norm_k", k, 1e-6)
    o = F.scaled_dot_product_attention(q, k, v, dropout_p=0.0, is_causal=False)
    o = o.transpose(1, 2).reshape(B, -1, heads * hd).to(q.dtype)
    return O._lin(sd, p + ".to_out.0", o)


def timeit(fn, reps, warm=2):
    for _ in range(warm):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--clips", type=int, default=8)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--reps", type=int, default=2)
    a = ap.parse_args()
    torch.set_grad_enabled(False)
    dev = torch.device("cuda:0")
    cfg = configs.DEFAULT
    m = hv.AMDModel_New(**cfg)
    sd = synth.synth_state_dict(((k, v.shape) for k, v in m.state_dict().items()), 0)
    m.load_state_dict(sd)
    m.to(dev).set_compute_dtype(torch.bfloat16)
    sdc = {k: v.to(dev) for k, v in sd.items()}
    O.attention = sdpa_attention
    orc = O.AMDOracle(sdc, cfg["enc_nhead"], cfg["diffusion_attn_num_heads"], 16, 2, cfg["motion_token_num"])
    video, ref_img = synth.split_clip(synth.synth_clip_latents(a.clips, seed=1).to(dev))
    z0 = torch.randn(a.clips * 16, 4, 32, 32, device=dev)
    out = {"config": "configs[1]: default 695M model, batch %d, encode + %d Euler steps, one B200" % (a.clips, a.steps)}

    def lib_bf16():
        with torch.autocast("cuda", dtype=torch.bfloat16):
            return orc.sample(video, ref_img, sample_step=a.steps, z0=z0)[1]

    def lib_fp32():
        return orc.sample(video, ref_img, sample_step=a.steps, z0=z0)[1]

    def ours():
        return m.sample(video, ref_img, sample_step=a.steps)[1]

    ms = timeit(ours, a.reps)
    out["hivae_b200_bf16"] = {"ms": round(ms, 1), "clips_per_s": round(a.clips / ms * 1e3, 3)}
    ms = timeit(lib_bf16, a.reps, warm=1)
    out["torch_eager_autocast_bf16"] = {"ms": round(ms, 1), "clips_per_s": round(a.clips / ms * 1e3, 3)}
    torch.backends.cuda.matmul.allow_tf32 = False
    ms = timeit(lib_fp32, 1, warm=1)
    out["torch_eager_fp32"] = {"ms": round(ms, 1), "clips_per_s": round(a.clips / ms * 1e3, 3)}
    out["speedup_vs_eager_autocast_bf16"] = round(out["hivae_b200_bf16"]["clips_per_s"] / out["torch_eager_autocast_bf16"]["clips_per_s"], 3)
    # numerical agreement of the two bf16 paths on the same inputs (fixed noise)
    orig = m.scheduler.get_train_tuple
    m.scheduler.get_train_tuple = lambda z0_=None, z1=None, time_step=None: orig(z0=z0, z1=z1, time_step=time_step)
    a_, b_ = ours().float().flatten(), lib_bf16().float().flatten()
    out["max_abs_rel_ours_vs_eager_bf16"] = float((a_ - b_).abs().max() / b_.abs().max())

    # ---- per-op library times at the DiT shapes (128 frames x 274 tokens, D = 1024, 16 heads)
    ops_out = {}
    frames, L, D, H = 16 * a.clips, 274, 1024, 16
    M = frames * L
    q = torch.randn(frames, H, L, 64, device=dev, dtype=torch.bfloat16)
    ms = timeit(lambda: F.scaled_dot_product_attention(q, q, q), 20)
    fl = 4.0 * frames * H * L * L * 64
    ops_out["sdpa_bf16_%dx%d_L%d" % (frames, H, L)] = {"ms": round(ms, 4), "tflops": round(fl / ms / 1e9, 1)}
    qkv = torch.randn(M, 3 * D, device=dev, dtype=torch.bfloat16)
    att = torch.empty(M, D, device=dev, dtype=torch.bfloat16)
    from hivae_b200 import engine
    g = engine.dense_geom(frames, L)
    ms = timeit(lambda: ops.attention(qkv[:, :D], qkv[:, D:2 * D], qkv[:, 2 * D:], att, g.seqs, H, 64, L, L, g.qmap, g.kvmap, 0.125), 20)
    ops_out["hivae_attn_tc_same_shape"] = {"ms": round(ms, 4), "tflops": round(fl / ms / 1e9, 1)}
    x16 = torch.randn(M, D, device=dev, dtype=torch.bfloat16)
    for name, N, K in (("qkv", 3 * D, D), ("out", D, D), ("ff1", 4 * D, D), ("ff2", D, 4 * D)):
        xin = torch.randn(M, K, device=dev, dtype=torch.bfloat16)
        w = torch.randn(N, K, device=dev, dtype=torch.bfloat16)
        b = torch.randn(N, device=dev, dtype=torch.bfloat16)
        ms = timeit(lambda: F.linear(xin, w, b), 20)
        ops_out["cublaslt_linear_bf16_%s_%dx%dx%d" % (name, M, N, K)] = {"ms": round(ms, 4), "tflops": round(2.0 * M * N * K / ms / 1e9, 1)}
    x32 = torch.randn(M, D, device=dev)
    w32, b32 = torch.ones(D, device=dev), torch.zeros(D, device=dev)
    ms = timeit(lambda: F.layer_norm(x32, (D,), w32, b32).to(torch.bfloat16), 20)
    ops_out["aten_layernorm_fp32_to_bf16_%dx%d" % (M, D)] = {"ms": round(ms, 4), "gbs_at_6B_per_element": round(M * D * 6 / ms / 1e6, 1)}
    res = torch.randn(frames, L, D, device=dev)
    o16 = torch.empty(M, D, device=dev, dtype=torch.bfloat16)
    ms = timeit(lambda: ops.ln_modulate(ops.Rows(res, 0, L), w32, b32, 1e-5, o16), 20)
    ops_out["hivae_ln_modulate_same_shape"] = {"ms": round(ms, 4), "gbs_at_6B_per_element": round(M * D * 6 / ms / 1e6, 1)}
    out["ops"] = ops_out
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
