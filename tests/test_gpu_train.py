"""GPU tier (-m gpu): the training path (BASELINE config 5) against
  * tests/golden/train_step_small.safetensors — one step of train_amd.py:459-496 run by the UNMODIFIED reference on CPU
    (loss, global gradient norm, the norm of every parameter gradient, 31 named gradients in full, their AdamW deltas),
  * the oracle restatement under torch autograd (checked against the reference's autograd when the fixture was made),
  * tests/golden/small_masked.safetensors — random_masking with the reference's noise.

Tolerances: fp32 mode 1e-4, bf16 mode 2e-2 — max-abs error over max-abs of the reference tensor, floored at 1e-5 of the
global gradient norm for gradients that are mathematically zero (norm_k.bias: softmax is invariant to a constant added to
every key), with the RMS-relative figure printed beside it.
"""
import json
import os

import pytest
import torch

from oracle import amd_oracle as O
from hivae_b200 import configs, synth

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TOL = {torch.float32: 1e-4, torch.bfloat16: 2e-2}


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available(), "the gpu tier needs a CUDA device"
    from hivae_b200 import _lib
    _lib.require_device()
    return torch.device("cuda:0")


def _small(dev, dtype, cfg=None):
    import hivae_b200 as hv
    m = hv.AMDModel_New(**(cfg or configs.SMALL))
    sd = synth.synth_state_dict(((k, v.shape) for k, v in m.state_dict().items()), 0)
    m.load_state_dict(sd)
    return m.to(dev).set_compute_dtype(dtype), sd


def _fix(m, z0, steps, n):
    ts = steps.repeat_interleave(16)
    orig = m.scheduler.get_train_tuple
    m.scheduler.get_train_tuple = lambda z0_=None, z1=None, time_step=None, **kw: orig(z0=z0.to(z1.device), z1=z1, time_step=time_step)
    m.prepare_timestep = lambda batch_size, device, time_step=None: (steps if batch_size == n else ts).to(device)


def _err(a, b, floor=0.0):
    a, b = a.detach().float().cpu().reshape(-1), b.detach().float().cpu().reshape(-1)
    mx = float((a - b).abs().max() / max(float(b.abs().max()), floor, 1e-30))
    rms = float((a - b).norm() / max(float(b.norm()), floor, 1e-30))
    return mx, rms


def _report(rows, tol, what):
    rows.sort(key=lambda r: -r[1])
    print("\n%s: worst of %d tensors (max-abs rel, rms rel)" % (what, len(rows)))
    for name, mx, rms in rows[:8]:
        print("  %-72s %.2e %.2e" % (name, mx, rms))
    bad = [r for r in rows if not r[1] <= tol]
    assert not bad, "%s: %d tensors above %.0e, worst %s" % (what, len(bad), tol, bad[0])


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_train_step_golden(dev, golden, dtype):
    """loss.backward() through the differentiable AMDModel_New.forward == the reference's autograd (SURVEY 8c golden #6)"""
    g = golden("train_step_small")
    names = json.load(open(os.path.join(GOLDEN, "train_step_small_names.json")))
    tol = TOL[dtype]
    m, _ = _small(dev, dtype)
    assert sorted(k for k, _ in m.named_parameters()) == sorted(names)  # registration order differs, the key set is the contract
    n = g["steps"].numel()
    video, ref_img = synth.split_clip(synth.synth_clip_latents(n, seed=1))
    _fix(m, g["z0"], g["steps"], n)
    with torch.enable_grad():
        pre, vel, ld = m(video.to(dev), ref_img.to(dev))
        assert pre.requires_grad and ld["loss"].requires_grad
        ld["loss"].backward()
    assert _err(pre, g["pre"])[0] <= tol and _err(vel, g["vel"])[0] <= 1e-6
    assert abs(float(ld["loss"]) - float(g["loss"][0])) <= tol * float(g["loss"][0])
    assert abs(float(ld["rec_loss"]) - float(g["loss"][2])) <= tol * float(g["loss"][2])
    gn = float(g["grad_norm"])
    grads = {k: p.grad for k, p in m.named_parameters()}
    assert all(v is not None for v in grads.values())
    total = torch.sqrt(sum((v.double() ** 2).sum() for v in grads.values()))
    print("\nloss %.6f (ref %.6f)  grad-norm %.6f (ref %.6f)" % (float(ld["loss"]), float(g["loss"][0]), float(total), gn))
    assert abs(float(total) - gn) <= tol * gn
    # the norm of EVERY parameter gradient, then 31 gradients element by element
    rows = []
    for k, ref in zip(names, g["grad_norms_all"]):
        rows.append((k, abs(float(grads[k].norm()) - float(ref)) / max(float(ref), 1e-3 * gn), 0.0))
    _report(rows, tol, "per-parameter gradient norms (%s)" % dtype)
    rows = [(k[5:],) + _err(grads[k[5:]], v, 1e-5 * gn) for k, v in g.items() if k.startswith("grad.")]
    _report(rows, tol, "named gradients (%s)" % dtype)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_fused_train_step_golden(dev, golden, dtype):
    """optim.TrainStep (forward -> backward into the flat arena -> clip -> AdamW) == the reference's step: the parameter
    deltas after one step, and a second step still sees the updated weights (prepared copies are refreshed)"""
    from hivae_b200 import optim
    g = golden("train_step_small")
    tol = TOL[dtype]
    m, sd = _small(dev, dtype)
    n = g["steps"].numel()
    video, ref_img = synth.split_clip(synth.synth_clip_latents(n, seed=1))
    video, ref_img = video.to(dev), ref_img.to(dev)
    ts = g["steps"].repeat_interleave(16).to(dev)
    step = optim.TrainStep(m, lr=1e-4, max_grad_norm=1.0, distributed=False)
    out = step.step(video, ref_img, time_step=ts, noise=g["z0"].to(dev))
    gn = float(g["grad_norm"])
    assert abs(float(out["loss"]) - float(g["loss"][0])) <= tol * float(g["loss"][0])
    assert abs(float(out["grad_norm"]) - gn) <= tol * gn
    # AdamW's first step is -lr * g / (|g| + eps) - lr * wd * p.  Where |g| ~ eps = 1e-8 (clipped gradients of the small
    # layers) the step is proportional to g, so it inherits the gradient tolerance (1e-4 / 2e-2); where |g| >> eps it is
    # -lr * sign(g).  Its direction is ill-conditioned for near-zero elements, so the deltas are compared element by element
    # where the reference gradient is at least 5 % of the tensor's largest; one ulp of the parameter is allowed on top
    # (the delta is a difference of two fp32 parameters).  fp32 also checks the rms over all elements.
    cur = dict(m.named_parameters())
    rows = []
    for k, v in g.items():
        if k.startswith("delta."):
            name = k[6:]
            d = cur[name].detach().cpu() - sd[name]
            gr = g["grad." + name]
            mask = gr.abs() > 0.05 * gr.abs().max()
            ulp = 1.2e-7 * float(sd[name].abs().max())
            mx = max(0.0, float((d - v)[mask].abs().max()) - ulp) / float(v[mask].abs().max())
            rows.append((name, mx, _err(d, v)[1]))
    _report(list(rows), 2e-4 if dtype == torch.float32 else 2e-2, "AdamW parameter deltas where |g| > 5 %% of max (%s)" % dtype)
    if dtype == torch.float32:
        _report([(k, rms, rms) for k, mx, rms in rows], 2e-3, "AdamW parameter deltas, rms over all elements (fp32)")
    # second step: it must see the UPDATED weights (stale prepared copies would reproduce the first loss exactly), and the
    # inference forward on the weights after two steps must agree with the oracle evaluated on those weights
    out2 = step.step(video, ref_img, time_step=ts, noise=g["z0"].to(dev))
    assert float(out2["loss"]) != float(out["loss"])
    sd2 = {k: v.detach().cpu().clone() for k, v in m.state_dict().items()}
    orc = O.AMDOracle(sd2, configs.SMALL["enc_nhead"], configs.SMALL["diffusion_attn_num_heads"], 16, 2, configs.SMALL["motion_token_num"])
    _fix(m, g["z0"], g["steps"], n)
    with torch.no_grad():
        pre, vel, ld = m(video, ref_img)
        opre, ovel, old = orc.forward(video.cpu(), ref_img.cpu(), ts.cpu(), g["z0"])
    assert _err(pre, opre)[0] <= tol, "forward after two optimizer steps runs on stale prepared weights"


def test_train_vs_oracle_autograd_other_batch(dev):
    """a different batch / steps / noise than the fixture, fp32 mode, checked against the oracle's autograd on the host"""
    m, sd = _small(dev, torch.float32)
    n = 1
    video, ref_img = synth.split_clip(synth.synth_clip_latents(n, seed=5))
    z0 = torch.randn(16, 4, 32, 32, generator=torch.Generator().manual_seed(9))
    steps = torch.tensor([250.0])
    _fix(m, z0, steps, n)
    with torch.enable_grad():
        pre, vel, ld = m(video.to(dev), ref_img.to(dev))
        ld["loss"].backward()
        sdg = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
        orc = O.AMDOracle(sdg, configs.SMALL["enc_nhead"], configs.SMALL["diffusion_attn_num_heads"], 16, 2, configs.SMALL["motion_token_num"])
        opre, ovel, old = orc.forward(video, ref_img, steps.repeat_interleave(16), z0)
        old["loss"].backward()
    gn = float(torch.sqrt(sum((v.grad.double() ** 2).sum() for v in sdg.values())))
    rows = [(k,) + _err(p.grad, sdg[k].grad, 1e-5 * gn) for k, p in m.named_parameters()]
    _report(rows, 1e-4, "all gradients vs oracle autograd (fp32)")


def test_frozen_camera_encoder(dev, golden):
    """--frozen_name camera_motion_encoder (scripts/amd_new.sh:66): frozen parameters get no gradient, camera_down behind
    the frozen encoder still does, and every other gradient is unchanged"""
    g = golden("train_step_small")
    m, _ = _small(dev, torch.float32)
    for k, p in m.named_parameters():
        if "camera_motion_encoder" in k:
            p.requires_grad_(False)
    n = g["steps"].numel()
    video, ref_img = synth.split_clip(synth.synth_clip_latents(n, seed=1))
    _fix(m, g["z0"], g["steps"], n)
    with torch.enable_grad():
        pre, vel, ld = m(video.to(dev), ref_img.to(dev))
        ld["loss"].backward()
    gn = float(g["grad_norm"])
    for k, p in m.named_parameters():
        if "camera_motion_encoder" in k:
            assert p.grad is None
        elif "grad." + k in g:
            assert _err(p.grad, g["grad." + k], 1e-5 * gn)[0] <= 1e-4, k


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_random_masking_golden(dev, golden, dtype, monkeypatch):
    """camera_mask_ratio / object_mask_ratio (train_amd.py:459, transformer.py:164-192, 558-586) with the reference's noise:
    forward() in both modes (no_grad and differentiable) and encode()"""
    from hivae_b200 import transformer as T
    g = golden("small_masked")
    tol = TOL[dtype]
    m, _ = _small(dev, dtype)
    video, ref_img = synth.split_clip(synth.synth_clip_latents(2, seed=1))
    video, ref_img = video.to(dev), ref_img.to(dev)
    _fix(m, g["z0"], g["steps"], 2)

    def feed(*noises):
        it = iter(noises)
        monkeypatch.setattr(T, "mask_noise", lambda N, L, device: next(it).to(device))

    feed(g["enc_cam_noise"], g["enc_obj_noise"])
    cam, osrc, otgt = m.encode(video, ref_img, camera_mask_ratio=0.5, object_mask_ratio=0.8)
    assert cam.shape == g["enc_camera"].shape
    assert _err(cam, g["enc_camera"])[0] <= tol and _err(osrc, g["enc_obj_src"])[0] <= tol and _err(otgt, g["enc_obj_tgt"])[0] <= tol
    for grad_mode in (False, True):
        feed(g["cam_noise"], g["obj_noise"])
        torch.manual_seed(1234)  # the two ratio draws of forward() come from the CPU generator
        with torch.set_grad_enabled(grad_mode):
            pre, vel, ld = m(video, ref_img, camera_mask_ratio=0.5, object_mask_ratio=0.8)
            assert pre.requires_grad == grad_mode
            if grad_mode:
                ld["loss"].backward()
                assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in m.parameters())
        assert _err(pre, g["pre"])[0] <= tol, grad_mode
        assert abs(float(ld["loss"]) - float(g["loss"][0])) <= tol * float(g["loss"][0])


def test_temporal_encoder_random_masking(dev, monkeypatch):
    """MotionEncoderLearnTokenTemporalTransformer.random_masking (transformer.py:369-397; AMDModel's encoder) against the oracle
    with the same noise"""
    import hivae_b200 as hv
    from hivae_b200 import transformer as T
    kw = dict(img_height=32, img_width=32, img_inchannel=4, img_patch_size=2, motion_token_num=4, motion_channel=16,
              need_norm_out=True, video_frames=4, num_attention_heads=2, attention_head_dim=64, num_layers=2)
    m = hv.MotionEncoderLearnTokenTemporalTransformer(**kw)
    sd = synth.synth_state_dict(((k, v.shape) for k, v in m.state_dict().items()), 3)
    m.load_state_dict(sd)
    m.to(dev)
    x = torch.randn(2, 8, 4, 32, 32, generator=torch.Generator().manual_seed(12))
    noise = torch.rand(16, 256, generator=torch.Generator().manual_seed(13))
    monkeypatch.setattr(T, "mask_noise", lambda N, L, device: noise.to(device))
    y = m(x.to(dev), mask_ratio=0.6)
    ref = O.temporal_motion_encoder({"e." + k: v for k, v in sd.items()}, "e", x, 2, 4, keep=O.mask_keep(noise, 0.6))
    assert y.shape == ref.shape and _err(y, ref)[0] <= 1e-4


def test_linear_grads_odd_token_count(dev):
    """a token count that is not a multiple of 8: the contraction dimension of the weight-gradient GEMM is
    zero-padded; bf16 operands, fp64 reference"""
    from hivae_b200 import backward as B
    from hivae_b200.engine import Ctx
    ctx = Ctx(dev, torch.bfloat16)
    gen = torch.Generator().manual_seed(3)
    for M, N, K in ((274, 256, 128), (3 * 18 + 1, 192, 64)):
        x = torch.randn(M, K, generator=gen).to(dev).bfloat16()
        w = (torch.randn(N, K, generator=gen) / K ** 0.5).to(dev).bfloat16()
        dy = torch.randn(M, N, generator=gen).to(dev).bfloat16()
        dw = torch.zeros(N, K, device=dev)
        db = torch.zeros(N, device=dev)
        dx = B.linear_grads(ctx, dy, x, w, dw, db)
        xd, wd, dyd = x.double(), w.double(), dy.double()
        assert _err(dx, dyd @ wd)[0] <= 1e-2 and _err(dw, dyd.t() @ xd)[0] <= 1e-2 and _err(db, dyd.sum(0))[0] <= 1e-2
        B.linear_grads(ctx, dy, x, w, dw, db)  # accumulates
        assert _err(dw, 2 * (dyd.t() @ xd))[0] <= 1e-2


def test_linear_tn_weight_gradient_gemm(dev):
    """hv_linear_bf16_tn: out (+)= At^T Bt on MN-major tcgen05 operands (dW = dY^T X without transposed copies) against fp64:
    single CTA and CTA-pair tiles, ragged token counts (TMA zero fill), column-sliced operands, split-K accumulation"""
    from hivae_b200 import ops
    gen = torch.Generator().manual_seed(11)
    for K, M, N, ks in ((1000, 256, 128, 1), (4099, 512, 384, 4), (777, 1024, 1024, 3), (64, 128, 64, 1), (35072 // 8, 3072, 1024, 8)):
        At = torch.randn(K, M, generator=gen).to(dev).bfloat16()
        Bt = torch.randn(K, N, generator=gen).to(dev).bfloat16()
        ref = At.double().t() @ Bt.double()
        out = torch.empty(M, N, device=dev)
        ops.linear_tn(At, Bt, ops.epi_store(out))
        assert _err(out, ref)[0] <= 2e-3, (K, M, N)
        acc = torch.ones(M, N, device=dev)
        e = ops.epi_residual(ops.Rows(acc.view(1, M, N), 0, M))
        e.k_split = ks
        ops.linear_tn(At, Bt, e)
        assert _err(acc, ref + 1.0)[0] <= 2e-3, (K, M, N, ks)
    wide = torch.randn(500, 768, generator=gen).to(dev).bfloat16()   # a column slice as At (the q / k / v blocks of d(qkv))
    Bt = torch.randn(500, 256, generator=gen).to(dev).bfloat16()
    out = torch.empty(256, 256, device=dev)
    ops.linear_tn(wide[:, 256:512], Bt, ops.epi_store(out))
    assert _err(out, wide[:, 256:512].double().t() @ Bt.double())[0] <= 2e-3


def test_block_forward_after_adamw_step_uses_new_weights(dev):
    """FlatArena + AdamW update the parameters through raw pointers; the prepared bf16 / packed copies of
    a block must follow"""
    import hivae_b200 as hv
    from hivae_b200 import optim
    torch.manual_seed(0)
    blk = hv.modules.BasicTransformerBlock(128, 2, 64, attention_bias=True).to(dev)
    blk.hv_compute_dtype = torch.bfloat16
    x = torch.randn(2, 24, 128, device=dev)
    y0 = blk(x).clone()
    opt = optim.AdamW(blk.parameters(), lr=1e-1)
    y1 = blk(x).clone()           # the arena re-homed every parameter
    assert torch.equal(y0, y1)
    opt.arena.grad.normal_()
    optim.clip_grad_norm_(opt.arena, 1.0)
    opt.step()
    y2 = blk(x)
    sd = {"b." + k: v.detach().cpu() for k, v in blk.state_dict().items()}
    ref = O.basic_block(sd, "b", x.cpu(), 2)
    assert _err(y2, ref)[0] <= 3e-2 and _err(y2, y0)[0] > 5e-2
    # a gradient re-bound behind the arena's back is picked up
    p = next(blk.parameters())
    before = p.detach().clone()
    blk.zero_grad(set_to_none=True)
    for q in blk.parameters():
        q.grad = torch.ones_like(q)
    opt.step()
    assert not torch.equal(before, p.detach())


def test_dual_output_epilogue_matches_two_launches(dev):
    """the training forward's dual-output GEMM epilogue (pre-activation + activation from ONE tcgen05 GEMM) is bit-identical
    to the two separate launches it replaces; bf16 TMA-store path, the fallback shapes and the fp32 SIMT path"""
    from hivae_b200 import ops
    gen = torch.Generator().manual_seed(5)
    for dtype, M, D in ((torch.bfloat16, 1000, 256), (torch.bfloat16, 300, 128), (torch.float32, 130, 128)):
        x = torch.randn(M, D, generator=gen).to(dev).to(dtype)
        w = (torch.randn(3 * D, D, generator=gen) / D ** 0.5).to(dev).to(dtype)
        b = torch.randn(3 * D, generator=gen).to(dev)
        nw, nb = (1 + 0.1 * torch.randn(64, generator=gen)).to(dev), (0.1 * torch.randn(64, generator=gen)).to(dev)
        new = lambda: torch.empty(M, 3 * D, dtype=dtype, device=dev)
        a1, p1, a2, p2 = new(), new(), new(), new()
        ops.linear(x, w, ops.epi_qknorm(a1, b, nw, nb, nw, nb, D, D, pre=p1))
        ops.linear(x, w, ops.epi_qknorm(a2, b, nw, nb, nw, nb, D, D))
        ops.linear(x, w, ops.epi_store(p2, b))
        assert torch.equal(a1, a2) and torch.equal(p1, p2), ("qknorm", dtype, M)
        ops.linear(x, w, ops.epi_store(a1, b, gelu=True, pre=p1))
        ops.linear(x, w, ops.epi_store(a2, b, gelu=True))
        assert torch.equal(a1, a2) and torch.equal(p1, p2), ("gelu", dtype, M)


def test_pipeline_windowed_sample(dev):
    """AMDPipeLine_single on VAE latents (model/pipeline.py:647-751): one window, and a 33-frame latent sequence cut into two
    windows reconstructed in one batched sample call == the oracle's sample of each window"""
    import hivae_b200 as hv
    m, sd = _small(dev, torch.float32)
    pipe = hv.AMDPipeLine_single(None, m, window=16, device=dev)
    z = synth.synth_clip_latents(1, frames=33, seed=4)[0]
    z0 = torch.randn(32, 4, 32, 32, generator=torch.Generator().manual_seed(6))
    orig = m.scheduler.get_train_tuple
    m.scheduler.get_train_tuple = lambda z0_=None, z1=None, time_step=None: orig(z0=z0[:z1.shape[0]].to(z1.device), z1=z1, time_step=time_step)
    out = pipe.sample_long(z, video_sample_step=3)
    assert out.shape == (1, 33, 4, 32, 32) and torch.equal(out[0, 0].cpu(), z[0])
    orc = O.AMDOracle(sd, configs.SMALL["enc_nhead"], configs.SMALL["diffusion_attn_num_heads"], 16, 2, configs.SMALL["motion_token_num"])
    for k in range(2):
        video = z[1 + 16 * k:17 + 16 * k].unsqueeze(0)
        ref = z[16 * k:16 * k + 1].unsqueeze(0).repeat(1, 16, 1, 1, 1)
        _, ozt, _ = orc.sample(video, ref, sample_step=3, z0=z0[16 * k:16 * k + 16])
        assert _err(out[0, 1 + 16 * k:17 + 16 * k], ozt.reshape(16, 4, 32, 32))[0] <= 1e-4, k
    one = pipe.sample_latents(z[:17], video_sample_step=3)
    assert _err(one[0, 1:], out[0, 1:17])[0] <= 1e-5


def test_graphed_train_step_equals_eager(dev, golden):
    """TrainStep.step_graphed (the whole step replayed from one CUDA graph, AdamW step count / lr in device memory) ==
    TrainStep.step, over three steps on changing batches; capturing leaves weights, moments and step count untouched; with
    torch's generator the time steps / noise change from replay to replay (nothing random is frozen into the graph)."""
    from hivae_b200 import optim
    g = golden("train_step_small")
    n = g["steps"].numel()
    batches = []
    for seed in (1, 2, 3):
        v, r = synth.split_clip(synth.synth_clip_latents(n, seed=seed))
        batches.append((v.to(dev), r.to(dev)))
    runs = {}
    for mode in ("eager", "graph"):
        m, sd = _small(dev, torch.float32)
        _fix(m, g["z0"].to(dev), g["steps"], n)  # fixed time steps and noise (already on the device): both modes see the same draws
        step = optim.TrainStep(m, lr=1e-4, max_grad_norm=1.0, distributed=False)
        losses = []
        if mode == "graph":
            step.capture(*batches[0])
            for k, p in m.named_parameters():
                assert torch.equal(p.detach().cpu(), sd[k]), "capture changed " + k
            assert step.opt.step_count == 0 and float(step.opt._state[0]) == 0.0
        for i, (v, r) in enumerate(batches):
            if i == 2:
                step.opt.set_lr(5e-5)  # a learning-rate change must reach the replayed graph
            out = step.step_graphed(v, r) if mode == "graph" else step.step(v, r)
            losses.append((float(out["loss"]), float(out["grad_norm"])))
        assert step.opt.step_count == 3
        runs[mode] = (losses, {k: p.detach().cpu().clone() for k, p in m.named_parameters()}, sd)
    (le, pe, sd), (lg, pg, _) = runs["eager"], runs["graph"]
    for (a, b), (c, d) in zip(le, lg):
        assert abs(a - c) <= 1e-5 * abs(a) and abs(b - d) <= 1e-4 * abs(b), (le, lg)
    assert len({round(x[0], 6) for x in le}) == 3  # the three batches give three different losses
    rows = []
    for k in pe:
        de, dg = pe[k] - sd[k], pg[k] - sd[k]
        rows.append((k, _err(dg, de)[1], _err(dg, de)[1]))
    _report(rows, 2e-3, "parameter deltas after three steps, graph replay vs eager (rms)")
    # real random draws: the replays must not repeat the capture-time time steps / noise
    m, _ = _small(dev, torch.float32)
    step = optim.TrainStep(m, lr=1e-4, max_grad_norm=1.0, distributed=False)
    torch.manual_seed(11)
    seen = [float(step.step_graphed(*batches[0])["loss"]) for _ in range(3)]
    assert len({round(x, 5) for x in seen}) == 3, seen


def test_sample_graphed_equals_sample(dev):
    """AMDModel_New.sample_graphed (encode + Euler loop replayed from one CUDA graph) == sample on the same generator state,
    on two different batches through the same graph; a weight change invalidates the graph"""
    from hivae_b200 import engine
    m, _ = _small(dev, torch.bfloat16)
    outs = {}
    for mode in ("eager", "graph", "graph"):
        res = []
        for seed in (1, 2):
            v, r = synth.split_clip(synth.synth_clip_latents(2, seed=seed))
            v, r = v.to(dev), r.to(dev)
            torch.manual_seed(100 + seed)
            fn = m.sample_graphed if mode == "graph" else m.sample
            with torch.no_grad():
                res.append(fn(v, r, sample_step=4)[1].float().cpu().clone())
        outs.setdefault(mode, []).append(res)
    e = outs["eager"][0]
    for res in outs["graph"]:
        for a, b in zip(res, e):
            assert torch.equal(a, b)
    assert not torch.equal(e[0], e[1])
    assert len(m._sample_graphs) == 1
    engine.bump_weight_epoch()
    v, r = synth.split_clip(synth.synth_clip_latents(2, seed=1))
    torch.manual_seed(101)
    with torch.no_grad():
        again = m.sample_graphed(v.to(dev), r.to(dev), sample_step=4)[1].float().cpu()
    assert len(m._sample_graphs) == 1 and torch.equal(again, e[0])
