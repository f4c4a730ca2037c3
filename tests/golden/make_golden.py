"""Generates tests/golden/*.safetensors from the UNMODIFIED reference (via oracle/shim)
and asserts that oracle/amd_oracle.py (the restatement that travels to the GPU box)
reproduces it.  Run in the build container only:  python tests/golden/make_golden.py
"""
import os
import sys

import torch
from safetensors.torch import save_file

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import amd_oracle as O  # noqa: E402
from oracle import ref_loader  # noqa: E402
from hivae_b200 import configs, synth  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
torch.set_grad_enabled(False)


def check(name, a, b, tol=2e-5):
    err = float((a - b).abs().max() / (b.abs().max() + 1e-12))
    print("  %-34s max-rel-err oracle vs reference = %.2e" % (name, err))
    assert err < tol, name
    return err


def build_ref_model(ref, cfg, seed):
    torch.manual_seed(0)
    m = ref.model_AMD.AMDModel_New(**cfg).eval()
    sd = synth.synth_state_dict(((k, v.shape) for k, v in m.state_dict().items()), seed)
    m.load_state_dict(sd)
    return m, sd


def main():
    ref = ref_loader.load()
    g = torch.Generator().manual_seed(7)

    # ---------------------------------------------------------------- signal ops
    fx = {}
    x = torch.randn(3, 2, 16, 24, generator=g)
    ll, hl, lh, hh = ref.wavelet.dwt_init(x)
    for n_, a, b in zip(("LL", "HL", "LH", "HH"), (ll, hl, lh, hh), O.haar_dwt2(x)):
        check("dwt " + n_, b, a, 1e-7)
    rec = ref.wavelet.iwt_init(torch.cat([ll, hl, lh, hh], 0))
    check("iwt", O.haar_idwt2(torch.cat([ll, hl, lh, hh], 0)), rec, 1e-7)
    fx.update(dwt_x=x, dwt_LL=ll, dwt_HL=hl, dwt_LH=lh, dwt_HH=hh, iwt_out=rec)
    v = torch.randn(1, 2, 32, 32, 32, generator=g)
    for d in (0.5, 0.6):
        lpf = ref.frequency_utils.gaussian_low_pass_filter([32, 32, 32], d, d)
        check("lpf %.1f" % d, O.gaussian_lpf([32, 32, 32], d, d), lpf, 1e-7)
        lo, hi = ref.frequency_utils.freq_3d_filter(v, lpf)
        olo, ohi = O.freq_split3d(v, lpf)
        check("freq lo %.1f" % d, olo, lo, 1e-6), check("freq hi %.1f" % d, ohi, hi, 1e-6)
        fx["freq_lo_%d" % int(d * 10)] = lo
        fx["freq_hi_%d" % int(d * 10)] = hi
    fx["freq_x"] = v
    v2 = torch.randn(1, 2, 8, 12, 20, generator=g)  # ragged even sizes, anisotropic cut-offs
    lpf2 = ref.frequency_utils.gaussian_low_pass_filter([8, 12, 20], 0.3, 0.7)
    lo2, hi2 = ref.frequency_utils.freq_3d_filter(v2, lpf2)
    fx.update(freq2_x=v2, freq2_lo=lo2, freq2_hi=hi2)
    # rectified flow
    rf = ref.rectified_flow.RectifiedFlow(1000)
    z1, z0 = torch.randn(6, 4, 8, 8, generator=g), torch.randn(6, 4, 8, 8, generator=g)
    ts = torch.tensor([0., 1000., 250., 500., 999., 37.])
    zt, tgt = rf.get_train_tuple(z0=z0, z1=z1, time_step=ts)
    ozt, otg = O.rf_train_tuple(z1, ts, z0)
    check("rf zt", ozt, zt, 1e-7), check("rf target", otg, tgt, 1e-7)
    rec_t = rf.get_target_with_zt_vel(zt, tgt, ts)
    check("rf rec", O.rf_target(zt, tgt, ts), rec_t, 1e-7)
    fx.update(rf_z1=z1, rf_z0=z0, rf_ts=ts, rf_zt=zt, rf_target=tgt, rf_rec=rec_t)
    # RoPE attention (model/attention.py)
    att = ref.attention.Attention(128, 40, 4).eval()
    sd_a = synth.synth_state_dict(((k, v_.shape) for k, v_ in att.state_dict().items()), 3)
    att.load_state_dict(sd_a)
    xa = torch.randn(2, 24, 128, generator=g)
    mask = torch.tril(torch.ones(24, 24))
    for tag, kw in (("rope", dict()), ("norope", dict(inference=False)), ("rope_mask", dict(mask=mask, start_pos=3))):
        ya = att(xa, **kw)
        oa = O.rope_attention(xa, sd_a["wq.weight"], sd_a["wq.bias"], sd_a["wk.weight"], sd_a["wk.bias"],
                              sd_a["wv.weight"], sd_a["wv.bias"], 4, 40, **kw)
        check("rope attention " + tag, oa, ya, 1e-5)
        fx["ropeattn_" + tag] = ya
    fx["ropeattn_x"] = xa
    save_file({k: v_.contiguous() for k, v_ in fx.items()}, os.path.join(OUT, "signal_ops.safetensors"))

    # ---------------------------------------------------------------- blocks (D=128, 2 heads)
    fb = {}
    M = ref.modules
    D, H, E = 128, 2, 512
    blocks = {
        "basic": M.BasicTransformerBlock(D, H, 64, attention_bias=True),
        "cross": M.BasicCrossTransformerBlock(D, D, H, 64, attention_bias=True),
        "joint": M.AMDTransformerBlock(D, H, 64, E, attention_bias=True),
        "dit": M.BasicDiTBlock(D, H, 64, E, attention_bias=True),
    }
    xh = torch.randn(3, 10, D, generator=g)
    xe = torch.randn(3, 37, D, generator=g)
    te = torch.randn(3, E, generator=g)
    fb.update(x_hid=xh, x_enc=xe, temb=te)
    for name, blk in blocks.items():
        blk.eval()
        sdb = synth.synth_state_dict(((k, v_.shape) for k, v_ in blk.state_dict().items()), 11)
        blk.load_state_dict(sdb)
        sdp = {"b." + k: v_ for k, v_ in sdb.items()}
        if name == "basic":
            y = blk(xe)
            check("block basic", O.basic_block(sdp, "b", xe, H), y)
            fb["basic_out"] = y
        elif name == "cross":
            y = blk(xh, xe)
            check("block cross", O.basic_block(sdp, "b", xh, H, enc=xe), y)
            fb["cross_out"] = y
        elif name == "joint":
            yh, ye = blk(xh, te, xe)
            oh, oe = O.joint_block(sdp, "b", xh, xe, te, H)
            check("block joint hid", oh, yh), check("block joint enc", oe, ye)
            fb["joint_hid"], fb["joint_enc"] = yh, ye
        else:
            y = blk(xe, te)
            check("block dit", O.dit_block(sdp, "b", xe, te, H), y)
            fb["dit_out"] = y
    save_file({k: v_.contiguous() for k, v_ in fb.items()}, os.path.join(OUT, "blocks.safetensors"))

    # ---------------------------------------------------------------- small model (config 1)
    cfg = configs.SMALL
    m, sd = build_ref_model(ref, cfg, seed=0)
    z = synth.synth_clip_latents(1, seed=1)
    video, ref_img = synth.split_clip(z)
    z0 = torch.randn(16, 4, 32, 32, generator=torch.Generator().manual_seed(2))
    orc = O.AMDOracle(sd, cfg["enc_nhead"], cfg["diffusion_attn_num_heads"], 16, 2, cfg["motion_token_num"])
    cam, osrc, otgt = m.encode(video, ref_img)
    ocam, oosrc, ootgt = orc.encode(video, ref_img)
    check("small encode camera", ocam, cam), check("small encode obj src", oosrc, osrc), check("small encode obj tgt", ootgt, otgt)
    # reference sample draws its noise with randn_like -> patch it in through the scheduler's z0 argument
    orig = m.scheduler.get_train_tuple
    m.scheduler.get_train_tuple = lambda z0_=None, z1=None, time_step=None: orig(z0=z0, z1=z1, time_step=time_step)
    meta = m.sample(video, ref_img, sample_step=4, return_meta_info=True)
    zi_d, zt_d, zj_d = m.decode(video, ref_img, cam, osrc, otgt, start_step=1000, sample_step=4)
    check("ref decode == ref sample", zt_d, meta["sample"], 1e-7)
    trace = []
    ozi, ozt, ozj = orc.sample(video, ref_img, sample_step=4, z0=z0, trace=trace)
    for i, (a, b) in enumerate(zip(trace, meta["pre_cache"])):
        check("small sample pre[%d]" % i, a, b)
    check("small sample zt", ozt, meta["sample"])
    # training forward with explicit time_step (per-clip step repeated over frames) and noise
    ts = torch.tensor([437.0]).repeat_interleave(16)
    m.prepare_timestep = lambda batch_size, device, time_step=None: torch.tensor([437.0])[:batch_size].to(device) if batch_size == 1 else ts
    pre, vel, ld = m(video, ref_img)
    opre, ovel, old = orc.forward(video, ref_img, ts, z0)
    check("small forward pre", opre, pre), check("small forward vel", ovel, vel)
    check("small forward loss", old["loss"], ld["loss"]), check("small forward rec_loss", old["rec_loss"], ld["rec_loss"])
    save_file({k_: v_.contiguous() for k_, v_ in {
        "z0": z0, "camera": cam, "obj_src": osrc, "obj_tgt": otgt,
        **{"pre_%d" % i: p for i, p in enumerate(meta["pre_cache"])},
        "zt": meta["sample"].contiguous(), "fwd_pre": pre, "fwd_vel": vel,
        "fwd_loss": torch.stack([ld["loss"], ld["diff_loss"], ld["rec_loss"]]),
    }.items()}, os.path.join(OUT, "small_model.safetensors"))
    print("  small PSNR(zt,zj) = %.3f dB" % O.psnr(meta["sample"], meta["zj"]))

    # ---------------------------------------------------------------- default model: encode + one DiT forward, fp32
    if os.environ.get("HIVAE_GOLDEN_DEFAULT", "1") == "1":
        cfg = configs.DEFAULT
        m, sd = build_ref_model(ref, cfg, seed=0)
        orc = O.AMDOracle(sd, cfg["enc_nhead"], cfg["diffusion_attn_num_heads"], 16, 2, cfg["motion_token_num"])
        import json
        with open(os.path.join(OUT, "state_dict_keys_default.json"), "w") as f:
            json.dump({k: list(v_.shape) for k, v_ in m.state_dict().items()}, f, indent=0)
        cam, osrc, otgt = m.encode(video, ref_img)
        ocam, oosrc, ootgt = orc.encode(video, ref_img)
        check("default encode camera", ocam, cam, 1e-4), check("default encode obj", ootgt, otgt, 1e-4)
        zi, zj = ref_img.flatten(0, 1), video.flatten(0, 1)
        step = torch.ones(16) * 600
        zt, _ = O.rf_train_tuple(zj, step, z0)
        pre = m.diffusion_transformer(object_motion_source_hidden_states=osrc, object_motion_target_hidden_states=otgt,
                                      camera_motion_target_hidden_states=cam, image_hidden_states=torch.cat([zi, zt], 1),
                                      timestep=step)
        check("default dit forward", orc.dit(zi, zt, step, cam, osrc, otgt), pre, 1e-4)
        save_file({k_: v_.contiguous() for k_, v_ in
                   {"camera": cam, "obj_src": osrc, "obj_tgt": otgt, "dit_pre_600": pre}.items()},
                  os.path.join(OUT, "default_model.safetensors"))
    print("golden fixtures written to", OUT)


if __name__ == "__main__":
    main()
