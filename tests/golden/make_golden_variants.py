"""Golden fixture for the sibling denoiser: AMDModel_New(diffusion_model_type='default', use_camera=False) =
AMDDiffusionTransformerModelTempMotion (transformer.py:1483-1743), from the UNMODIFIED reference via oracle/shim.
Run in the build container only:  python tests/golden/make_golden_variants.py"""
import os
import sys

import torch
from safetensors.torch import save_file

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from make_golden import OUT, build_ref_model, check  # noqa: E402
from oracle import amd_oracle as O  # noqa: E402
from oracle import ref_loader  # noqa: E402
from hivae_b200 import configs, synth  # noqa: E402

torch.set_grad_enabled(False)


def main():
    ref = ref_loader.load()
    cfg = configs.SMALL_OBJECT_ONLY
    m, sd = build_ref_model(ref, cfg, seed=0)
    assert type(m.diffusion_transformer).__name__ == "AMDDiffusionTransformerModelTempMotion"
    video, ref_img = synth.split_clip(synth.synth_clip_latents(1, seed=1))
    z0 = torch.randn(16, 4, 32, 32, generator=torch.Generator().manual_seed(2))
    orc = O.AMDOracle(sd, cfg["enc_nhead"], cfg["diffusion_attn_num_heads"], 16, 2, cfg["motion_token_num"])
    orig = m.scheduler.get_train_tuple
    m.scheduler.get_train_tuple = lambda z0_=None, z1=None, time_step=None: orig(z0=z0, z1=z1, time_step=time_step)
    meta = m.sample(video, ref_img, sample_step=3, return_meta_info=True)
    trace = []
    _, ozt, _ = orc.sample(video, ref_img, sample_step=3, z0=z0, trace=trace)
    for i, (a, b) in enumerate(zip(trace, meta["pre_cache"])):
        check("object-only sample pre[%d]" % i, a, b)
    check("object-only sample zt", ozt, meta["sample"])
    # training forward: 'default' keeps the per-FRAME timesteps it is given (no per-clip redraw, model_AMD.py:1962)
    ts = torch.arange(16).float() * 60.0 + 13.0
    m.prepare_timestep = lambda batch_size, device, time_step=None: ts.to(device)
    pre, vel, ld = m(video, ref_img)
    opre, ovel, old = orc.forward(video, ref_img, ts, z0)
    check("object-only forward pre", opre, pre), check("object-only forward loss", old["loss"], ld["loss"])
    save_file({k: v.contiguous() for k, v in {
        "z0": z0, "ts": ts, **{"pre_%d" % i: p for i, p in enumerate(meta["pre_cache"])}, "zt": meta["sample"].contiguous(),
        "fwd_pre": pre, "fwd_loss": torch.stack([ld["loss"], ld["diff_loss"], ld["rec_loss"]])}.items()},
        os.path.join(OUT, "small_object_only.safetensors"))
    with open(os.path.join(OUT, "state_dict_keys_small_object_only.json"), "w") as f:
        import json
        json.dump({k: list(v.shape) for k, v in m.state_dict().items()}, f, indent=0)
    print("written", os.path.join(OUT, "small_object_only.safetensors"))
    temporal_encoder(ref)
    cross_and_refimg(ref)
    imgspatial_dit(ref)
    amd_model(ref)


def amd_model(ref):
    """a18 sibling AMDModel (model_AMD.py:33-594): forward + sample for two runnable configurations"""
    out = {}
    gen = torch.Generator().manual_seed(41)
    n, t = 2, 4
    video, ref_img, vg, rg = (torch.randn(n, t, 4, 32, 32, generator=gen) * 0.7 for _ in range(4))
    mask = (torch.rand(n, 2 * t, 4, 32, 32, generator=gen) > 0.3).float()
    z0 = torch.randn(n * t, 4, 32, 32, generator=gen)
    out.update(video=video, ref_img=ref_img, vg=vg, rg=rg, mask=mask, z0=z0)
    for name, cfg in (("temporal", configs.AMDMODEL_TEMPORAL), ("legacy", configs.AMDMODEL_LEGACY)):
        torch.manual_seed(0)
        m = ref.model_AMD.AMDModel(**cfg).eval()
        sd = synth.synth_state_dict(((k, v.shape) for k, v in m.state_dict().items()), 7)
        m.load_state_dict(sd)
        orc = O.AMDModelOracle(sd, cfg)
        orig = m.scheduler.get_train_tuple
        m.scheduler.get_train_tuple = lambda z0_=None, z1=None, time_step=None: orig(z0=z0, z1=z1, time_step=time_step)
        ts_clip = torch.tensor([700.0, 250.0])
        ts = ts_clip.repeat_interleave(t) if cfg["diffusion_model_type"] != "default" else torch.arange(n * t).float() * 100 + 50
        m.prepare_timestep = (lambda batch_size, device, time_step=None: ts_clip) if cfg["diffusion_model_type"] != "default" \
            else (lambda batch_size, device, time_step=None: ts)
        cm = mask if cfg.get("use_mask") else None
        pre, vel, ld = m(video, ref_img, video_grey=vg, ref_img_grey=rg, camera_mask=cm)
        opre, ovel, old = orc.forward(video, ref_img, ts, z0, vg, rg, cm)
        check("AMDModel %s forward pre" % name, opre, pre), check("AMDModel %s loss" % name, old["loss"], ld["loss"])
        zi, zt, zj = m.sample(video, ref_img, video_grey=vg, ref_img_grey=rg, sample_step=2)
        check("AMDModel %s sample zt" % name, orc.sample(video, ref_img, 2, z0, vg, rg), zt)
        out.update({name + "_ts": ts, name + "_pre": pre.contiguous(), name + "_loss": torch.stack([ld["loss"], ld["rec_loss"]]), name + "_zt": zt.contiguous()})
        with open(os.path.join(OUT, "state_dict_keys_amdmodel_%s.json" % name), "w") as f:
            import json
            json.dump({k: list(v.shape) for k, v in m.state_dict().items()}, f, indent=0)
    save_file(out, os.path.join(OUT, "amd_model.safetensors"))


DITS = dict(num_attention_heads=2, attention_head_dim=64, out_channels=4, num_layers=8, image_width=32, image_height=32,
            image_patch_size=2, image_in_channels=8, motion_token_num=4, motion_in_channels=16, motion_target_num_frame=4)


def imgspatial_dit(ref):
    """a17 sibling AMDDiffusionTransformerModelImgSpatial (transformer.py:2044-2356): 'plus', and 'decouple' with and
    without a camera source (8 layers so that the swap after layer 5 is exercised)."""
    g = torch.Generator().manual_seed(21)
    N, L = 8, 4  # two clips of four frames
    image, ts = torch.randn(N, 8, 32, 32, generator=g), torch.tensor([900.0] * 4 + [350.0] * 4)
    ct, cs, os_, ot = (torch.randn(N, L, 16, generator=g) for _ in range(4))
    out = dict(image=image, ts=ts, ct=ct, cs=cs, os=os_, ot=ot)
    for mt in ("plus", "decouple"):
        torch.manual_seed(0)
        m = ref.transformer.AMDDiffusionTransformerModelImgSpatial(motion_type=mt, **DITS).eval()
        sd = synth.synth_state_dict(((k, v.shape) for k, v in m.state_dict().items()), 5)
        m.load_state_dict(sd)
        osd = {"d." + k: v for k, v in sd.items()}
        cases = {"plus": [("obj", None, os_, ot)], "decouple": [("cam", None, None, None), ("cam_obj", None, os_, ot), ("camsrc_obj", cs, os_, ot)]}[mt]
        for name, c_src, o_s, o_t in cases:
            y = m(ct, image, ts, camera_motion_source_hidden_states=c_src, object_motion_source_hidden_states=o_s,
                  object_motion_target_hidden_states=o_t)
            yo = O.dit_imgspatial_forward(osd, "d", ct, image, ts, 2, 4, 2, 4, mt, c_src, o_s, o_t)
            check("ImgSpatial %s/%s" % (mt, name), yo, y)
            out["y_%s_%s" % (mt, name)] = y.contiguous()
    # the oldest sibling, AMDDiffusionTransformerModel (transformer.py:1010-1259): 9 layers so that layers 6, 7 run in
    # both passes and layer 8 only in the object pass
    kw = {k: v for k, v in DITS.items() if k != "motion_target_num_frame"}
    kw["num_layers"] = 9
    for mt in ("plus", "decouple"):
        torch.manual_seed(0)
        m2 = ref.transformer.AMDDiffusionTransformerModel(motion_type=mt, **kw).eval()
        sd = synth.synth_state_dict(((k, v.shape) for k, v in m2.state_dict().items()), 6)
        m2.load_state_dict(sd)
        osd = {"d." + k: v for k, v in sd.items()}
        cases = {"plus": [("camsrc_obj", cs, os_, ot)], "decouple": [("cam_obj", None, os_, ot), ("camsrc_obj", cs, os_, ot)]}[mt]
        for name, c_src, o_s, o_t in cases:
            y = m2(ct, image, ts, camera_motion_source_hidden_states=c_src, object_motion_source_hidden_states=o_s,
                   object_motion_target_hidden_states=o_t)
            yo = O.dit_imgspatial_forward(osd, "d", ct, image, ts, 2, 4, 2, 4, mt, c_src, o_s, o_t, legacy=True)
            check("legacy DiT %s/%s" % (mt, name), yo, y)
            out["legacy_%s_%s" % (mt, name)] = y.contiguous()
    with open(os.path.join(OUT, "state_dict_keys_legacy_dit.json"), "w") as f:
        import json
        json.dump({k: list(v.shape) for k, v in m2.state_dict().items()}, f, indent=0)
    save_file(out, os.path.join(OUT, "imgspatial_dit.safetensors"))
    with open(os.path.join(OUT, "state_dict_keys_imgspatial_dit.json"), "w") as f:
        import json
        json.dump({k: list(v.shape) for k, v in m.state_dict().items()}, f, indent=0)


def cross_and_refimg(ref):
    """a18: AMDModel_New.sample_cross (model_AMD.py:2198-2377) and .sample_with_refimg_motion (2701-2785), small config"""
    cfg = configs.SMALL
    m, sd = build_ref_model(ref, cfg, seed=0)
    orc = O.AMDOracle(sd, cfg["enc_nhead"], cfg["diffusion_attn_num_heads"], 16, 2, cfg["motion_token_num"])
    v1, ref_img = synth.split_clip(synth.synth_clip_latents(1, seed=1))
    v2, _ = synth.split_clip(synth.synth_clip_latents(1, seed=4))
    z0 = torch.randn(16, 4, 32, 32, generator=torch.Generator().manual_seed(2))
    orig = m.scheduler.get_train_tuple
    m.scheduler.get_train_tuple = lambda z0_=None, z1=None, time_step=None: orig(z0=z0, z1=z1, time_step=time_step)
    zi, zt, zj = m.sample_cross(v1, v2, ref_img, sample_step=3)
    ozi, ozt, ozj = orc.sample_cross(v1, v2, ref_img, sample_step=3, z0=z0)
    check("sample_cross zt", ozt, zt), check("sample_cross zj", ozj, zj)
    cam, osrc, otgt = m.encode(v2, ref_img)
    grey = ref_img * 0.5 + v2 * 0.25  # any (n, t, c, h, w) tensor serves as the "grey" reference clip
    ri, rt, rj = m.sample_with_refimg_motion(ref_img, cam, otgt, sample_step=3, ref_img_grey=grey)
    oi, ot, oj = orc.sample_with_refimg_motion(ref_img, cam, otgt, grey, sample_step=3, z0=z0)
    check("refimg_motion zt", ot, rt)
    assert rj.shape == oj.shape == (16, 4, 32, 32)
    save_file({k: v.contiguous() for k, v in dict(z0=z0, cross_zt=zt, cam=cam, otgt=otgt, grey=grey, refimg_zt=rt).items()},
              os.path.join(OUT, "small_cross_refimg.safetensors"))


TENC = dict(img_height=32, img_width=32, img_inchannel=4, img_patch_size=2, motion_token_num=4, motion_channel=16,
            need_norm_out=True, video_frames=4, num_attention_heads=2, attention_head_dim=64, num_layers=2)


def temporal_encoder(ref):
    """a16: MotionEncoderLearnTokenTemporalTransformer (transformer.py:195-367), n = 1 (target = frames T/2..T-1)
    and n = 2 (where the reference's dim-0 chunk regroups the second CLIP instead)."""
    torch.manual_seed(0)
    m = ref.transformer.MotionEncoderLearnTokenTemporalTransformer(**TENC).eval()
    sd = synth.synth_state_dict(((k, v.shape) for k, v in m.state_dict().items()), 3)
    m.load_state_dict(sd)
    out = {}
    for n in (1, 2):
        x = torch.randn(n, 8, 4, 32, 32, generator=torch.Generator().manual_seed(10 + n))
        y = m(x)
        check("temporal encoder n=%d" % n, O.temporal_motion_encoder({"e." + k: v for k, v in sd.items()}, "e", x, 2, 4), y)
        out["x%d" % n], out["y%d" % n] = x, y.contiguous()
    save_file(out, os.path.join(OUT, "temporal_encoder.safetensors"))
    with open(os.path.join(OUT, "state_dict_keys_temporal_encoder.json"), "w") as f:
        import json
        json.dump({k: list(v.shape) for k, v in m.state_dict().items()}, f, indent=0)


def camera_and_rec_models(ref):
    """f3: AMDModel_Camera (model_AMD.py:894-1661) and AMDModel_Rec (2840-3125) in their runnable configurations"""
    import json
    g = torch.Generator().manual_seed(21)
    # ---- AMDModel_Camera, both denoisers
    out = {}
    n, t = 1, 4
    video, ref_img = torch.randn(n, t, 4, 32, 32, generator=g) * 0.7, torch.randn(n, t, 4, 32, 32, generator=g) * 0.7
    vg, rg = torch.randn(n, t, 4, 32, 32, generator=g) * 0.7, torch.randn(n, t, 4, 32, 32, generator=g) * 0.7
    mask = (torch.rand(n, 2 * t, 1, 32, 32, generator=g) > 0.3).float()
    z0 = torch.randn(n * t, 4, 32, 32, generator=g)
    ts = torch.tensor([411.0]).repeat_interleave(t)
    out.update(video=video, ref_img=ref_img, video_grey=vg, ref_img_grey=rg, camera_mask=mask, z0=z0)
    for dmt in ("spatial", "default"):
        cfg = dict(configs.AMDMODEL_CAMERA, diffusion_model_type=dmt)
        torch.manual_seed(0)
        m = ref.model_AMD.AMDModel_Camera(**cfg).eval()
        sd = synth.synth_state_dict(((k, v.shape) for k, v in m.state_dict().items()), 8)
        m.load_state_dict(sd)
        orig = m.scheduler.get_train_tuple
        m.scheduler.get_train_tuple = lambda z0_=None, z1=None, time_step=None, orig=orig: orig(z0=z0, z1=z1, time_step=time_step)
        m.prepare_timestep = lambda batch_size, device, time_step=None: (torch.tensor([411.0]) if batch_size == n else ts).to(device)
        pre, vel, ld = m(video, ref_img, video_grey=vg, ref_img_grey=rg, camera_mask=mask)
        orc = O.AMDModelCameraOracle(sd, cfg)
        opre, ovel, old = orc.forward(video, ref_img, ts, z0, vg, rg, mask)
        check("AMDModel_Camera %s forward pre" % dmt, opre, pre), check("AMDModel_Camera %s loss" % dmt, old["loss"], ld["loss"])
        zi, zt, zj = m.sample(video, ref_img, video_grey=vg, ref_img_grey=rg, camera_mask=mask, sample_step=3)
        ozt, ozj = orc.sample(video, ref_img, 3, z0, vg, rg, mask)
        check("AMDModel_Camera %s sample zt" % dmt, ozt, zt), check("AMDModel_Camera %s zj" % dmt, ozj, zj)
        out.update({"%s_pre" % dmt: pre, "%s_loss" % dmt: torch.stack([ld["loss"], ld["rec_loss"]]), "%s_zt" % dmt: zt.contiguous(),
                    "%s_zj" % dmt: zj.contiguous()})
        with open(os.path.join(OUT, "state_dict_keys_amdmodel_camera_%s.json" % dmt), "w") as f:
            json.dump({k: list(v.shape) for k, v in m.state_dict().items()}, f, indent=0)
    save_file({k: v.contiguous() for k, v in out.items()}, os.path.join(OUT, "amd_model_camera.safetensors"))
    # ---- AMDModel_Rec
    cfg = configs.AMDMODEL_REC
    torch.manual_seed(0)
    m = ref.model_AMD.AMDModel_Rec(**cfg).eval()
    sd = synth.synth_state_dict(((k, v.shape) for k, v in m.state_dict().items()), 9)
    m.load_state_dict(sd)
    pre, zj, ld = m(video, ref_img)
    opre, ozj, old = O.AMDModelRecOracle(sd, cfg).forward(video, ref_img)
    check("AMDModel_Rec forward pre", opre, pre), check("AMDModel_Rec loss", old["loss"], ld["loss"])
    save_file({k: v.contiguous() for k, v in dict(video=video, ref_img=ref_img, pre=pre, loss=ld["loss"].reshape(1)).items()},
              os.path.join(OUT, "amd_model_rec.safetensors"))
    with open(os.path.join(OUT, "state_dict_keys_amdmodel_rec.json"), "w") as f:
        json.dump({k: list(v.shape) for k, v in m.state_dict().items()}, f, indent=0)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "camera_rec":
        torch.set_grad_enabled(False)
        camera_and_rec_models(ref_loader.load())
    else:
        main()
