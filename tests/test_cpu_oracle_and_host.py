"""CPU tier (-m "not gpu"): the oracle against the committed golden vectors (generated from the
unmodified reference, tests/golden/make_golden.py), host-side logic, and the C-ABI surface."""
import json
import os
import re

import numpy as np
import pytest
import torch

from oracle import amd_oracle as O
from hivae_b200 import configs, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
torch.set_grad_enabled(False)


def rel(a, b):
    return float((a - b).abs().max() / (b.abs().max() + 1e-12))


# ------------------------------------------------------------------ oracle vs golden
def test_oracle_signal_ops(golden):
    g = golden("signal_ops")
    for name, o in zip(("LL", "HL", "LH", "HH"), O.haar_dwt2(g["dwt_x"])):
        assert torch.equal(o, g["dwt_" + name])
    bands = torch.cat([g["dwt_LL"], g["dwt_HL"], g["dwt_LH"], g["dwt_HH"]], 0)
    assert torch.equal(O.haar_idwt2(bands), g["iwt_out"])
    assert rel(g["iwt_out"], g["dwt_x"]) < 1e-6  # DWT -> IWT round trip
    for d in (5, 6):
        lo, hi = O.freq_split3d(g["freq_x"], O.gaussian_lpf([32, 32, 32], d / 10, d / 10))
        assert rel(lo, g["freq_lo_%d" % d]) < 1e-6 and rel(hi, g["freq_hi_%d" % d]) < 1e-6
        assert rel(lo + hi, g["freq_x"]) < 1e-5
    zt, tgt = O.rf_train_tuple(g["rf_z1"], g["rf_ts"], g["rf_z0"])
    assert torch.equal(zt, g["rf_zt"]) and torch.equal(tgt, g["rf_target"])
    assert torch.equal(O.rf_target(zt, tgt, g["rf_ts"]), g["rf_rec"])


def test_oracle_haar_pyramid_roundtrip():
    x = torch.randn(2, 3, 64, 48, generator=torch.Generator().manual_seed(0))
    for levels in (1, 2, 3):
        bands, ll = O.haar_pyramid(x, levels)
        assert ll.shape[-2:] == (64 >> levels, 48 >> levels)
        assert rel(O.haar_pyramid_inverse(bands, ll), x) < 1e-6


def test_oracle_blocks(golden):
    g = golden("blocks")
    D, H = 128, 2
    xh, xe, te = g["x_hid"], g["x_enc"], g["temb"]
    from hivae_b200 import modules as M
    blocks = {"basic": M.BasicTransformerBlock(D, H, 64, attention_bias=True),
              "cross": M.BasicCrossTransformerBlock(D, D, H, 64, attention_bias=True),
              "joint": M.AMDTransformerBlock(D, H, 64, 512, attention_bias=True),
              "dit": M.BasicDiTBlock(D, H, 64, 512, attention_bias=True)}
    for name, blk in blocks.items():
        # the product modules expose exactly the reference's parameter names -> same synthetic weights
        sd = {"b." + k: v for k, v in synth.synth_state_dict(((k, v.shape) for k, v in blk.state_dict().items()), 11).items()}
        if name == "basic":
            assert rel(O.basic_block(sd, "b", xe, H), g["basic_out"]) < 2e-5
        elif name == "cross":
            assert rel(O.basic_block(sd, "b", xh, H, enc=xe), g["cross_out"]) < 2e-5
        elif name == "joint":
            oh, oe = O.joint_block(sd, "b", xh, xe, te, H)
            assert rel(oh, g["joint_hid"]) < 2e-5 and rel(oe, g["joint_enc"]) < 2e-5
        else:
            assert rel(O.dit_block(sd, "b", xe, te, H), g["dit_out"]) < 2e-5


def _small_oracle():
    import hivae_b200 as hv
    cfg = configs.SMALL
    m = hv.AMDModel_New(**cfg)
    sd = synth.synth_state_dict(((k, v.shape) for k, v in m.state_dict().items()), 0)
    return O.AMDOracle(sd, cfg["enc_nhead"], cfg["diffusion_attn_num_heads"], 16, 2, cfg["motion_token_num"]), m, sd


def test_oracle_small_model(golden):
    """BASELINE.json config 1: small AMD encoder + rectified-flow decoder, fp32 CPU, one clip."""
    g = golden("small_model")
    orc, _, _ = _small_oracle()
    video, ref_img = synth.split_clip(synth.synth_clip_latents(1, seed=1))
    cam, osrc, otgt = orc.encode(video, ref_img)
    assert rel(cam, g["camera"]) < 1e-5 and rel(osrc, g["obj_src"]) < 1e-5 and rel(otgt, g["obj_tgt"]) < 1e-5
    trace = []
    zi, zt, zj = orc.decode(video, ref_img, cam, osrc, otgt, 1000, 4, z0=g["z0"], trace=trace)
    for i, p in enumerate(trace):
        assert rel(p, g["pre_%d" % i]) < 1e-5
    assert rel(zt, g["zt"]) < 1e-5
    assert abs(O.psnr(zt, zj) - O.psnr(g["zt"], zj)) < 0.05
    ts = torch.tensor([437.0]).repeat_interleave(16)
    pre, vel, ld = orc.forward(video, ref_img, ts, g["z0"])
    assert rel(pre, g["fwd_pre"]) < 1e-5 and rel(vel, g["fwd_vel"]) < 1e-6
    assert abs(float(ld["loss"]) - float(g["fwd_loss"][0])) < 1e-5 * float(g["fwd_loss"][0])


def test_oracle_rope_attention(golden):
    g = golden("signal_ops")
    sd = synth.synth_state_dict([("wq.weight", (128, 128)), ("wq.bias", (128,)), ("wk.weight", (128, 128)), ("wk.bias", (128,)),
                                 ("wv.weight", (128, 128)), ("wv.bias", (128,)), ("out.weight", (128, 128)), ("out.bias", (128,))], 3)
    x = g["ropeattn_x"]
    args = (x, sd["wq.weight"], sd["wq.bias"], sd["wk.weight"], sd["wk.bias"], sd["wv.weight"], sd["wv.bias"], 4, 40)
    assert rel(O.rope_attention(*args), g["ropeattn_rope"]) < 1e-5
    assert rel(O.rope_attention(*args, inference=False), g["ropeattn_norope"]) < 1e-5
    assert rel(O.rope_attention(*args, mask=torch.tril(torch.ones(24, 24)), start_pos=3), g["ropeattn_rope_mask"]) < 1e-5


TENC = dict(img_height=32, img_width=32, img_inchannel=4, img_patch_size=2, motion_token_num=4, motion_channel=16,
            need_norm_out=True, video_frames=4, num_attention_heads=2, attention_head_dim=64, num_layers=2)  # = make_golden_variants.TENC


# ------------------------------------------------------------------ host logic (no GPU)
def test_state_dict_keys_match_reference():
    """drop-in contract: a reference model.safetensors must load unchanged."""
    import hivae_b200 as hv
    with open(os.path.join(ROOT, "tests", "golden", "state_dict_keys_default.json")) as f:
        ref = json.load(f)
    m = hv.AMD_models["AMD_N"](**{k: v for k, v in configs.DEFAULT.items()
                                  if k not in ("enc_nhead", "enc_ndim", "diffusion_attn_head_dim", "diffusion_attn_num_heads",
                                               "diffusion_out_channels", "diffusion_num_layers")})
    mine = {k: list(v.shape) for k, v in m.state_dict().items()}
    assert mine == ref
    assert abs(sum(int(np.prod(s)) for s in mine.values()) - 695.00e6) < 0.01e6


def test_state_dict_keys_object_only_variant():
    """diffusion_model_type='default' (AMDDiffusionTransformerModelTempMotion, transformer.py:1483-1743):
    same keys/shapes as the unmodified reference; unknown types raise like the reference's list lookup."""
    import hivae_b200 as hv
    with open(os.path.join(ROOT, "tests", "golden", "state_dict_keys_small_object_only.json")) as f:
        ref = json.load(f)
    m = hv.AMDModel_New(**configs.SMALL_OBJECT_ONLY)
    assert type(m.diffusion_transformer).__name__ == "AMDDiffusionTransformerModelTempMotion"
    assert {k: list(v.shape) for k, v in m.state_dict().items()} == ref
    assert not any(".spatial_blocks." in k or ".camera_transformer_blocks." in k for k in ref)
    with pytest.raises(IndexError):
        hv.AMDModel_New(**dict(configs.SMALL, diffusion_model_type="nope"))


def test_state_dict_keys_temporal_encoder():
    """a16 MotionEncoderLearnTokenTemporalTransformer (transformer.py:195-296): keys/shapes of the unmodified reference."""
    import hivae_b200 as hv
    with open(os.path.join(ROOT, "tests", "golden", "state_dict_keys_temporal_encoder.json")) as f:
        ref = json.load(f)
    m = hv.MotionEncoderLearnTokenTemporalTransformer(**TENC)
    assert {k: list(v.shape) for k, v in m.state_dict().items()} == ref
    assert m.motion_temporal_embedding.shape == (1, 16, 128)
    with pytest.raises(NotImplementedError):
        hv.AMDTransformerMotionBlock(128, 2, 64, time_embed_dim=32)


DITS = dict(num_attention_heads=2, attention_head_dim=64, out_channels=4, num_layers=8, image_width=32, image_height=32,
            image_patch_size=2, image_in_channels=8, motion_token_num=4, motion_in_channels=16, motion_target_num_frame=4)  # = make_golden_variants.DITS


def test_state_dict_keys_imgspatial_dit():
    """a17 sibling AMDDiffusionTransformerModelImgSpatial (transformer.py:2044-2184): keys/shapes of the unmodified reference."""
    import hivae_b200 as hv
    with open(os.path.join(ROOT, "tests", "golden", "state_dict_keys_imgspatial_dit.json")) as f:
        ref = json.load(f)
    m = hv.AMDDiffusionTransformerModelImgSpatial(motion_type="decouple", **DITS)
    assert {k: list(v.shape) for k, v in m.state_dict().items()} == ref
    # the oldest sibling AMDDiffusionTransformerModel (transformer.py:1010-1128): no temporal blocks
    with open(os.path.join(ROOT, "tests", "golden", "state_dict_keys_legacy_dit.json")) as f:
        ref = json.load(f)
    kw = dict({k: v for k, v in DITS.items() if k != "motion_target_num_frame"}, num_layers=9)
    m = hv.AMDDiffusionTransformerModel(motion_type="plus", **kw)
    assert {k: list(v.shape) for k, v in m.state_dict().items()} == ref


def test_state_dict_keys_amdmodel():
    """a18 sibling AMDModel (model_AMD.py:33-256): keys/shapes of the unmodified reference for two runnable configurations;
    the branches the reference cannot run raise the same kind of error."""
    import hivae_b200 as hv
    for name, cfg in (("temporal", configs.AMDMODEL_TEMPORAL), ("legacy", configs.AMDMODEL_LEGACY)):
        with open(os.path.join(ROOT, "tests", "golden", "state_dict_keys_amdmodel_%s.json" % name)) as f:
            ref = json.load(f)
        m = hv.AMDModel(**cfg)
        assert {k: list(v.shape) for k, v in m.state_dict().items()} == ref, name
    with pytest.raises(IndexError):
        hv.AMDModel(**dict(configs.AMDMODEL_LEGACY, diffusion_model_type="nope"))
    with pytest.raises(NotImplementedError):
        hv.AMDModel(**dict(configs.AMDMODEL_LEGACY, diffusion_model_type="dual"))


def test_sincos_tables_match_oracle():
    from hivae_b200 import embeddings as E
    assert np.array_equal(E.sincos_2d(256, 16, 16), O.sincos_2d(256, 16, 16))
    assert np.array_equal(E.sincos_2d(128, 4, 4), O.sincos_2d(128, 4, 4))
    assert np.array_equal(E.sincos_1d(512, np.arange(82)), O.sincos_1d(512, np.arange(82)))


def test_circulant_factorisation_equals_fft_split(golden):
    """the separable circular-convolution form used by the CUDA kernel == the reference's FFT split"""
    from hivae_b200 import frequency_utils as F
    g = golden("signal_ops")
    x = g["freq_x"].double()
    for d in (0.5, 0.6):
        ct, ch, cw = (torch.from_numpy(F._circulant(F._gauss_1d(32, d))) for _ in range(3))
        lo = torch.einsum("ai,bj,ck,...ijk->...abc", ct, ch, cw, x)
        assert rel(lo.float(), g["freq_lo_%d" % int(d * 10)]) < 1e-6
    x2 = g["freq2_x"].double()  # ragged even sizes, anisotropic cut-offs, untagged mask -> rank-1 factorisation path
    lpf2 = O.gaussian_lpf([8, 12, 20], 0.3, 0.3)
    mats = [torch.from_numpy(F._circulant(v)) for v in F._factor_mask(lpf2)]
    lo2 = torch.einsum("ai,bj,ck,...ijk->...abc", *mats, x2)
    assert rel(lo2.float(), O.freq_split3d(g["freq2_x"], lpf2)[0]) < 1e-6
    with pytest.raises(NotImplementedError):
        F._circulant(F._gauss_1d(7, 0.5))
    m = F.gaussian_low_pass_filter([32, 32, 32], 0.6, 0.6)
    assert torch.equal(m, O.gaussian_lpf([32, 32, 32], 0.6, 0.6)) and m._hv_gauss == (0.6, 0.6)
    assert torch.count_nonzero(F.gaussian_low_pass_filter([4, 4, 4], 0.0, 0.5)) == 0


def test_rectified_flow_host(golden):
    from hivae_b200 import RectifiedFlow
    g = golden("signal_ops")
    rf = RectifiedFlow(1000)
    zt, tgt = rf.get_train_tuple(z0=g["rf_z0"], z1=g["rf_z1"], time_step=g["rf_ts"])
    assert torch.equal(zt, g["rf_zt"]) and torch.equal(tgt, g["rf_target"])
    assert torch.equal(rf.get_target_with_zt_vel(zt, tgt, g["rf_ts"]), g["rf_rec"])
    out = rf.sample_loop(lambda z, t: torch.ones_like(z), torch.zeros(2, 1, 2, 2), sample_step=4)
    assert torch.allclose(out, torch.ones(2, 1, 2, 2))  # 4 Euler steps of dt = 1/4
    assert O.sample_steps(1000, 4) == [1000, 750, 500, 250]


def test_config_roundtrip(tmp_path):
    import hivae_b200 as hv
    m = hv.AMDModel_New(**configs.SMALL)
    m.save_config(str(tmp_path))
    cfg = hv.AMDModel_New.load_config(str(tmp_path))
    m2 = hv.AMDModel_New.from_config(cfg)
    assert [k for k, _ in m2.state_dict().items()] == [k for k, _ in m.state_dict().items()]
    with pytest.raises(NotImplementedError):
        hv.AMDModel_New(**dict(configs.SMALL, diffusion_model_type="default"))


def test_synth_weights_are_order_independent():
    a = synth.synth_state_dict([("x.weight", (4, 8)), ("y.norm.weight", (8,))], 0)
    b = synth.synth_state_dict([("y.norm.weight", (8,)), ("x.weight", (4, 8))], 0)
    assert torch.equal(a["x.weight"], b["x.weight"]) and torch.equal(a["y.norm.weight"], b["y.norm.weight"])
    assert abs(float(a["y.norm.weight"].mean()) - 1.0) < 0.2


# ------------------------------------------------------------------ C-ABI surface
def test_cabi_exports_every_declared_symbol():
    from hivae_b200 import _lib
    hdr = open(os.path.join(ROOT, "include", "hivae_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(hv_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 20
    lib = _lib.load()
    for name in declared:
        assert hasattr(lib, name), name
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    assert lib.hv_version() == 100
    assert isinstance(lib.hv_last_error(), bytes)


def test_no_cpu_fallback():
    """the product path must fail loudly without a GPU, never fall back to torch/CPU"""
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import hivae_b200 as hv
    with pytest.raises(RuntimeError):
        hv.dwt_init(torch.randn(1, 1, 4, 4))
    with pytest.raises(RuntimeError):
        hv.BasicTransformerBlock(128, 2, 64, attention_bias=True)(torch.randn(1, 4, 128))
    from hivae_b200 import backward as Bk
    blk = hv.BasicTransformerBlock(128, 2, 64, attention_bias=True)
    blk.hv_compute_dtype = torch.bfloat16
    with pytest.raises(RuntimeError):   # the differentiable block has no torch / CPU path either
        Bk.basic_block_apply(blk, torch.randn(1, 8, 128, requires_grad=True))
    with pytest.raises(RuntimeError):
        Bk.transpose2d(torch.randn(4, 4))


def test_product_does_not_import_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "hivae_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle|import_module\(.oracle", src, flags=re.M), f


# ------------------------------------------------------------------ checkpoint compatibility
def test_reference_format_checkpoint_loads(tmp_path):
    """`checkpoint-<step>/model.safetensors` written by safetensors.torch.save_model — from the UNMODIFIED reference module
    when /root/reference is present (build container), otherwise from a module with the same key contract — loads with
    safetensors.torch.load_model into hivae_b200.AMDModel_New (amd_inference.py:99-109), and find_latest_checkpoint /
    model_load_pretrain (model/utils.py:178-185, 217-237) behave like the reference's."""
    import hivae_b200 as hv
    from safetensors.torch import load_model, save_model
    from oracle import ref_loader
    cfg = configs.SMALL
    if ref_loader.available():
        ref = ref_loader.load()
        torch.manual_seed(3)
        src = ref.model_AMD.AMDModel_New(**cfg)
    else:
        torch.manual_seed(3)
        src = hv.AMDModel_New(**cfg)
    with torch.no_grad():
        for p in src.parameters():
            p.add_(0.01 * torch.randn_like(p))
    for step in (500, 12000, 3000):
        os.makedirs(tmp_path / ("checkpoint-%d" % step))
    (tmp_path / "checkpoint-best").mkdir()
    path = hv.find_latest_checkpoint(str(tmp_path))
    assert path == str(tmp_path / "checkpoint-12000" / "model.safetensors")
    save_model(src, path)
    m = hv.AMDModel_New(**cfg)
    missing, unexpected = load_model(m, path)
    assert not missing and not unexpected
    want = src.state_dict()
    for k, v in m.state_dict().items():
        assert torch.equal(v, want[k]), k
    # filtered load: everything but the denoiser (the reference's keyword filter), strict=False
    m2 = hv.AMDModel_New(**cfg)
    before = {k: v.clone() for k, v in m2.state_dict().items()}
    hv.model_load_pretrain(m2, path, not_load_keyword="diffusion_transformer")
    for k, v in m2.state_dict().items():
        assert torch.equal(v, before[k] if "diffusion_transformer" in k else want[k]), k
    with pytest.raises(ValueError):
        hv.find_latest_checkpoint(str(tmp_path / "checkpoint-best"))
    # config round trip next to the weights (train_amd.py:175, amd_inference.py:99-107)
    m.save_config(str(tmp_path))
    m3 = hv.AMDModel_New.from_config(hv.AMDModel_New.load_config(str(tmp_path)))
    assert sorted(m3.state_dict()) == sorted(want)
    hv.frozen_model(m3, "camera_motion_encoder")
    assert all(p.requires_grad != ("camera_motion_encoder" in k) for k, p in m3.named_parameters())


def test_state_dict_keys_and_oracle_camera_rec_models(golden):
    """f3: AMDModel_Camera / AMDModel_Rec (model_AMD.py:894-1661, 2840-3125): state-dict keys/shapes of the unmodified
    reference, every AMD_models factory name of the reference exists, and the oracle restatements reproduce the committed
    fixtures (forward / 3-step sample of AMDModel_Camera with both denoisers, forward of AMDModel_Rec)."""
    import hivae_b200 as hv
    assert sorted(hv.AMD_models) == sorted(["AMD_S", "AMD_S_Camera", "AMD_N", "AMD_L", "AMD_S_Rec", "AMD_S_RecSplit"])  # model_AMD.py:3203-3210
    g = golden("amd_model_camera")
    ts = torch.tensor([411.0]).repeat_interleave(4)
    for dmt in ("spatial", "default"):
        cfg = dict(configs.AMDMODEL_CAMERA, diffusion_model_type=dmt)
        with open(os.path.join(ROOT, "tests", "golden", "state_dict_keys_amdmodel_camera_%s.json" % dmt)) as f:
            ref = json.load(f)
        m = hv.AMDModel_Camera(**cfg)
        assert {k: list(v.shape) for k, v in m.state_dict().items()} == ref, dmt
        sd = synth.synth_state_dict(((k, v.shape) for k, v in m.state_dict().items()), 8)
        orc = O.AMDModelCameraOracle(sd, cfg)
        pre, vel, ld = orc.forward(g["video"], g["ref_img"], ts, g["z0"], g["video_grey"], g["ref_img_grey"], g["camera_mask"])
        assert float((pre - g[dmt + "_pre"]).abs().max() / g[dmt + "_pre"].abs().max()) < 2e-5
        zt, zj = orc.sample(g["video"], g["ref_img"], 3, g["z0"], g["video_grey"], g["ref_img_grey"], g["camera_mask"])
        assert float((zt - g[dmt + "_zt"]).abs().max() / g[dmt + "_zt"].abs().max()) < 2e-5
    with pytest.raises(NotImplementedError):
        hv.AMDModel_Camera(**dict(configs.AMDMODEL_CAMERA, use_motiontemporal=True))  # not runnable upstream either
    gr = golden("amd_model_rec")
    with open(os.path.join(ROOT, "tests", "golden", "state_dict_keys_amdmodel_rec.json")) as f:
        ref = json.load(f)
    m = hv.AMDModel_Rec(**configs.AMDMODEL_REC)
    assert {k: list(v.shape) for k, v in m.state_dict().items()} == ref
    sd = synth.synth_state_dict(((k, v.shape) for k, v in m.state_dict().items()), 9)
    pre, zj, ld = O.AMDModelRecOracle(sd, configs.AMDMODEL_REC).forward(gr["video"], gr["ref_img"])
    assert float((pre - gr["pre"]).abs().max() / gr["pre"].abs().max()) < 2e-5
    with pytest.raises(AttributeError):
        m.sample(gr["video"], gr["ref_img"])


def test_oracle_random_masking_matches_reference_fixture(golden):
    """random_masking (transformer.py:164-192, 558-586) in the oracle: with the reference's noise tensors the masked encode
    of the small model reproduces what the unmodified reference returned (tests/golden/small_masked.safetensors)"""
    g = golden("small_masked")
    cfg = configs.SMALL
    import hivae_b200 as hv
    m = hv.AMDModel_New(**cfg)
    sd = synth.synth_state_dict(((k, v.shape) for k, v in m.state_dict().items()), 0)
    orc = O.AMDOracle(sd, cfg["enc_nhead"], cfg["diffusion_attn_num_heads"], 16, 2, cfg["motion_token_num"])
    video, ref_img = synth.split_clip(synth.synth_clip_latents(2, seed=1))
    cam, osrc, otgt = orc.encode(video, ref_img, cam_keep=O.mask_keep(g["enc_cam_noise"], 0.5), obj_keep=O.mask_keep(g["enc_obj_noise"], 0.8))
    assert cam.shape == g["enc_camera"].shape == (2, 16, 8, 8)
    for a, b in ((cam, g["enc_camera"]), (osrc, g["enc_obj_src"]), (otgt, g["enc_obj_tgt"])):
        assert float((a - b).abs().max() / b.abs().max()) < 2e-5


def test_freq_split_mma_index_arithmetic_emulated():
    """tools/emulate_freq_mma.py walks the loops of freq_split3d_mma_kernel (csrc/signal.cu) lane by lane with exact
    products placed through the PTX m16n8k8 fragment layout: the n-permuted T pass, the k-permuted W pass and the swizzled
    slab must reproduce three plain matrix products for every template instance, and write every output exactly once"""
    import importlib.util
    spec = importlib.util.spec_from_file_location("emulate_freq_mma", os.path.join(ROOT, "tools", "emulate_freq_mma.py"))
    emu = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(emu)
    for T, N in emu.INSTANCES:
        err, unwritten, rec = emu.emulate(T, N)
        assert err < 1e-12 and unwritten == 0 and rec < 1e-12, (T, N, err, unwritten, rec)


def test_freq_split_3xtf32_error_budget():
    """numerical design of the tensor-core frequency split, on the CPU: operands split as in split_tf32 (csrc/signal.cu: hi =
    round-to-nearest TF32 by integer add + mask, lo = x - hi with its 13 low mantissa bits ignored by the tensor core),
    products hi.hi + hi.lo + lo.hi accumulated in fp32 over the three circulant passes of frequency_utils.py:84-108; the
    result has to stay an order of magnitude inside the 1e-5 bound of the golden test"""
    from hivae_b200 import frequency_utils as F

    def split(a):
        bits = a.astype(np.float32).view(np.uint32)
        hi = ((bits + np.uint32(0x1000)) & np.uint32(0xFFFFE000)).view(np.float32)
        lo = ((a.astype(np.float32) - hi).view(np.uint32) & np.uint32(0xFFFFE000)).view(np.float32)   # hardware truncation
        return hi, lo

    def mm3(c, x):     # c (n, n) applied along axis 0 of x (n, m); fp32 accumulation
        ch_, cl_ = split(c)
        xh, xl = split(x)
        return (cl_ @ xh + ch_ @ xl + ch_ @ xh).astype(np.float32)

    rng = np.random.default_rng(3)
    T = H = W = 32
    x = rng.standard_normal((T, H, W)).astype(np.float32)
    mats = [F._circulant(F._gauss_1d(n, d)).astype(np.float32) for n, d in ((T, 0.6), (H, 0.6), (W, 0.6))]
    y = mm3(mats[0], x.reshape(T, H * W)).reshape(T, H, W)
    y = mm3(mats[1], y.transpose(1, 0, 2).reshape(H, T * W)).reshape(H, T, W).transpose(1, 0, 2)
    y = mm3(mats[2], np.ascontiguousarray(y.transpose(2, 0, 1)).reshape(W, T * H)).reshape(W, T, H).transpose(1, 2, 0)
    ref = np.einsum('at,thw->ahw', mats[0].astype(np.float64), x.astype(np.float64))
    ref = np.einsum('bh,ahw->abw', mats[1].astype(np.float64), ref)
    ref = np.einsum('cw,abw->abc', mats[2].astype(np.float64), ref)
    err = np.abs(y - ref).max() / np.abs(ref).max()
    assert err < 1e-6, err
    # and against the oracle's FFT formulation of the reference
    olo, _ = O.freq_split3d(torch.from_numpy(x)[None, None], O.gaussian_lpf([T, H, W], 0.6, 0.6))
    assert float(np.abs(y - olo[0, 0].numpy()).max() / np.abs(ref).max()) < 2e-6
