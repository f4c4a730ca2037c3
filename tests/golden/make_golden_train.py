"""Round-2 fixtures from the UNMODIFIED reference (via oracle/shim); build container only.

  python tests/golden/make_golden_train.py train     -> train_step_small.safetensors   (SURVEY 8c golden #6)
  python tests/golden/make_golden_train.py sample20  -> default_sample20.safetensors   (BASELINE config 2, 20 Euler steps)
  python tests/golden/make_golden_train.py masked    -> small_masked.safetensors       (random_masking, host-drawn noise)

train: one step of train_amd.py:459-496 on the small config in fp32 on CPU:
  forward(video, ref_img) with a fixed (time_step, noise) -> loss = diff_loss -> loss.backward()
  -> clip_grad_norm_(params, 1.0) -> torch.optim.AdamW(lr=1e-4).step()
and stores loss, the global gradient norm, the per-parameter gradient norms of EVERY parameter, a set of named
gradients in full and their post-step parameter deltas.  The oracle restatement (oracle/amd_oracle.py) is checked
against the reference's autograd for the same step so that it can serve as the checker on the GPU box.
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

# parameters whose full gradient / AdamW delta is stored (every structural role of the path once)
NAMED = [
    "diffusion_transformer.proj_out.weight",
    "diffusion_transformer.norm_out.linear.weight",
    "diffusion_transformer.norm_final.weight",
    "diffusion_transformer.spatial_blocks.1.norm1.linear.bias",
    "diffusion_transformer.spatial_blocks.0.attn1.to_q.weight",
    "diffusion_transformer.object_transformer_blocks.1.norm2.linear.weight",
    "diffusion_transformer.object_transformer_blocks.0.attn1.norm_k.weight",
    "diffusion_transformer.object_transformer_blocks.0.ff.net.0.proj.weight",
    "diffusion_transformer.camera_transformer_blocks.0.attn1.to_out.0.weight",
    "diffusion_transformer.camera_transformer_blocks.1.ff.net.2.bias",
    "diffusion_transformer.image_patch_embed.proj.weight",
    "diffusion_transformer.time_embedding.linear_1.weight",
    "diffusion_transformer.time_embedding.linear_2.bias",
    "diffusion_transformer.camera_motion_patch_embed.weight",
    "diffusion_transformer.object_motion_patch_embed.weight",
    "diffusion_transformer.source_token",
    "diffusion_transformer.target_token",
    "object_motion_encoder.motion_token",
    "object_motion_encoder.motion_embed.weight",
    "object_motion_encoder.patch_embed.proj.weight",
    "object_motion_encoder.transformer_blocks.0.attn1.to_v.weight",
    "object_motion_encoder.transformer_blocks.1.norm2.weight",
    "object_motion_encoder.proj_out.weight",
    "camera_motion_encoder.motion_token",
    "camera_motion_encoder.patch_embed.proj.weight",
    "camera_motion_encoder.transformer_blocks.0.attn1.to_k.weight",
    "camera_motion_encoder.transformer_blocks.1.ff.net.0.proj.bias",
    "camera_motion_encoder.norm_final.bias",
    "camera_motion_encoder.proj_out.weight",
    "camera_down.0.weight",
    "camera_down.1.bias",
]


def relerr(a, b):
    return float((a - b).abs().max() / (b.abs().max() + 1e-30))


def build_ref_model(ref, cfg, seed):
    torch.manual_seed(0)
    m = ref.model_AMD.AMDModel_New(**cfg)
    sd = synth.synth_state_dict(((k, v.shape) for k, v in m.state_dict().items()), seed)
    m.load_state_dict(sd)
    return m, sd


def fix_noise(m, z0):
    orig = m.scheduler.get_train_tuple
    m.scheduler.get_train_tuple = lambda z0_=None, z1=None, time_step=None: orig(z0=z0, z1=z1, time_step=time_step)


def train_golden(ref, n_clips=2):
    cfg = configs.SMALL
    m, sd = build_ref_model(ref, cfg, seed=0)
    m.train()
    z = synth.synth_clip_latents(n_clips, seed=1)
    video, ref_img = synth.split_clip(z)
    z0 = torch.randn(n_clips * 16, 4, 32, 32, generator=torch.Generator().manual_seed(2))
    steps = torch.tensor([437.0, 81.0, 903.0, 655.0][:n_clips])
    ts = steps.repeat_interleave(16)
    fix_noise(m, z0)
    m.prepare_timestep = lambda batch_size, device, time_step=None: (steps if batch_size == n_clips else ts).to(device)
    pre, vel, ld = m(video, ref_img)
    ld["loss"].backward()
    names = [k for k, _ in m.named_parameters()]
    grads = {k: p.grad.detach().clone() for k, p in m.named_parameters()}
    missing = [k for k in names if grads.get(k) is None]
    assert not missing, missing
    total = torch.nn.utils.clip_grad_norm_(m.parameters(), 1.0)
    before = {k: p.detach().clone() for k, p in m.named_parameters()}
    opt = torch.optim.AdamW(m.parameters(), lr=1e-4)
    opt.step()
    delta = {k: (p.detach() - before[k]) for k, p in m.named_parameters()}
    print("  reference train step: loss %.6f rec_loss %.6f grad-norm %.6f (%d parameters tensors)" %
          (float(ld["loss"]), float(ld["rec_loss"]), float(total), len(names)))

    # the oracle restatement under torch autograd must reproduce the reference's gradients
    sdg = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    orc = O.AMDOracle(sdg, cfg["enc_nhead"], cfg["diffusion_attn_num_heads"], 16, 2, cfg["motion_token_num"])
    with torch.enable_grad():
        opre, ovel, old = orc.forward(video, ref_img, ts, z0)
        old["loss"].backward()
    # norm_k.bias has a mathematically ZERO gradient (a constant added to every key shifts all scores of a query equally,
    # softmax is invariant), what autograd returns there is rounding noise: tensors are compared relative to
    # max(|reference|, 1e-5 * global gradient norm)
    worst, floor = 0.0, 1e-5 * float(total)
    for k in names:
        e = float((sdg[k].grad - grads[k]).abs().max() / max(float(grads[k].abs().max()), floor))
        worst = max(worst, e)
        assert e < 2e-4, (k, e, float(grads[k].abs().max()))
    print("  oracle autograd vs reference autograd: worst per-tensor max-rel-err %.2e over %d tensors" % (worst, len(names)))
    fx = {"z0": z0, "steps": steps, "pre": pre.detach(), "vel": vel.detach(),
          "loss": torch.stack([ld["loss"].detach(), ld["diff_loss"].detach(), ld["rec_loss"].detach()]),
          "grad_norm": total.detach().reshape(1),
          "grad_norms_all": torch.stack([grads[k].norm() for k in names])}
    for k in NAMED:
        assert k in grads, k
        fx["grad." + k] = grads[k]
        fx["delta." + k] = delta[k]
    save_file({k: v.contiguous() for k, v in fx.items()}, os.path.join(OUT, "train_step_small.safetensors"))
    import json
    with open(os.path.join(OUT, "train_step_small_names.json"), "w") as f:
        json.dump(names, f, indent=0)
    print("  wrote train_step_small.safetensors")


def sample20_golden(ref):
    torch.set_grad_enabled(False)
    cfg = configs.DEFAULT
    m, sd = build_ref_model(ref, cfg, seed=0)
    m.eval()
    video, ref_img = synth.split_clip(synth.synth_clip_latents(1, seed=1))
    z0 = torch.randn(16, 4, 32, 32, generator=torch.Generator().manual_seed(2))
    fix_noise(m, z0)
    import time
    t0 = time.time()
    meta = m.sample(video, ref_img, sample_step=20, return_meta_info=True)
    print("  reference default-config 20-step sample: %.1f s on %d threads" % (time.time() - t0, torch.get_num_threads()))
    fx = {"z0": z0, "zt": meta["sample"].contiguous()}
    for i in (0, 6, 13, 19):
        fx["pre_%d" % i] = meta["pre_cache"][i]
    fx["psnr"] = torch.tensor([O.psnr(meta["sample"], meta["zj"])])
    print("  PSNR(zt, zj) = %.3f dB" % float(fx["psnr"]))
    save_file({k: v.contiguous() for k, v in fx.items()}, os.path.join(OUT, "default_sample20.safetensors"))
    print("  wrote default_sample20.safetensors")


def masked_golden(ref):
    """random_masking (transformer.py:164-192, 558-586): the noise tensors are drawn with torch.rand on the CPU
    generator in the reference's call order (forward: camera ratio draw, object ratio draw, camera encoder noise,
    object encoder noise), so a host that makes the same calls in the same order reproduces the kept indices."""
    torch.set_grad_enabled(False)
    cfg = configs.SMALL
    m, sd = build_ref_model(ref, cfg, seed=0)
    m.eval()
    video, ref_img = synth.split_clip(synth.synth_clip_latents(2, seed=1))
    z0 = torch.randn(32, 4, 32, 32, generator=torch.Generator().manual_seed(2))
    steps = torch.tensor([437.0, 81.0])
    ts = steps.repeat_interleave(16)
    fix_noise(m, z0)
    m.prepare_timestep = lambda batch_size, device, time_step=None: (steps if batch_size == 2 else ts).to(device)
    torch.manual_seed(1234)
    pre, vel, ld = m(video, ref_img, camera_mask_ratio=0.5, object_mask_ratio=0.8)
    torch.manual_seed(1234)
    cr = (0.6 + 0.4 * torch.rand(1)).item() * 0.5
    orr = (0.5 * torch.rand(1)).item() * 0.8
    cam_noise = torch.rand(2, 16)
    obj_noise = torch.rand(64, 256)
    fx = {"z0": z0, "steps": steps, "pre": pre, "vel": vel, "loss": torch.stack([ld["loss"], ld["rec_loss"]]),
          "ratios": torch.tensor([cr, orr], dtype=torch.float64), "cam_noise": cam_noise, "obj_noise": obj_noise}
    torch.manual_seed(1234)
    cam, osrc, otgt = m.encode(video, ref_img, camera_mask_ratio=0.5, object_mask_ratio=0.8)
    torch.manual_seed(1234)
    fx.update(enc_camera=cam, enc_obj_src=osrc, enc_obj_tgt=otgt, enc_cam_noise=torch.rand(2, 16), enc_obj_noise=torch.rand(64, 256))
    print("  masked forward: camera tokens %s, loss %.6f" % (tuple(cam.shape), float(ld["loss"])))
    save_file({k: v.contiguous() for k, v in fx.items()}, os.path.join(OUT, "small_masked.safetensors"))
    print("  wrote small_masked.safetensors")


if __name__ == "__main__":
    what = sys.argv[1:] or ["train", "masked", "sample20"]
    ref = ref_loader.load()
    for w in what:
        {"train": train_golden, "sample20": sample20_golden, "masked": masked_golden}[w](ref)
