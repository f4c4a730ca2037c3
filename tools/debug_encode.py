"""stage-by-stage comparison of the small model's encode against the CPU oracle (debug aid)"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import hivae_b200 as hv
from hivae_b200 import configs, synth, ops
from hivae_b200.frequency_utils import freq_split_gauss
from oracle import amd_oracle as O
torch.set_grad_enabled(False)
dev = torch.device("cuda:0")
rel = lambda a, b: float((a.float().cpu() - b).abs().max() / (b.abs().max() + 1e-12))
cfgname = sys.argv[1] if len(sys.argv) > 1 else "small"
cfg = configs.CONFIGS[cfgname]
m = hv.AMDModel_New(**cfg)
sd = synth.synth_state_dict(((k, v.shape) for k, v in m.state_dict().items()), 0)
m.load_state_dict(sd); m.to(dev)
orc = O.AMDOracle(sd, cfg["enc_nhead"], cfg["diffusion_attn_num_heads"], 16, 2, cfg["motion_token_num"])
video, ref_img = synth.split_clip(synth.synth_clip_latents(1, seed=1))
n, t, c, h, w = video.shape
rv = torch.cat([ref_img, video], 1)
lpf = O.gaussian_lpf([2 * t, h, w], 0.6, 0.6)
lo, _ = O.freq_split3d(rv.permute(0, 2, 1, 3, 4), lpf)
lf = lo.permute(0, 2, 1, 3, 4)[:, t:]
glo, _ = freq_split_gauss(rv.to(dev).permute(0, 2, 1, 3, 4).contiguous(), 0.6, 0.6)
print("freq lo", rel(glo, lo))
olf = O.camera_down(orc.sd, "camera_down", lf.flatten(0, 1))
glf = m._camera_down(lf.flatten(0, 1).contiguous().to(dev))
print("camera_down", rel(glf, olf))
olf5 = olf.view(n, t, *olf.shape[1:])
ocam = O.camera_encoder(orc.sd, "camera_motion_encoder", olf5, orc.eh, orc.T, 2, False)
gcam = m.camera_motion_encoder(olf5.to(dev))
print("camera enc", rel(gcam, ocam))
oobj = O.object_encoder(orc.sd, "object_motion_encoder", rv, orc.eh, 2, False)
gobj = m.object_motion_encoder(rv.to(dev))
print("object enc", rel(gobj, oobj), [rel(gobj[:, i], oobj[:, i]) for i in range(0, 32, 5)])
# one-layer variants: truncate blocks
for L in (0, 1):
    e = m.object_motion_encoder
    saved = e.transformer_blocks
    e.transformer_blocks = torch.nn.ModuleList(list(saved)[:L]); e._hv_key = None
    sd2 = {k: v for k, v in orc.sd.items() if not any(k.startswith("object_motion_encoder.transformer_blocks.%d." % j) for j in range(L, 99))}
    print("object enc with %d layers" % L, rel(e(rv.to(dev)), O.object_encoder(sd2, "object_motion_encoder", rv, orc.eh, 2, False)))
    e.transformer_blocks = saved; e._hv_key = None
    e = m.camera_motion_encoder
    saved = e.transformer_blocks
    e.transformer_blocks = torch.nn.ModuleList(list(saved)[:L]); e._hv_key = None
    sd2 = {k: v for k, v in orc.sd.items() if not any(k.startswith("camera_motion_encoder.transformer_blocks.%d." % j) for j in range(L, 99))}
    print("camera enc with %d layers" % L, rel(e(olf5.to(dev)), O.camera_encoder(sd2, "camera_motion_encoder", olf5, orc.eh, orc.T, 2, False)))
    e.transformer_blocks = saved; e._hv_key = None
# DiT with oracle motions
cam, osrc, otgt = orc.encode(video, ref_img)
zi, zj = ref_img.flatten(0, 1), video.flatten(0, 1)
z0 = torch.randn(16, 4, 32, 32, generator=torch.Generator().manual_seed(2))
step = torch.ones(16) * 600
zt, _ = O.rf_train_tuple(zj, step, z0)
opre = orc.dit(zi, zt, step, cam, osrc, otgt)
gpre = m.diffusion_transformer(torch.cat([zi, zt], 1).to(dev), step.to(dev), camera_motion_target_hidden_states=cam.to(dev),
                               object_motion_source_hidden_states=osrc.to(dev), object_motion_target_hidden_states=otgt.to(dev))
print("dit", rel(gpre, opre))
