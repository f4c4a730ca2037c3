"""AMDModel_New — the model-level API of the reference (model/model_AMD.py:1663-2838) on the
hivae_b200 kernels: same constructor kwargs, same state-dict keys (a reference `model.safetensors`
loads unchanged), same encode / decode / sample / forward signatures and return conventions.

Supported configuration = the one every shipped script uses (scripts/amd_new.sh:37-67):
diffusion_model_type='spatial', use_grey=False, use_regularizers=False,
need_motion_transformer=False; use_filter / use_camera_down / use_camera / use_object free.
"""
import inspect
import json
import os

import numpy as np
import torch
from torch import nn

from . import engine, ops
from .frequency_utils import freq_split_gauss
from .rectified_flow import RectifiedFlow
from .transformer import (AMDDiffusionTransformerModel, AMDDiffusionTransformerModelImgSpatial, AMDReconstructTransformerModel,
                          AMDDiffusionTransformerModelImgSpatialTempMotion, AMDDiffusionTransformerModelTempMotion,
                          MotionEncoderLearnTokenOnlyTemporalTransformer, MotionEncoderLearnTokenTemporalTransformer,
                          MotionEncoderLearnTokenTransformer)


def l2(pre, gt):
    """model/loss.py:13-14 (sm_100a reduction kernel on the device, torch on CPU tensors)."""
    if pre.is_cuda:
        from . import optim
        return optim.mse_loss(pre, gt)
    return torch.nn.functional.mse_loss(pre, gt)


class _ConfigModel(nn.Module):
    """The slice of diffusers' ConfigMixin/ModelMixin the reference relies on
    (train_amd.py:175,224; amd_inference.py:99-107): records kwargs, from_config / save_config."""
    config_name = "config.json"

    def _register(self, frame):
        sig = inspect.signature(type(self).__init__).parameters
        self.config = {k: frame[k] for k in sig if k not in ("self", "kwargs") and k in frame}

    @classmethod
    def from_config(cls, config, **kwargs):
        cfg = {k: v for k, v in dict(config).items() if not k.startswith("_")}
        cfg.update(kwargs)
        return cls(**cfg)

    @classmethod
    def load_config(cls, path, **kwargs):
        p = path if path.endswith(".json") else os.path.join(path, cls.config_name)
        with open(p) as f:
            return json.load(f)

    def save_config(self, save_directory, **kwargs):
        os.makedirs(save_directory, exist_ok=True)
        with open(os.path.join(save_directory, self.config_name), "w") as f:
            json.dump({"_class_name": type(self).__name__, **self.config}, f, indent=2)

    @property
    def device(self):
        return next(self.parameters()).device

    @property
    def dtype(self):
        return next(self.parameters()).dtype


class AMDModel_New(_ConfigModel):
    _supports_gradient_checkpointing = True

    def __init__(self, image_inchannel=4, image_height=32, image_width=32, video_frames=16, scheduler_num_step=1000,
                 use_filter=False, filter_num=0.4, use_grey=False, use_camera_down=False, use_regularizers=False,
                 use_motiontemporal=True, klloss_weight=0.005, use_mask=False, motion_type="plus", use_camera=True,
                 use_object=True, object_motion_token_num=12, object_motion_token_channel=128, object_enc_num_layers=8,
                 enc_nhead=8, enc_ndim=64, enc_dropout=0.0, motion_need_norm_out=False, camera_motion_token_num=12,
                 camera_motion_token_channel=128, camera_enc_num_layers=8, motion_token_num=12, motion_token_channel=128,
                 need_motion_transformer=False, motion_transformer_attn_head_dim=64, motion_transformer_attn_num_heads=16,
                 motion_transformer_num_layers=4, diffusion_model_type="default", diffusion_attn_head_dim=64,
                 diffusion_attn_num_heads=16, diffusion_out_channels=4, diffusion_num_layers=16, image_patch_size=2,
                 motion_patch_size=1, motion_drop_ratio=0.0, extract_motion_with_motion_transformer=False, **kwargs):
        super().__init__()
        self._register(locals())
        if diffusion_model_type not in ("spatial", "default"):
            raise IndexError(diffusion_model_type)  # model_AMD.py:1841-1842
        if diffusion_model_type == "default" and (use_camera or not use_object):
            # the reference's 'default' denoiser takes no camera stream (its forward has no such argument,
            # transformer.py:1631-1640) and dereferences the object tokens unconditionally (1656)
            raise NotImplementedError("diffusion_model_type='default' needs use_camera=False, use_object=True")
        if use_grey or need_motion_transformer or use_regularizers:
            raise NotImplementedError("use_grey / need_motion_transformer / use_regularizers are outside the hot path")
        self.num_step = scheduler_num_step
        self.scheduler = RectifiedFlow(num_steps=scheduler_num_step)
        self.diffusion_model_type = diffusion_model_type
        self.target_frame = video_frames
        self.camera_motion_token_channel = camera_motion_token_channel
        self.object_motion_token_channel = object_motion_token_channel
        self.motion_token_channel = motion_token_channel
        self.filter_num, self.use_camera, self.use_object = filter_num, use_camera, use_object
        self.use_filter, self.use_grey, self.use_camera_down = use_filter, use_grey, use_camera_down
        if use_camera:
            ch, cw = (image_height // 4, image_width // 4) if use_camera_down else (image_height, image_width)
            self.camera_motion_encoder = MotionEncoderLearnTokenOnlyTemporalTransformer(
                img_height=ch, img_width=cw, img_inchannel=image_inchannel, img_patch_size=image_patch_size,
                motion_token_num=camera_motion_token_num, motion_channel=camera_motion_token_channel,
                need_norm_out=motion_need_norm_out, video_frames=self.target_frame, num_attention_heads=enc_nhead,
                attention_head_dim=enc_ndim, num_layers=camera_enc_num_layers, dropout=enc_dropout, attention_bias=True)
        if use_object:
            self.object_motion_encoder = MotionEncoderLearnTokenTransformer(
                img_height=image_height, img_width=image_width, img_inchannel=image_inchannel,
                img_patch_size=image_patch_size, motion_token_num=object_motion_token_num,
                motion_channel=object_motion_token_channel, need_norm_out=motion_need_norm_out,
                num_attention_heads=enc_nhead, attention_head_dim=enc_ndim, num_layers=object_enc_num_layers,
                dropout=enc_dropout, attention_bias=True)
        if use_camera_down:
            self.camera_down = nn.Sequential(
                nn.Conv2d(in_channels=4, out_channels=16, kernel_size=3, stride=2, padding=1),
                nn.Conv2d(in_channels=16, out_channels=4, kernel_size=3, stride=1, padding=1),
                nn.MaxPool2d(kernel_size=2, stride=2, padding=0))
        dit_cls = AMDDiffusionTransformerModelImgSpatialTempMotion if diffusion_model_type == "spatial" else AMDDiffusionTransformerModelTempMotion
        self.diffusion_transformer = dit_cls(
            num_attention_heads=diffusion_attn_num_heads, attention_head_dim=diffusion_attn_head_dim,
            out_channels=diffusion_out_channels, num_layers=diffusion_num_layers, use_camera=use_camera,
            use_object=use_object, image_width=image_width, image_height=image_height, image_patch_size=image_patch_size,
            image_in_channels=image_inchannel * 2, motion_token_num=motion_token_num,
            camera_motion_in_channels=camera_motion_token_channel, object_motion_in_channels=object_motion_token_channel,
            motion_target_num_frame=self.target_frame)

    # ------------------------------------------------------------------ helpers
    def set_compute_dtype(self, dtype):
        """bf16 = the reference's autocast(bf16) semantics (bf16 GEMM/attention operands, fp32
        accumulation, fp32 LayerNorm statistics and residual stream); fp32 = the parity mode."""
        for m in self.modules():
            m.hv_compute_dtype = dtype
        return self

    def _camera_down(self, x):
        c0, c1 = self.camera_down[0], self.camera_down[1]
        f = lambda t: t.detach().float().contiguous()
        y = ops.conv3x3(x, f(c0.weight), f(c0.bias), 2)
        y = ops.conv3x3(y, f(c1.weight), f(c1.bias), 1)
        return ops.maxpool2(y)

    def _encode(self, video, ref_img, d_lo, camera_mask_ratio=None, object_mask_ratio=None):
        """shared front half of forward / sample / encode (model_AMD.py:1877-1944, 2037-2095, 2391-2455)"""
        n, t, c, h, w = video.shape
        assert video.shape == ref_img.shape, f"video.shape:{video.shape}should be equal to ref_img.shape:{ref_img.shape}"
        rv = torch.cat([ref_img, video], dim=1).float().contiguous()  # (n, 2t, c, h, w)
        cam = osrc = otgt = None
        if self.use_filter:
            # n t c h w -> n c t h w: the FFT volume is (2t, h, w) per (n, c)
            vol = rv.permute(0, 2, 1, 3, 4).contiguous()
            lo, _hi = freq_split_gauss(vol, d_lo, d_lo)
            lf_video = lo.permute(0, 2, 1, 3, 4)[:, t:].contiguous()  # low-frequency target half
            if self.use_camera:
                if self.use_camera_down:
                    lf = self._camera_down(lf_video.view(n * t, c, h, w))
                    lf_video = lf.view(n, t, *lf.shape[1:])
                cam = self.camera_motion_encoder(lf_video, camera_mask_ratio)
            if self.use_object:
                obj = self.object_motion_encoder(rv, object_mask_ratio)  # NOTE: unfiltered input (model_AMD.py:1926)
                osrc, otgt = obj[:, :t].flatten(0, 1).contiguous(), obj[:, t:].flatten(0, 1).contiguous()
                assert osrc.shape == otgt.shape
        else:
            # reference path without the filter only drives the camera encoder (model_AMD.py:1936-1944)
            x = rv
            if self.use_camera_down:
                y = self._camera_down(x.view(n * 2 * t, c, h, w))
                x = y.view(n, 2 * t, *y.shape[1:])
            cam = self.camera_motion_encoder(x[:, t:].contiguous(), camera_mask_ratio)
        return cam, osrc, otgt

    def prepare_timestep(self, batch_size, device, time_step=None):
        if time_step is not None:
            return time_step.to(device)
        return torch.randint(0, self.num_step + 1, (batch_size,)).to(device)

    def _dit(self, zi, zt, time_step, cam, osrc, otgt):
        dt_ = self.diffusion_transformer
        M = dt_.set_motion(zi.shape[0], zi.device, cam if self.use_camera else None,
                           osrc if self.use_object else None, otgt if self.use_object else None)
        pre = torch.empty_like(zt)
        dt_.run(M, zi, zt, time_step, pre=pre)
        return pre

    # ------------------------------------------------------------------ public API
    def forward(self, video, ref_img, video_grey=None, ref_img_grey=None, time_step=None, return_meta_info=False,
                camera_mask_ratio=None, object_mask_ratio=None):
        """Training-step forward (model_AMD.py:1844-2017): returns (pre, vel, {'loss','diff_loss','rec_loss'}).
        With autograd enabled and trainable parameters the result is differentiable (`loss.backward()` runs the
        hand-chained backward of hivae_b200.train on the sm_100a kernels); under torch.no_grad() it is the plain forward.
        Random draws are the reference's torch calls in the reference's order (mask ratios, masking noise, time steps,
        rectified-flow noise), so a seeded run sees the same random numbers."""
        n, t, c, h, w = video.shape
        device = video.device
        assert video.shape == ref_img.shape, f"video.shape:{video.shape}should be equal to ref_img.shape:{ref_img.shape}"
        if camera_mask_ratio is not None:
            camera_mask_ratio = (0.6 + 0.4 * torch.rand(1)).item() * camera_mask_ratio   # model_AMD.py:1857-1858
        if object_mask_ratio is not None:
            object_mask_ratio = (0.5 * torch.rand(1)).item() * object_mask_ratio           # model_AMD.py:1861-1862
        grad = torch.is_grad_enabled() and any(p.requires_grad for p in self.parameters())
        if not grad:
            with torch.no_grad():
                cam, osrc, otgt = self._encode(video, ref_img, 0.6, camera_mask_ratio, object_mask_ratio)
        else:
            cam_idx, obj_idx = self._mask_indices(n, t, device, camera_mask_ratio, object_mask_ratio)
        zi, zj = ref_img.flatten(0, 1).float().contiguous(), video.flatten(0, 1).float().contiguous()
        if time_step is None:
            time_step = self.prepare_timestep(batch_size=zj.shape[0], device=device)
        if self.diffusion_model_type != "default":  # 'spatial' mode redraws one step per clip (model_AMD.py:1962-1964)
            time_step = self.prepare_timestep(batch_size=n, device=device).repeat_interleave(t)
        zt, vel = self.scheduler.get_train_tuple(z1=zj, time_step=time_step)
        zt = zt.contiguous()
        if grad:
            from . import train
            pre = train.forward_with_grad(self, video, ref_img, time_step.float(), zt, cam_idx, obj_idx)
            diff_loss = train.mse(pre, vel)
        else:
            with torch.no_grad():
                pre = self._dit(zi, zt, time_step.float(), cam, osrc, otgt)
                diff_loss = l2(pre, vel)
        with torch.no_grad():
            rec_loss = l2(self.scheduler.get_target_with_zt_vel(zt, pre.detach(), time_step), zj)
        if return_meta_info:
            return {"zi": zi, "zj": zj, "zt": zt, "pre": pre, "time_step": time_step}
        return pre, vel, {"loss": diff_loss, "diff_loss": diff_loss, "rec_loss": rec_loss}

    def _mask_indices(self, n, t, device, camera_mask_ratio, object_mask_ratio):
        """kept-token indices of the two encoders' random_masking, drawn in the reference's order (camera encoder first,
        model_AMD.py:1920-1926) with the reference's calls (torch.rand on the device, argsort)"""
        from .transformer import _mask_idx as mask_indices
        from .transformer import mask_noise
        cam_idx = obj_idx = None
        if camera_mask_ratio is not None and self.use_camera:
            cam_idx = mask_indices(mask_noise(n, self.camera_motion_encoder.img_token_len, device), camera_mask_ratio)
        if object_mask_ratio is not None and self.use_object and self.use_filter:
            obj_idx = mask_indices(mask_noise(n * 2 * t, self.object_motion_encoder.img_token_len, device), object_mask_ratio)
        return cam_idx, obj_idx

    @torch.no_grad()
    def encode(self, video, ref_img, video_grey=None, ref_img_grey=None, camera_mask_ratio=None, object_mask_ratio=None):
        """model_AMD.py:2379-2462 -> (camera_target_motion (n,t,s,dc), object_source_motion, object_target_motion (n*t,l,do))."""
        return self._encode(video, ref_img, 0.6, camera_mask_ratio, object_mask_ratio)

    @torch.no_grad()
    def decode(self, video, ref_img, camera_target_motion, object_source_motion, object_target_motion, time_step=None,
               start_step=None, sample_step=50, return_meta_info=False):
        """Euler sampling loop (model_AMD.py:2464-2536): (zi, zt, zj), each (n,t,c,h,w)."""
        if start_step is None:
            raise ValueError("decode() needs start_step (the reference raises a TypeError here, model_AMD.py:2476)")
        n, t = video.shape[:2]
        zi, zj = ref_img.flatten(0, 1).float().contiguous(), video.flatten(0, 1).float().contiguous()
        zi, zt, zj, meta = self._euler(zi, zj, camera_target_motion if self.use_camera else None,
                                       object_source_motion if self.use_object else None,
                                       object_target_motion if self.use_object else None,
                                       start_step, sample_step, return_meta_info)
        shp = (n, t) + tuple(zi.shape[1:])
        zi, zt, zj = zi.view(shp), zt.view(shp), zj.view(shp)
        if return_meta_info:
            return dict(meta, zi=zi, zj=zj, sample=zt)
        return zi, zt, zj

    def _euler(self, zi, zj, cam, osrc, otgt, start_step, sample_step, return_meta_info):
        """noise zj to start_step, then the Euler loop shared by decode / sample_cross / sample_with_refimg_motion
        (model_AMD.py:2124-2160, 2326-2356, 2745-2768); flat (n*t, c, h, w) tensors in and out"""
        time_step = torch.ones((zj.shape[0],), device=zj.device) * start_step
        zt, _vel = self.scheduler.get_train_tuple(z1=zj, time_step=time_step)
        zt = zt.contiguous().clone()
        step_seq = np.linspace(0, start_step, num=sample_step + 1, endpoint=True, dtype=int)
        step_seq = list(reversed(step_seq[1:]))
        dt = 1.0 / sample_step
        dit = self.diffusion_transformer
        M = dit.set_motion(zi.shape[0], zi.device, cam, osrc, otgt)
        pre_cache, sample_cache = [], []
        ts = torch.empty(zt.shape[0], dtype=torch.float32, device=zt.device)
        for i in step_seq:
            ts.fill_(float(i))
            pre = torch.empty_like(zt) if return_meta_info else None
            dit.run(M, zi, zt, ts, pre=pre, zt=zt, dt=dt)  # zt += dt * pre fused into the head kernel
            if return_meta_info:
                pre_cache.append(pre)
                sample_cache.append(zt.clone())
        return zi, zt, zj, {"pre_cache": pre_cache, "sample_cache": sample_cache, "step_seq": step_seq}

    @torch.no_grad()
    def sample_cross(self, video_1, video_2, ref_img, video_grey_1=None, video_grey_2=None, ref_img_grey=None, sample_step=50,
                     object_mask_ratio=None, camera_mask_ratio=None, start_step=None, return_meta_info=False):
        """model_AMD.py:2198-2377: camera motion from video_1 (low band, cut-off 0.5 over ITS t-frame volume, 2219-2221),
        noised ground truth and reference image from video_2.  The reference also encodes video_2's object motion
        (2256-2260) but does not hand it to the denoiser (2349-2351); that unused encoder pass is not run here."""
        n, t, c, h, w = video_1.shape
        if start_step is None:
            start_step = self.scheduler.num_step
        assert start_step <= self.scheduler.num_step, "start_step cant be larger than scheduler.num_step"
        if not (self.use_filter and self.use_camera) or self.diffusion_model_type != "spatial":
            raise NotImplementedError("sample_cross needs use_filter, use_camera and the 'spatial' denoiser "
                                      "(the reference's other branches read undefined names, model_AMD.py:2264-2272)")
        lo, _hi = freq_split_gauss(video_1.float().permute(0, 2, 1, 3, 4).contiguous(), 0.5, 0.5)
        lf_video = lo.permute(0, 2, 1, 3, 4).contiguous()
        if self.use_camera_down:
            lf = self._camera_down(lf_video.view(n * t, c, h, w))
            lf_video = lf.view(n, t, *lf.shape[1:])
        cam = self.camera_motion_encoder(lf_video, camera_mask_ratio)
        zi, zj = ref_img.flatten(0, 1).float().contiguous(), video_2.flatten(0, 1).float().contiguous()
        zi, zt, zj, meta = self._euler(zi, zj, cam, None, None, start_step, sample_step, return_meta_info)
        shp = (n, t) + tuple(zi.shape[1:])
        zi, zt, zj = zi.view(shp), zt.view(shp), zj.view(shp)
        if return_meta_info:
            return dict(meta, zi=zi, zj=zj, sample=zt)
        return zi, zt, zj

    @torch.no_grad()
    def sample_with_refimg_motion(self, ref_img, camera_target_motion=None, object_target_motion=None, sample_step=10,
                                  return_meta_info=False, ref_img_grey=None):
        """model_AMD.py:2701-2785: generate from given target motions; the object SOURCE motion comes from the high
        band (cut-off 0.6 over the t-frame volume) of `ref_img_grey`.  Returns (zi, zt, zj) with zj left flat
        (n*t, c, h, w) as the reference does (2774-2775 rearrange only zi and zt)."""
        if return_meta_info:
            raise NameError("name 'target_motion' is not defined (the reference's meta-info branch, model_AMD.py:2782)")
        n, t, c, h, w = ref_img.shape
        start_step = self.scheduler.num_step
        _lo, hi = freq_split_gauss(ref_img_grey.float().permute(0, 2, 1, 3, 4).contiguous(), 0.6, 0.6)
        hf_refimg = hi.permute(0, 2, 1, 3, 4).contiguous()
        osrc = self.object_motion_encoder(hf_refimg).flatten(0, 1).contiguous()
        assert osrc.shape == object_target_motion.shape, \
            f"object_source_motion.shape {osrc.shape} != object_target_motion.shape {object_target_motion.shape}"
        zi = ref_img.flatten(0, 1).float().contiguous()
        zi, zt, zj, _meta = self._euler(zi, zi, camera_target_motion, osrc, object_target_motion, start_step, sample_step, False)
        shp = (n, t) + tuple(zi.shape[1:])
        return zi.view(shp), zt.view(shp), zj

    @torch.no_grad()
    def sample(self, video, ref_img, video_grey=None, ref_img_grey=None, sample_step=50, object_mask_ratio=None,
               camera_mask_ratio=None, start_step=None, return_meta_info=False):
        """encode + decode (model_AMD.py:2020-2195)."""
        if start_step is None:
            start_step = self.scheduler.num_step
        assert start_step <= self.scheduler.num_step, "start_step cant be larger than scheduler.num_step"
        cam, osrc, otgt = self._encode(video, ref_img, 0.6, camera_mask_ratio, object_mask_ratio)
        return self.decode(video, ref_img, cam, osrc, otgt, start_step=start_step, sample_step=sample_step,
                           return_meta_info=return_meta_info)


    @torch.no_grad()
    def sample_graphed(self, video, ref_img, sample_step=50, start_step=None):
        """sample() replayed from ONE CUDA graph per (batch shape, sample_step, start_step, weights): encode + the whole Euler
        loop are ~270 launches per step through ctypes, which bounds small batches by the host (1-2 clips per GPU: SURVEY 7
        step 6).  Everything in sample() already runs on the device without host synchronisation (the start noise comes from
        torch's graph-safe CUDA generator), so the capture is the eager code unchanged.  Returns (zi, zt, zj) like sample();
        the tensors are owned by the graph and overwritten by the next call with the same key."""
        from . import engine
        dev = video.device
        key = (tuple(video.shape), tuple(ref_img.shape), video.dtype, ref_img.dtype, int(sample_step), start_step,
               engine.params_version(self))
        cache = self.__dict__.setdefault("_sample_graphs", {})
        ent = cache.get(key)
        if ent is None:
            for k in [k for k in cache if k[:6] == key[:6]]:
                del cache[k]   # same shapes, older weights: the graph would read stale prepared weight copies
            gv, gr = video.clone(), ref_img.clone()
            rng = torch.cuda.get_rng_state(dev)
            side = torch.cuda.Stream(dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):  # eager warm-up: weight copies, workspaces and kernel attributes exist before the capture
                self.sample(gv, gr, sample_step=sample_step, start_step=start_step)
            torch.cuda.current_stream(dev).wait_stream(side)
            torch.cuda.synchronize(dev)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                out = self.sample(gv, gr, sample_step=sample_step, start_step=start_step)
            torch.cuda.set_rng_state(rng, dev)
            ent = cache[key] = (graph, gv, gr, out)
        graph, gv, gr, out = ent
        gv.copy_(video, non_blocking=True)
        gr.copy_(ref_img, non_blocking=True)
        graph.replay()
        return out

    # ------------------------------------------------------------------ small helpers of the reference class
    def prepare_encoder_input(self, video):
        """model_AMD.py:2804-2814: consecutive frame pairs stacked on the channel axis, (b*(t-1), 2c, h, w)"""
        assert len(video.shape) == 5, f"only support video data : 5D tensor , but got {video.shape}"
        return torch.cat([video[:, :-1], video[:, 1:]], dim=2).flatten(0, 1)

    @torch.no_grad()
    def unpatchify(self, x, patch_size):
        """model_AMD.py:2816-2830: tokens (N, S, p*p*C) with (p, q, c) feature order -> (N, C, H, W)"""
        return ops.unpatchify(x, patch_size, order=0).to(x.dtype)

    def reset_infer_num_frame(self, num):
        """model_AMD.py:2832-2838 (the reference then reads a non-existent `use_motiontemporal`; the
        frame count of the denoiser is what matters for AMDModel_New and is set before that)"""
        old_num = self.diffusion_transformer.target_frame
        self.diffusion_transformer.target_frame = num
        print(f"* Reset infer frame from {old_num} to {self.diffusion_transformer.target_frame} *")


class AMDModel(_ConfigModel):
    """The older two-encoder model (model_AMD.py:33-594): object motion from the HIGH band and camera motion from the
    LOW band of the grey clip, both over all 2t frames, one single-stream denoiser ('default' ->
    AMDDiffusionTransformerModel, 'spatial' -> AMDDiffusionTransformerModelImgSpatial).  Built for the runnable
    configurations: use_regularizers=False, need_motion_transformer=False; use_filter needs use_grey (the reference reads
    an undefined name otherwise, 291 / 462) and use_camera_down needs use_filter (314 / 472)."""
    _supports_gradient_checkpointing = True

    def __init__(self, image_inchannel=4, image_height=32, image_width=32, video_frames=16, scheduler_num_step=1000,
                 use_filter=False, filter_num=0.4, high_filter_num=0.6, use_grey=False, use_camera_down=False,
                 use_regularizers=False, use_motiontemporal=True, klloss_weight=0.005, use_mask=False, motion_type="plus",
                 object_motion_token_num=12, object_motion_token_channel=128, object_enc_num_layers=8, enc_nhead=8,
                 enc_ndim=64, enc_dropout=0.0, motion_need_norm_out=False, camera_motion_token_num=12,
                 camera_motion_token_channel=128, camera_enc_num_layers=8, motion_token_num=12, motion_token_channel=128,
                 need_motion_transformer=False, motion_transformer_attn_head_dim=64, motion_transformer_attn_num_heads=16,
                 motion_transformer_num_layers=4, diffusion_model_type="default", diffusion_attn_head_dim=64,
                 diffusion_attn_num_heads=16, diffusion_out_channels=4, diffusion_num_layers=16, image_patch_size=2,
                 motion_patch_size=1, motion_drop_ratio=0.0, extract_motion_with_motion_transformer=False, **kwargs):
        super().__init__()
        self._register(locals())
        if use_regularizers or need_motion_transformer:
            raise NotImplementedError("use_regularizers / need_motion_transformer are out of scope (false in every shipped script)")
        self.num_step = scheduler_num_step
        self.scheduler = RectifiedFlow(num_steps=scheduler_num_step)
        self.need_motion_transformer, self.extract_motion_with_motion_transformer = need_motion_transformer, extract_motion_with_motion_transformer
        self.diffusion_model_type, self.target_frame, self.use_motiontemporal = diffusion_model_type, video_frames, use_motiontemporal
        self.camera_motion_token_channel, self.object_motion_token_channel = camera_motion_token_channel, object_motion_token_channel
        self.motion_token_channel, self.filter_num, self.high_filter_num = motion_token_channel, filter_num, high_filter_num
        self.use_filter, self.use_grey, self.use_camera_down, self.use_mask = use_filter, use_grey, use_camera_down, use_mask
        self.use_regularizers, self.klloss_weight = use_regularizers, klloss_weight
        if camera_motion_token_channel != motion_token_channel:
            self.camera_motion_map = nn.Linear(camera_motion_token_channel, motion_token_channel)
        if object_motion_token_channel != motion_token_channel:
            self.object_motion_map = nn.Linear(object_motion_token_channel, motion_token_channel)
        enc = dict(img_inchannel=image_inchannel, img_patch_size=image_patch_size, need_norm_out=motion_need_norm_out,
                   num_attention_heads=enc_nhead, attention_head_dim=enc_ndim, dropout=enc_dropout, attention_bias=True)
        obj = dict(img_height=image_height, img_width=image_width, motion_token_num=object_motion_token_num,
                   motion_channel=object_motion_token_channel, num_layers=object_enc_num_layers)
        cam = dict(motion_token_num=camera_motion_token_num, motion_channel=camera_motion_token_channel, num_layers=camera_enc_num_layers)
        if use_motiontemporal:
            self.object_motion_encoder = MotionEncoderLearnTokenTemporalTransformer(video_frames=video_frames, **obj, **enc)
            self.camera_motion_encoder = MotionEncoderLearnTokenTemporalTransformer(img_height=image_height, img_width=image_width,
                                                                                    video_frames=video_frames, **cam, **enc)
        else:
            self.object_motion_encoder = MotionEncoderLearnTokenTransformer(**obj, **enc)
            self.camera_motion_encoder = MotionEncoderLearnTokenTransformer(img_height=image_height // 4, img_width=image_width // 4,
                                                                            **cam, **enc)  # sized for the camera_down output (186-187)
        if use_camera_down:
            self.camera_down = nn.Sequential(nn.Conv2d(4, 16, kernel_size=3, stride=2, padding=1),
                                             nn.Conv2d(16, 4, kernel_size=3, stride=1, padding=1), nn.MaxPool2d(2, 2, 0))
        dit = dict(num_attention_heads=diffusion_attn_num_heads, attention_head_dim=diffusion_attn_head_dim,
                   out_channels=diffusion_out_channels, num_layers=diffusion_num_layers, motion_type=motion_type,
                   image_width=image_width, image_height=image_height, image_patch_size=image_patch_size,
                   image_in_channels=image_inchannel * 2, motion_token_num=motion_token_num, motion_in_channels=motion_token_channel)
        if diffusion_model_type == "default":
            self.diffusion_transformer = AMDDiffusionTransformerModel(**dit)
        elif diffusion_model_type == "spatial":
            self.diffusion_transformer = AMDDiffusionTransformerModelImgSpatial(motion_target_num_frame=video_frames, **dit)
        elif diffusion_model_type == "dual":
            raise NotImplementedError("AMDDiffusionTransformerModelDualStream is not runnable in the reference (modules.py:1250)")
        else:
            raise IndexError

    set_compute_dtype = AMDModel_New.set_compute_dtype
    prepare_timestep = AMDModel_New.prepare_timestep
    _camera_down = AMDModel_New._camera_down

    def _motions(self, video, ref_img, video_grey, ref_img_grey, camera_mask, mask_ratio):
        """shared front half of forward / sample (model_AMD.py:272-343, 441-500) -> four (n*t, l, d) motion tensors"""
        n, t, c, h, w = video.shape
        rv = torch.cat([ref_img, video], dim=1).float().contiguous()
        if self.use_filter:
            if not self.use_grey:
                raise NameError("name 'refimg_and_video_grey' is not defined")  # model_AMD.py:291, 462
            grey = torch.cat([ref_img_grey, video_grey], dim=1).float().permute(0, 2, 1, 3, 4).contiguous()
            lo, _ = freq_split_gauss(grey, self.filter_num, self.filter_num)
            _, hi = freq_split_gauss(grey, self.high_filter_num, self.high_filter_num)
            lf, hf = lo.permute(0, 2, 1, 3, 4).contiguous(), hi.permute(0, 2, 1, 3, 4).contiguous()
            if self.use_mask and camera_mask is not None:
                lf = (lf * camera_mask).contiguous()
            if self.use_camera_down:
                y = self._camera_down(lf.view(n * 2 * t, c, h, w))
                lf = y.view(n, 2 * t, *y.shape[1:])
            obj, cam = self.object_motion_encoder(hf, mask_ratio), self.camera_motion_encoder(lf, mask_ratio)
        else:
            if self.use_camera_down:
                raise UnboundLocalError("LF_refimg_and_video")  # model_AMD.py:314, 472
            obj, cam = self.object_motion_encoder(rv, mask_ratio), self.camera_motion_encoder(rv, mask_ratio)
        osrc, otgt = obj[:, :t].flatten(0, 1).contiguous(), obj[:, t:].flatten(0, 1).contiguous()
        csrc, ctgt = cam[:, :t].flatten(0, 1).contiguous(), cam[:, t:].flatten(0, 1).contiguous()
        assert osrc.shape == otgt.shape and csrc.shape == ctgt.shape
        lin = lambda m, x: torch.nn.functional.linear(x, m.weight, m.bias) if not x.is_cuda else _device_linear(m, x)
        if self.camera_motion_token_channel != self.motion_token_channel:
            csrc, ctgt = lin(self.camera_motion_map, csrc), lin(self.camera_motion_map, ctgt)
        if self.object_motion_token_channel != self.motion_token_channel:
            osrc, otgt = lin(self.object_motion_map, osrc), lin(self.object_motion_map, otgt)
        return cam, obj, csrc, ctgt, osrc, otgt

    def _dit(self, zi, zt, time_step, m):
        _, _, csrc, ctgt, osrc, otgt = m
        return self.diffusion_transformer(ctgt, (zi, zt), time_step, camera_motion_source_hidden_states=csrc,
                                          object_motion_source_hidden_states=osrc, object_motion_target_hidden_states=otgt)

    @torch.no_grad()
    def forward(self, video, ref_img, video_grey=None, ref_img_grey=None, camera_mask=None, time_step=None,
                return_meta_info=False, mask_ratio=None):
        """model_AMD.py:258-423 (inference kernels; no autograd yet) -> (pre, vel, {'loss','diff_loss','rec_loss'})"""
        n, t, c, h, w = video.shape
        assert video.shape == ref_img.shape, f"video.shape:{video.shape}should be equal to ref_img.shape:{ref_img.shape}"
        if mask_ratio is not None:
            mask_ratio = torch.rand(1).item() * mask_ratio      # model_AMD.py:271-272
        m = self._motions(video, ref_img, video_grey, ref_img_grey, camera_mask, mask_ratio)
        zi, zj = ref_img.flatten(0, 1).float().contiguous(), video.flatten(0, 1).float().contiguous()
        if time_step is None:
            time_step = self.prepare_timestep(batch_size=zj.shape[0], device=video.device)
        if self.diffusion_model_type != "default":  # one step per clip, redrawn even when a time_step was given (370-372)
            time_step = self.prepare_timestep(batch_size=n, device=video.device).repeat_interleave(t)
        zt, vel = self.scheduler.get_train_tuple(z1=zj, time_step=time_step)
        zt = zt.contiguous()
        pre = self._dit(zi, zt, time_step.float(), m)
        diff_loss = l2(pre, vel)
        rec_loss = l2(self.scheduler.get_target_with_zt_vel(zt, pre, time_step), zj)
        if return_meta_info:
            return {"camera_motion": m[0], "object_motion": m[1], "zi": zi, "zj": zj, "zt": zt, "pre": pre, "time_step": time_step}
        return pre, vel, {"loss": diff_loss, "diff_loss": diff_loss, "rec_loss": rec_loss}

    @torch.no_grad()
    def sample(self, video, ref_img, video_grey=None, ref_img_grey=None, camera_mask=None, sample_step=50, mask_ratio=None,
               start_step=None, return_meta_info=False):
        """model_AMD.py:426-594 -> (zi, zt, zj), each (n, t, c, h, w).  (The reference's camera_mask is not applied in
        sample either, 441-478; its meta-info branch reads an undefined `target_motion`, 590.)"""
        if return_meta_info:
            raise NameError("name 'target_motion' is not defined")
        n, t = video.shape[:2]
        if start_step is None:
            start_step = self.scheduler.num_step
        assert start_step <= self.scheduler.num_step, "start_step cant be larger than scheduler.num_step"
        m = self._motions(video, ref_img, video_grey, ref_img_grey, None, mask_ratio)  # the ratio is used as given (436-438)
        zi, zj = ref_img.flatten(0, 1).float().contiguous(), video.flatten(0, 1).float().contiguous()
        zt, _ = self.scheduler.get_train_tuple(z1=zj, time_step=torch.ones((zj.shape[0],), device=zj.device) * start_step)
        zt = zt.contiguous().clone()
        step_seq = list(reversed(np.linspace(0, start_step, num=sample_step + 1, endpoint=True, dtype=int)[1:]))
        dt = 1.0 / sample_step
        ts = torch.empty(zt.shape[0], dtype=torch.float32, device=zt.device)
        for i in step_seq:
            ts.fill_(float(i))
            ops.axpy(dt, self._dit(zi, zt, ts, m).contiguous(), zt)  # zt += dt * pre
        shp = (n, t) + tuple(zi.shape[1:])
        return zi.view(shp), zt.view(shp), zj.view(shp)


class AMDModel_Camera(_ConfigModel):
    """The camera-only model (model_AMD.py:894-1661): ONE motion encoder over the LOW band of [ref | video] (the grey clip when
    use_grey), camera source / target tokens only, a single-stream denoiser; the target of the flow is the low band itself and
    the reference image is `ref_img_grey` (1192-1193).  Built for the runnable configuration: use_filter (the reference reads
    undefined names without it, 1136-1141), use_motiontemporal=False (the temporal branch encodes t frames and then splits
    them into [:t] / [t:], 1147-1163), no camera_down (the encoder is sized for the full image, 960-972), motion_type
    'decouple' ('plus' adds object tokens that this model never passes)."""
    _supports_gradient_checkpointing = True

    def __init__(self, image_inchannel=4, image_height=32, image_width=32, video_frames=16, scheduler_num_step=1000,
                 use_filter=False, filter_num=0.4, use_grey=False, use_camera_down=False, use_regularizers=False,
                 use_motiontemporal=True, klloss_weight=0.005, use_mask=True, motion_type="plus", object_motion_token_num=12,
                 object_motion_token_channel=128, object_enc_num_layers=8, enc_nhead=8, enc_ndim=64, enc_dropout=0.0,
                 motion_need_norm_out=False, camera_motion_token_num=12, camera_motion_token_channel=128, camera_enc_num_layers=8,
                 motion_token_num=12, motion_token_channel=128, need_motion_transformer=False, motion_transformer_attn_head_dim=64,
                 motion_transformer_attn_num_heads=16, motion_transformer_num_layers=4, diffusion_model_type="default",
                 diffusion_attn_head_dim=64, diffusion_attn_num_heads=16, diffusion_out_channels=4, diffusion_num_layers=16,
                 image_patch_size=2, motion_patch_size=1, motion_drop_ratio=0.0, extract_motion_with_motion_transformer=False, **kwargs):
        super().__init__()
        self._register(locals())
        if use_regularizers or need_motion_transformer:
            raise NotImplementedError("use_regularizers / need_motion_transformer are out of scope (false in every shipped script)")
        if use_motiontemporal or use_camera_down or not use_filter or motion_type != "decouple":
            raise NotImplementedError("AMDModel_Camera is runnable in the reference only with use_filter=True, use_motiontemporal=False, "
                                      "use_camera_down=False, motion_type='decouple' (model_AMD.py:960-972, 1136-1163)")
        self.num_step = scheduler_num_step
        self.scheduler = RectifiedFlow(num_steps=scheduler_num_step)
        self.diffusion_model_type, self.target_frame = diffusion_model_type, video_frames
        self.camera_motion_token_channel, self.motion_token_channel = camera_motion_token_channel, motion_token_channel
        self.filter_num, self.use_filter, self.use_grey, self.use_mask = filter_num, use_filter, use_grey, use_mask
        if camera_motion_token_channel != motion_token_channel:
            self.camera_motion_map = nn.Linear(camera_motion_token_channel, motion_token_channel)
        self.camera_motion_encoder = MotionEncoderLearnTokenTransformer(
            img_height=image_height, img_width=image_width, img_inchannel=image_inchannel, img_patch_size=image_patch_size,
            motion_token_num=camera_motion_token_num, motion_channel=camera_motion_token_channel, need_norm_out=motion_need_norm_out,
            num_attention_heads=enc_nhead, attention_head_dim=enc_ndim, num_layers=camera_enc_num_layers, dropout=enc_dropout,
            attention_bias=True)
        dit = dict(num_attention_heads=diffusion_attn_num_heads, attention_head_dim=diffusion_attn_head_dim,
                   out_channels=diffusion_out_channels, num_layers=diffusion_num_layers, motion_type=motion_type,
                   image_width=image_width, image_height=image_height, image_patch_size=image_patch_size,
                   image_in_channels=image_inchannel * 2, motion_token_num=motion_token_num, motion_in_channels=motion_token_channel)
        if diffusion_model_type == "default":
            self.diffusion_transformer = AMDDiffusionTransformerModel(**dit)
        elif diffusion_model_type == "spatial":
            self.diffusion_transformer = AMDDiffusionTransformerModelImgSpatial(motion_target_num_frame=video_frames, **dit)
        else:
            raise IndexError

    set_compute_dtype = AMDModel_New.set_compute_dtype
    prepare_timestep = AMDModel_New.prepare_timestep

    def _motions(self, video, ref_img, video_grey, ref_img_grey, camera_mask, mask_ratio):
        n, t, c, h, w = video.shape
        src = torch.cat([ref_img_grey, video_grey], dim=1) if self.use_grey else torch.cat([ref_img, video], dim=1)
        lo, _ = freq_split_gauss(src.float().permute(0, 2, 1, 3, 4).contiguous(), self.filter_num, self.filter_num)
        lf = lo.permute(0, 2, 1, 3, 4).contiguous()
        gt = lf[:, t:].contiguous()
        if self.use_mask:
            lf = (lf * camera_mask).contiguous()
        cam = self.camera_motion_encoder(lf, mask_ratio)
        csrc, ctgt = cam[:, :t].flatten(0, 1).contiguous(), cam[:, t:].flatten(0, 1).contiguous()
        assert csrc.shape == ctgt.shape
        if self.camera_motion_token_channel != self.motion_token_channel:
            csrc, ctgt = _device_linear(self.camera_motion_map, csrc), _device_linear(self.camera_motion_map, ctgt)
        return cam, csrc, ctgt, gt.flatten(0, 1).contiguous()

    def _dit(self, zi, zt, time_step, csrc, ctgt):
        return self.diffusion_transformer(ctgt, (zi, zt), time_step, camera_motion_source_hidden_states=csrc)

    @torch.no_grad()
    def forward(self, video, ref_img, video_grey=None, ref_img_grey=None, camera_mask=None, time_step=None,
                return_meta_info=False, mask_ratio=None):
        """model_AMD.py:1086-1224 -> (pre, vel, {'loss','diff_loss','rec_loss'})"""
        n, t, c, h, w = video.shape
        assert video.shape == ref_img.shape, f"video.shape:{video.shape}should be equal to ref_img.shape:{ref_img.shape}"
        if mask_ratio is not None:
            mask_ratio = torch.rand(1).item() * mask_ratio
        cam, csrc, ctgt, zj = self._motions(video, ref_img, video_grey, ref_img_grey, camera_mask, mask_ratio)
        zi = ref_img_grey.flatten(0, 1).float().contiguous()
        if time_step is None:
            time_step = self.prepare_timestep(batch_size=zj.shape[0], device=video.device)
        if self.diffusion_model_type != "default":
            time_step = self.prepare_timestep(batch_size=n, device=video.device).repeat_interleave(t)
        zt, vel = self.scheduler.get_train_tuple(z1=zj, time_step=time_step)
        zt = zt.contiguous()
        pre = self._dit(zi, zt, time_step.float(), csrc, ctgt)
        diff_loss = l2(pre, vel)
        rec_loss = l2(self.scheduler.get_target_with_zt_vel(zt, pre, time_step), zj)
        if return_meta_info:
            return {"camera_motion": cam, "zi": zi, "zj": zj, "zt": zt, "pre": pre, "time_step": time_step}
        return pre, vel, {"loss": diff_loss, "diff_loss": diff_loss, "rec_loss": rec_loss}

    @torch.no_grad()
    def sample(self, video, ref_img, video_grey=None, ref_img_grey=None, camera_mask=None, sample_step=50, mask_ratio=None,
               start_step=None, return_meta_info=False):
        """model_AMD.py:1226-1365 -> (zi, zt, zj), each (n, t, c, h, w); zj is the low band the flow reconstructs"""
        n, t = video.shape[:2]
        if start_step is None:
            start_step = self.scheduler.num_step
        assert start_step <= self.scheduler.num_step, "start_step cant be larger than scheduler.num_step"
        _cam, csrc, ctgt, zj = self._motions(video, ref_img, video_grey, ref_img_grey, camera_mask, mask_ratio)
        zi = ref_img_grey.flatten(0, 1).float().contiguous()
        zt, _ = self.scheduler.get_train_tuple(z1=zj, time_step=torch.ones((zj.shape[0],), device=zj.device) * start_step)
        zt = zt.contiguous().clone()
        step_seq = list(reversed(np.linspace(0, start_step, num=sample_step + 1, endpoint=True, dtype=int)[1:]))
        dt = 1.0 / sample_step
        ts = torch.empty(zt.shape[0], dtype=torch.float32, device=zt.device)
        pre_cache, sample_cache = [], []
        for i in step_seq:
            ts.fill_(float(i))
            pre = self._dit(zi, zt, ts, csrc, ctgt).contiguous()
            ops.axpy(dt, pre, zt)
            if return_meta_info:
                pre_cache.append(pre)
                sample_cache.append(zt.clone())
        shp = (n, t) + tuple(zi.shape[1:])
        zi, zt, zj = zi.view(shp), zt.view(shp), zj.view(shp)
        if return_meta_info:
            return {"zi": zi, "zj": zj, "sample": zt, "pre_cache": pre_cache, "sample_cache": sample_cache, "step_seq": step_seq, "motion": ctgt}
        return zi, zt, zj


class AMDModel_Rec(_ConfigModel):
    """The reconstruction model (model_AMD.py:2840-3125): object + camera motion from two MotionEncoderLearnTokenTransformer
    over [ref | video], added; one pass of AMDReconstructTransformerModel on [zi | learnable zt_token]; loss = l2(pre, video).
    `sample` is not runnable in the reference (it reads a non-existent `self.motion_encoder`, 3020) and raises the same
    AttributeError here."""
    _supports_gradient_checkpointing = True

    def __init__(self, image_inchannel=4, image_height=32, image_width=32, video_frames=16, scheduler_num_step=1000,
                 motion_token_num=12, motion_token_channel=128, enc_num_layers=8, enc_nhead=8, enc_ndim=64, enc_dropout=0.0,
                 motion_need_norm_out=True, need_motion_transformer=False, motion_transformer_attn_head_dim=64,
                 motion_transformer_attn_num_heads=16, motion_transformer_num_layers=4, diffusion_attn_head_dim=64,
                 diffusion_attn_num_heads=16, diffusion_out_channels=4, diffusion_num_layers=16, image_patch_size=2,
                 motion_patch_size=1, motion_drop_ratio=0.0, **kwargs):
        super().__init__()
        self._register(locals())
        if need_motion_transformer:
            raise NotImplementedError("need_motion_transformer is out of scope (false in every shipped script)")
        self.num_step = scheduler_num_step
        self.scheduler = RectifiedFlow(num_steps=scheduler_num_step)
        self.need_motion_transformer = need_motion_transformer
        self.zt_token = nn.Parameter(torch.randn(1, image_inchannel, image_height, image_width) * 0.02)
        enc = dict(img_height=image_height, img_width=image_width, img_inchannel=image_inchannel, img_patch_size=image_patch_size,
                   motion_token_num=motion_token_num, motion_channel=motion_token_channel, need_norm_out=motion_need_norm_out,
                   num_attention_heads=enc_nhead, attention_head_dim=enc_ndim, num_layers=enc_num_layers, dropout=enc_dropout,
                   attention_bias=True)
        self.object_motion_encoder = MotionEncoderLearnTokenTransformer(**enc)
        self.camera_motion_encoder = MotionEncoderLearnTokenTransformer(**enc)
        self.transformer = AMDReconstructTransformerModel(
            num_attention_heads=diffusion_attn_num_heads, attention_head_dim=diffusion_attn_head_dim, out_channels=diffusion_out_channels,
            num_layers=diffusion_num_layers, image_width=image_width, image_height=image_height, image_patch_size=image_patch_size,
            image_in_channels=image_inchannel * 2, motion_token_num=motion_token_num, motion_in_channels=motion_token_channel)

    set_compute_dtype = AMDModel_New.set_compute_dtype

    @torch.no_grad()
    def forward(self, video, ref_img, time_step=None, return_meta_info=False):
        """model_AMD.py:2940-3007 -> (pre, zj, {'loss', 'rec_loss'})"""
        n, t, c, h, w = video.shape
        assert video.shape == ref_img.shape, f"video.shape:{video.shape}should be equal to ref_img.shape:{ref_img.shape}"
        rv = torch.cat([ref_img, video], dim=1).float().contiguous()
        obj = self.object_motion_encoder(rv)
        osrc, otgt = obj[:, :t].flatten(0, 1), obj[:, t:].flatten(0, 1)
        assert osrc.shape == otgt.shape
        cam = self.camera_motion_encoder(rv)
        csrc, ctgt = cam[:, :t].flatten(0, 1), cam[:, t:].flatten(0, 1)
        assert csrc.shape == ctgt.shape
        zi, zj = ref_img.flatten(0, 1).float().contiguous(), video.flatten(0, 1).float().contiguous()
        zt = self.zt_token.detach().float().repeat(zj.shape[0], 1, 1, 1).contiguous()
        pre = self.transformer((osrc + csrc).contiguous(), (otgt + ctgt).contiguous(), (zi, zt))
        rec_loss = l2(pre, zj)
        if return_meta_info:
            return {"camera_motion": cam, "object_motion": obj, "zi": zi, "zj": zj, "zt": zt, "pre": pre, "time_step": time_step}
        return pre, zj, {"loss": rec_loss, "rec_loss": rec_loss}

    def sample(self, video, ref_img, sample_step=50, start_step=None, return_meta_info=False):
        raise AttributeError("'AMDModel_Rec' object has no attribute 'motion_encoder'")  # model_AMD.py:3020


def _device_linear(m, x):
    """tiny motion-map Linear (model_AMD.py:124-127) on the SIMT GEMM"""
    from .engine import Ctx, small_linear
    ctx = Ctx(x.device, torch.float32)
    y = small_linear(ctx, x.reshape(-1, x.shape[-1]).float().contiguous(), m.weight.detach().float().contiguous(),
                     m.bias.detach().float().contiguous(), "motion_map")
    return y.view(*x.shape[:-1], -1).clone()


def AMD_N(**kwargs):
    """model_AMD.py:3151-3161."""
    return AMDModel_New(enc_nhead=8, enc_ndim=64, diffusion_attn_head_dim=64, diffusion_attn_num_heads=16,
                        diffusion_out_channels=4, diffusion_num_layers=12, **kwargs)


def AMD_S(**kwargs):
    """model_AMD.py:3127-3137."""
    return AMDModel(enc_nhead=8, enc_ndim=64, diffusion_attn_head_dim=64, diffusion_attn_num_heads=16, diffusion_out_channels=4,
                    diffusion_num_layers=12, **kwargs)


def AMD_S_Camera(**kwargs):
    """model_AMD.py:3139-3149."""
    return AMDModel_Camera(enc_nhead=8, enc_ndim=64, diffusion_attn_head_dim=64, diffusion_attn_num_heads=16, diffusion_out_channels=4,
                           diffusion_num_layers=12, **kwargs)


def AMD_L(**kwargs):
    """model_AMD.py:3163-3173.  96-wide attention heads: outside the 64-wide kernels of this package (raises NotImplementedError)."""
    return AMDModel(enc_num_layers=8, enc_nhead=16, enc_ndim=64, diffusion_attn_head_dim=96, diffusion_attn_num_heads=16,
                    diffusion_out_channels=4, diffusion_num_layers=16, **kwargs)


def AMD_S_Rec(**kwargs):
    """model_AMD.py:3176-3186."""
    return AMDModel_Rec(enc_num_layers=8, enc_nhead=8, enc_ndim=64, diffusion_attn_head_dim=64, diffusion_attn_num_heads=16,
                        diffusion_out_channels=4, diffusion_num_layers=12, **kwargs)


def AMD_S_RecSplit(**kwargs):
    """model_AMD.py:3189-3200 (`is_split` is swallowed by AMDModel_Rec's **kwargs in the reference too)."""
    return AMDModel_Rec(enc_num_layers=8, enc_nhead=8, enc_ndim=64, diffusion_attn_head_dim=64, diffusion_attn_num_heads=16,
                        diffusion_out_channels=4, diffusion_num_layers=12, is_split=True, **kwargs)


AMD_models = {"AMD_S": AMD_S, "AMD_S_Camera": AMD_S_Camera, "AMD_N": AMD_N, "AMD_L": AMD_L, "AMD_S_Rec": AMD_S_Rec,
              "AMD_S_RecSplit": AMD_S_RecSplit}
