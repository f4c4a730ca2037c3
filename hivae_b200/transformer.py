"""Motion encoders and the default rectified-flow DiT with the reference's class names, kwargs and
state-dict keys (reference: model/transformer.py:36-192, 406-586, 1746-2043), running on the
hivae_b200 sm_100a kernels.  No tensor is transposed between the spatial and temporal blocks: the
'(n t) s d <-> (n s) t d' regrouping of the reference is expressed as strided sequence maps.
"""
from typing import Optional, Union

import numpy as np
import torch
from torch import nn

from . import embeddings as E
from . import engine, ops
from .engine import AdaPack, Ctx
from .modules import (AdaLayerNorm, AMDTransformerBlock, AMDTransformerMotionBlock, BasicCrossTransformerBlock, BasicDiTBlock,
                      BasicTransformerBlock, PatchEmbed)
from .ops import Ada, Rows


class Timesteps(nn.Module):
    """diffusers Timesteps (parameter-free)."""

    def __init__(self, num_channels, flip_sin_to_cos, downscale_freq_shift):
        super().__init__()
        if not flip_sin_to_cos or downscale_freq_shift != 0:
            raise NotImplementedError("only flip_sin_to_cos=True, freq_shift=0 (the reference's setting)")
        self.num_channels = num_channels


class TimestepEmbedding(nn.Module):
    """diffusers TimestepEmbedding parameter layout: linear_1 -> SiLU -> linear_2."""

    def __init__(self, in_channels, time_embed_dim, act_fn="silu"):
        super().__init__()
        if act_fn != "silu":
            raise NotImplementedError(act_fn)
        self.linear_1 = nn.Linear(in_channels, time_embed_dim)
        self.act = nn.SiLU()
        self.linear_2 = nn.Linear(time_embed_dim, time_embed_dim)


class _Prepared:
    def weights_changed(self):
        """drop everything that was derived from the weights by kernels (token tables)"""
        if hasattr(self, "tabs"):
            self.tabs.clear()


class _HvModule(nn.Module):
    """caches the prepared (packed) weights per (device, compute dtype, parameter versions)"""

    def _prepared(self, device):
        dtype = engine.compute_dtype_for(self)
        key = (torch.device(device), dtype, engine.params_key(self))
        if getattr(self, "_hv_key", None) != key:
            self._hv_P = self._prepare(Ctx(device, dtype))
            self._hv_key = key
            self._hv_P.ctx.version = engine.params_version(self)
        elif self._hv_P.ctx.version != engine.params_version(self):  # parameters updated in place: refresh the packed copies
            self._hv_P.ctx.refresh()
            self._hv_P.weights_changed()
            self._hv_P.ctx.version = engine.params_version(self)
        return self._hv_P

    def _patch_params(self, ctx, pe):
        w = ctx.f32(pe.proj.weight)
        return w.reshape(w.shape[0], -1).contiguous(), ctx.f32(pe.proj.bias)


def mask_noise(N, L, device):
    """the noise of random_masking (transformer.py:173, 567): the reference's own call, so a seeded run draws the same
    numbers; tests replace this hook to inject the noise of a CPU run of the reference"""
    return torch.rand(N, L, device=device)


def _mask_idx(noise, mask_ratio):
    """transformer.py:170-181: ids_keep = argsort(noise)[:, :int(L (1 - mask_ratio))]"""
    keep = int(noise.shape[1] * (1 - mask_ratio))
    return torch.argsort(noise, dim=1)[:, :keep].to(torch.int32).contiguous()


def _final_norm_out(need_norm_out, m):
    """norm_out = LayerNorm(motion_channel, elementwise_affine=False) of the encoders (transformer.py:107-108, 157-158)"""
    if need_norm_out:  # motion_need_norm_out=False in every shipped AMD_N script; AMDModel_Rec's default
        c = m.shape[-1]
        if m.is_cuda and c % 128 == 0:
            rows = m.reshape(1, -1, c).float().contiguous()
            out = torch.empty(rows.shape[1], c, dtype=torch.float32, device=m.device)
            ops.ln_modulate(Rows(rows, 0, rows.shape[1]), None, None, 1e-5, out)
            return out.view(m.shape).to(m.dtype)
        m = torch.nn.functional.layer_norm(m, (c,))  # narrow token channels (16): a few hundred values
    return m


# ================================================================================================
class MotionEncoderLearnTokenTransformer(_HvModule):
    """Object-motion encoder (transformer.py:36-162): learnable tokens || patch tokens -> blocks ->
    first tokens -> LN -> Linear."""

    def __init__(self, img_height=32, img_width=32, img_inchannel=4, img_patch_size=2, motion_token_num=12,
                 motion_channel=128, need_norm_out=True, num_attention_heads=12, attention_head_dim=64, num_layers=8,
                 freq_shift=0, dropout=0.0, attention_bias=True, activation_fn="gelu-approximate",
                 norm_elementwise_affine=True, norm_eps=1e-5, spatial_interpolation_scale=1.0,
                 temporal_interpolation_scale=1.0):
        super().__init__()
        hidden_dim = num_attention_heads * attention_head_dim
        iph, ipw = img_height // img_patch_size, img_width // img_patch_size
        self.img_token_len = iph * ipw
        self.motion_token = nn.Parameter(torch.randn(1, motion_token_num, motion_channel) * 0.02)
        self.motion_embed = nn.Linear(motion_channel, hidden_dim)
        self.patch_embed = PatchEmbed(img_patch_size, img_inchannel, hidden_dim)
        self.embedding_dropout = nn.Dropout(dropout)
        self.register_buffer("pos_embedding", E.table(E.sincos_2d(hidden_dim, iph, ipw)), persistent=False)
        self.transformer_blocks = nn.ModuleList([
            BasicTransformerBlock(dim=hidden_dim, num_attention_heads=num_attention_heads,
                                  attention_head_dim=attention_head_dim, dropout=dropout, activation_fn=activation_fn,
                                  attention_bias=attention_bias, norm_elementwise_affine=norm_elementwise_affine,
                                  norm_eps=norm_eps) for _ in range(num_layers)])
        self.norm_final = nn.LayerNorm(hidden_dim, eps=norm_eps, elementwise_affine=norm_elementwise_affine)
        self.proj_out = nn.Linear(hidden_dim, motion_channel)
        self.need_norm_out = need_norm_out
        if need_norm_out:
            self.norm_out = nn.LayerNorm(motion_channel, eps=norm_eps, elementwise_affine=False)
        self.gradient_checkpointing = False

    def _prepare(self, ctx):
        P = _Prepared()
        P.ctx = ctx
        P.blocks = [engine.prepare_block(ctx, b, "basic") for b in self.transformer_blocks]
        P.pw, P.pb = self._patch_params(ctx, self.patch_embed)
        P.pos = ctx.f32(self.pos_embedding[0])
        P.tok = ctx.f32(self.motion_token[0])
        P.ew, P.eb = ctx.f32(self.motion_embed.weight), ctx.f32(self.motion_embed.bias)
        P.nfw, P.nfb = ctx.f32(self.norm_final.weight), ctx.f32(self.norm_final.bias)
        P.ow, P.ob = ctx.f32(self.proj_out.weight), ctx.f32(self.proj_out.bias)
        return P

    @torch.no_grad()
    def forward(self, img_hidden_states, mask_ratio=None):
        N, T, C, H, W = img_hidden_states.shape
        P = self._prepared(img_hidden_states.device)
        ctx = P.ctx
        D, p = P.ew.shape[0], self.patch_embed.patch_size
        Lm, S = P.tok.shape[0], self.img_token_len
        assert S == (H // p) * (W // p), "img_token_len should be equal!"
        # random_masking (transformer.py:164-192): every frame keeps the int(S (1 - ratio)) tokens with the smallest noise
        idx = _mask_idx(mask_noise(N * T, S, img_hidden_states.device), mask_ratio) if mask_ratio is not None else None
        keep = S if idx is None else idx.shape[1]
        frames, FR = N * T, Lm + keep
        x = img_hidden_states.reshape(frames, C, H, W).float().contiguous()
        res = ctx.buf("enc_res", (frames, FR, D), torch.float32)
        tok = engine.small_linear(ctx, P.tok, P.ew, P.eb, "enc_tok")          # (Lm, D), same for every frame
        ops.rows_scatter(tok, 0, 1, None, Rows(res, 0, Lm))
        if idx is None:
            ops.patch_embed(x, None, p, P.pw, P.pb, P.pos, None, res, FR, Lm, D)
        else:
            from .backward import rows_index
            full = ctx.buf("enc_full", (frames, S, D), torch.float32)
            ops.patch_embed(x, None, p, P.pw, P.pb, P.pos, None, full, S, 0, D)
            rows_index(full, res[:, Lm:], idx, frames, S, keep, D, S, FR)
        rows, geom = Rows(res, 0, FR), engine.dense_geom(frames, FR)
        for bp in P.blocks:
            engine.run_block(ctx, bp, rows, geom)
        mrows = Rows(res, 0, Lm)
        y = ops.ln_modulate(mrows, P.nfw, P.nfb, self.norm_final.eps, ctx.buf("enc_fin", (frames * Lm, D), torch.float32))
        m = engine.small_linear(ctx, y, P.ow, P.ob, "enc_out").view(N, T, Lm, -1).clone()
        return _final_norm_out(self.need_norm_out, m)


# ================================================================================================
class MotionEncoderLearnTokenTemporalTransformer(_HvModule):
    """Spatial + per-token temporal motion encoder (transformer.py:195-367; AMDModel's encoder when
    use_motiontemporal): every layer = BasicTransformerBlock over each frame's [motion tokens || patch tokens],
    then AMDTransformerMotionBlock over time for each motion token of the SECOND HALF of the flat (n t) frame
    batch (the reference's `chunk(2, dim=0)`, transformer.py:340 — the target frames when n == 1).  Both run in
    place on one fp32 residual buffer; the '(n t) l d -> (n l) t d' regrouping is a strided attention geometry."""

    def __init__(self, img_height=32, img_width=32, img_inchannel=4, img_patch_size=2, motion_token_num=12,
                 motion_channel=128, need_norm_out=True, video_frames=16, num_attention_heads=12, attention_head_dim=64,
                 num_layers=8, freq_shift=0, dropout=0.0, attention_bias=True, activation_fn="gelu-approximate",
                 norm_elementwise_affine=True, norm_eps=1e-5, spatial_interpolation_scale=1.0,
                 temporal_interpolation_scale=1.0):
        super().__init__()
        hidden_dim = num_attention_heads * attention_head_dim
        iph, ipw = img_height // img_patch_size, img_width // img_patch_size
        self.img_token_len, self.video_frames = iph * ipw, video_frames
        self.motion_token = nn.Parameter(torch.randn(1, motion_token_num, motion_channel) * 0.02)
        self.motion_embed = nn.Linear(motion_channel, hidden_dim)
        self.patch_embed = PatchEmbed(img_patch_size, img_inchannel, hidden_dim)
        self.embedding_dropout = nn.Dropout(dropout)
        self.register_buffer("pos_embedding", E.table(E.sincos_2d(hidden_dim, iph, ipw)), persistent=False)
        self.register_buffer("motion_temporal_embedding",
                             E.table(E.sincos_1d(hidden_dim, np.arange(video_frames * motion_token_num))), persistent=False)
        kw = dict(dim=hidden_dim, num_attention_heads=num_attention_heads, attention_head_dim=attention_head_dim,
                  dropout=dropout, activation_fn=activation_fn, attention_bias=attention_bias,
                  norm_elementwise_affine=norm_elementwise_affine, norm_eps=norm_eps)
        self.transformer_blocks = nn.ModuleList([BasicTransformerBlock(**kw) for _ in range(num_layers)])
        self.motion_blocks = nn.ModuleList([AMDTransformerMotionBlock(**kw) for _ in range(num_layers)])
        self.norm_final = nn.LayerNorm(hidden_dim, eps=norm_eps, elementwise_affine=norm_elementwise_affine)
        self.proj_out = nn.Linear(hidden_dim, motion_channel)
        self.need_norm_out = need_norm_out
        if need_norm_out:
            self.norm_out = nn.LayerNorm(motion_channel, eps=norm_eps, elementwise_affine=False)
        self.gradient_checkpointing = False

    def _prepare(self, ctx):
        P = MotionEncoderLearnTokenTransformer._prepare(self, ctx)
        P.mblocks = [engine.prepare_block(ctx, b, "basic") for b in self.motion_blocks]
        P.tpos = ctx.f32(self.motion_temporal_embedding[0])
        P.tabs = {}
        return P

    @staticmethod
    def _token_table(P, ctx, T):
        """initial motion tokens of one clip (T, Lm, D): embed(motion_token) for the first T/2 frames,
        + temporal sincos at position (f - T/2)*Lm + l for the others (transformer.py:305-313); cached per T"""
        if T not in P.tabs:
            Lm, D, half = P.tok.shape[0], P.ew.shape[0], T // 2
            tok = engine.small_linear(ctx, P.tok, P.ew, P.eb, "tenc_tok")
            tab = torch.empty(T, Lm, D, dtype=torch.float32, device=ctx.device)
            ops.rows_scatter(tok, 0, 1, None, Rows(tab[:half], 0, Lm))
            for f in range(half):
                ops.rows_scatter(tok, 0, 1, P.tpos[f * Lm:(f + 1) * Lm].contiguous(), Rows(tab[half + f:half + f + 1], 0, Lm))
            P.tabs[T] = tab
        return P.tabs[T]

    @torch.no_grad()
    def forward(self, img_hidden_states, mask_ratio=None):
        N, T, C, H, W = img_hidden_states.shape
        P = self._prepared(img_hidden_states.device)
        ctx = P.ctx
        D, p = P.ew.shape[0], self.patch_embed.patch_size
        Lm, S = P.tok.shape[0], self.img_token_len
        assert S == (H // p) * (W // p), "img_token_len should be equal!"
        assert T % 2 == 0 and (T // 2) * Lm <= P.tpos.shape[0]
        # random_masking (transformer.py:369-397): every frame keeps the int(S (1 - ratio)) image tokens with the smallest noise
        idx = _mask_idx(mask_noise(N * T, S, img_hidden_states.device), mask_ratio) if mask_ratio is not None else None
        keep = S if idx is None else idx.shape[1]
        frames, FR, half = N * T, Lm + keep, T // 2
        x = img_hidden_states.reshape(frames, C, H, W).float().contiguous()
        res = ctx.buf("tenc_res", (frames, FR, D), torch.float32)
        tab = self._token_table(P, ctx, T)
        for n in range(N):
            ops.rows_scatter(tab.view(T * Lm, D), Lm, 1, None, Rows(res[n * T:(n + 1) * T], 0, Lm))
        if idx is None:
            ops.patch_embed(x, None, p, P.pw, P.pb, P.pos, None, res, FR, Lm, D)
        else:
            from .backward import rows_index
            full = ctx.buf("tenc_full", (frames, S, D), torch.float32)
            ops.patch_embed(x, None, p, P.pw, P.pb, P.pos, None, full, S, 0, D)
            rows_index(full, res[:, Lm:], idx, frames, S, keep, D, S, FR)
        rows, geom = Rows(res, 0, FR), engine.dense_geom(frames, FR)
        trows, tgeom = Rows(res[frames // 2:], 0, Lm), engine.strided_geom(N, Lm, half)  # '(n t) l d -> (n l) t d'
        for bp, mp in zip(P.blocks, P.mblocks):
            engine.run_block(ctx, bp, rows, geom)
            engine.run_block(ctx, mp, trows, tgeom)
        y = ops.ln_modulate(Rows(res, 0, Lm), P.nfw, P.nfb, self.norm_final.eps, ctx.buf("tenc_fin", (frames * Lm, D), torch.float32))
        m = engine.small_linear(ctx, y, P.ow, P.ob, "tenc_out").view(N, T, Lm, -1).clone()
        return _final_norm_out(self.need_norm_out, m)


# ================================================================================================
class MotionEncoderLearnTokenOnlyTemporalTransformer(_HvModule):
    """Camera-motion encoder (transformer.py:406-556): per spatial location, T learnable temporal
    queries cross-attend to that location's T image tokens."""

    def __init__(self, img_height=32, img_width=32, img_inchannel=4, img_patch_size=2, motion_token_num=12,
                 motion_channel=128, need_norm_out=True, video_frames=16, num_attention_heads=12, attention_head_dim=64,
                 num_layers=8, freq_shift=0, dropout=0.0, attention_bias=True, activation_fn="gelu-approximate",
                 norm_elementwise_affine=True, norm_eps=1e-5, spatial_interpolation_scale=1.0,
                 temporal_interpolation_scale=1.0):
        super().__init__()
        hidden_dim = num_attention_heads * attention_head_dim
        iph, ipw = img_height // img_patch_size, img_width // img_patch_size
        self.img_token_len = iph * ipw
        self.video_frames = video_frames
        self.motion_token = nn.Parameter(torch.randn(1, motion_token_num, motion_channel) * 0.02)
        self.motion_embed = nn.Linear(motion_channel, hidden_dim)
        self.patch_embed = PatchEmbed(img_patch_size, img_inchannel, hidden_dim)
        self.embedding_dropout = nn.Dropout(dropout)
        self.register_buffer("pos_embedding", E.table(E.sincos_2d(hidden_dim, iph, ipw)), persistent=False)
        self.register_buffer("motion_temporal_embedding", E.table(E.sincos_1d(hidden_dim, np.arange(video_frames))),
                             persistent=False)
        self.transformer_blocks = nn.ModuleList([
            BasicCrossTransformerBlock(dim=hidden_dim, cross_dim=hidden_dim, num_attention_heads=num_attention_heads,
                                       attention_head_dim=attention_head_dim, dropout=dropout, activation_fn=activation_fn,
                                       attention_bias=attention_bias, norm_elementwise_affine=norm_elementwise_affine,
                                       norm_eps=norm_eps) for _ in range(num_layers)])
        self.norm_final = nn.LayerNorm(hidden_dim, eps=norm_eps, elementwise_affine=norm_elementwise_affine)
        self.proj_out = nn.Linear(hidden_dim, motion_channel)
        self.need_norm_out = need_norm_out
        if need_norm_out:
            self.norm_out = nn.LayerNorm(motion_channel, eps=norm_eps, elementwise_affine=False)
        self.gradient_checkpointing = False

    def _prepare(self, ctx):
        P = _Prepared()
        P.ctx = ctx
        P.blocks = [engine.prepare_block(ctx, b, "cross") for b in self.transformer_blocks]
        P.pw, P.pb = self._patch_params(ctx, self.patch_embed)
        P.pos = ctx.f32(self.pos_embedding[0])
        P.tpos = ctx.f32(self.motion_temporal_embedding[0])
        P.tok = ctx.f32(self.motion_token[0])
        P.ew, P.eb = ctx.f32(self.motion_embed.weight), ctx.f32(self.motion_embed.bias)
        P.nfw, P.nfb = ctx.f32(self.norm_final.weight), ctx.f32(self.norm_final.bias)
        P.ow, P.ob = ctx.f32(self.proj_out.weight), ctx.f32(self.proj_out.bias)
        P.tabs = {}
        return P

    @staticmethod
    def _token_table(P, ctx, T, S, D):
        """weights-only table of the initial query tokens of one clip, cached per T"""
        if (T, S) not in P.tabs:
            t0 = P.tok.shape[0]
            rep = T // t0
            tok = engine.small_linear(ctx, P.tok, P.ew, P.eb, "cam_tok")  # (t0, D)
            tt = torch.empty(1, T, D, dtype=torch.float32, device=ctx.device)  # tt[t] = tok[t // rep] + temporal[t]
            for f in range(t0):
                ops.rows_scatter(tok[f:f + 1], 0, 0, P.tpos[f * rep:(f + 1) * rep].contiguous(), Rows(tt, f * rep, rep))
            tab = torch.empty(1, T * S, D, dtype=torch.float32, device=ctx.device)
            ops.rows_scatter(tt.view(T, D), 1, 0, None, Rows(tab.view(T, S, D), 0, S))  # broadcast over s
            P.tabs[(T, S)] = tab
        return P.tabs[(T, S)]

    @torch.no_grad()
    def forward(self, img_hidden_states, mask_ratio=None):
        N, T, C, H, W = img_hidden_states.shape
        P = self._prepared(img_hidden_states.device)
        ctx = P.ctx
        D, p = P.ew.shape[0], self.patch_embed.patch_size
        S0, t0 = self.img_token_len, P.tok.shape[0]
        assert S0 == (H // p) * (W // p), "img_token_len should be equal!"
        assert T <= self.video_frames and T % t0 == 0
        x = img_hidden_states.reshape(N * T, C, H, W).float().contiguous()
        # random_masking (transformer.py:558-586): a clip keeps the same int(S (1 - ratio)) spatial locations in every frame
        idx = _mask_idx(mask_noise(N, S0, img_hidden_states.device), mask_ratio) if mask_ratio is not None else None
        S = S0 if idx is None else idx.shape[1]
        # image K/V tokens, dense rows ordered (n, t, s), + 2-D sincos + temporal sincos  (506-514)
        img = ctx.buf("cam_img", (N * T * S, D))
        if idx is None:
            ops.patch_embed(x, None, p, P.pw, P.pb, P.pos, P.tpos[:T].contiguous(), img.view(N * T, S, D), S, 0, D)
        else:
            from .backward import rows_index
            full = ctx.buf("cam_full", (N * T, S0, D), torch.float32)
            ops.patch_embed(x, None, p, P.pw, P.pb, P.pos, P.tpos[:T].contiguous(), full, S0, 0, D)
            kept = rows_index(full, ctx.buf("cam_kept", (N * T, S, D), torch.float32), idx.repeat_interleave(T, 0).contiguous(),
                              N * T, S0, S, D, S0, S)
            ops.convert(kept, img.view(N * T, S, D))
        # motion queries: token (t) = embed(motion_token)[t // (T/t0)] + temporal[t], identical for all (n, s)  (521-531)
        res = ctx.buf("cam_res", (N * T, S, D), torch.float32)                            # rows ordered (n, t, s)
        block = Rows(res.view(N, T * S, D), 0, T * S)
        tab = self._token_table(P, ctx, T, S, D)  # (T*S, D): row t*S + s = tok[t // (T/t0)] + temporal[t]
        ops.rows_scatter(tab.view(T * S, D), 0, 1, None, block)
        rows = Rows(res, 0, S)
        geom = engine.strided_geom(N, S, T)
        for bp in P.blocks:
            engine.run_block(ctx, bp, rows, geom, kv_dense=img)
        y = ops.ln_modulate(rows, P.nfw, P.nfb, self.norm_final.eps, ctx.buf("cam_fin", (N * T * S, D), torch.float32))
        m = engine.small_linear(ctx, y, P.ow, P.ob, "cam_out").view(N, T, S, -1).clone()  # already 'n t s c'
        return _final_norm_out(self.need_norm_out, m)


# ================================================================================================
class AMDDiffusionTransformerModelImgSpatialTempMotion(_HvModule):
    """Default rectified-flow denoiser (transformer.py:1746-2043): per layer an object joint block
    (image + object tokens), a camera joint block (image + camera tokens) and a temporal block."""

    _supports_gradient_checkpointing = True
    _has_temporal_blocks = True

    def __init__(self, num_attention_heads=20, attention_head_dim=64, out_channels: Optional[int] = 4, num_layers=12,
                 use_camera=False, use_object=False, image_width=32, image_height=32, image_patch_size=2,
                 image_in_channels: Optional[int] = 4, motion_token_num=12, camera_motion_in_channels: Optional[int] = 16,
                 object_motion_in_channels: Optional[int] = 64, motion_target_num_frame=16, flip_sin_to_cos=True,
                 freq_shift=0, time_embed_dim=512, dropout=0.0, attention_bias=True, temporal_compression_ratio=4,
                 activation_fn="gelu-approximate", timestep_activation_fn="silu", norm_elementwise_affine=True,
                 norm_eps=1e-5, spatial_interpolation_scale=1.0, temporal_interpolation_scale=1.0):
        super().__init__()
        hidden_dim = num_attention_heads * attention_head_dim
        iph, ipw = image_height // image_patch_size, image_width // image_patch_size
        self.image_patch_size, self.out_channels = image_patch_size, out_channels
        self.target_frame = motion_target_num_frame
        self.use_camera, self.use_object = use_camera, use_object
        self.image_patch_embed = PatchEmbed(patch_size=image_patch_size, in_channels=image_in_channels,
                                            embed_dim=hidden_dim, bias=True)
        self.embedding_dropout = nn.Dropout(dropout)
        self.register_buffer("pos_embedding", E.table(E.sincos_2d(hidden_dim, iph, ipw)), persistent=False)
        self.register_buffer("motion_pos_embedding", E.table(E.sincos_1d(hidden_dim, np.arange(2 + 2 * motion_token_num))),
                             persistent=False)
        self.register_buffer("img_temporal_embedding", E.table(E.sincos_1d(hidden_dim, np.arange(self.target_frame))),
                             persistent=False)
        self.time_proj = Timesteps(hidden_dim, flip_sin_to_cos, freq_shift)
        self.time_embedding = TimestepEmbedding(hidden_dim, time_embed_dim, timestep_activation_fn)
        blk = dict(dim=hidden_dim, num_attention_heads=num_attention_heads, attention_head_dim=attention_head_dim,
                   time_embed_dim=time_embed_dim, dropout=dropout, activation_fn=activation_fn,
                   attention_bias=attention_bias, norm_elementwise_affine=norm_elementwise_affine, norm_eps=norm_eps)
        if use_camera:
            self.camera_motion_patch_embed = nn.Linear(camera_motion_in_channels, hidden_dim)
            self.camera_transformer_blocks = nn.ModuleList([AMDTransformerBlock(**blk) for _ in range(num_layers)])
        if use_object:
            self.source_token = nn.Parameter(torch.zeros(1, 1, hidden_dim), requires_grad=True)
            self.target_token = nn.Parameter(torch.zeros(1, 1, hidden_dim), requires_grad=True)
            self.object_motion_patch_embed = nn.Linear(object_motion_in_channels, hidden_dim)
            self.object_transformer_blocks = nn.ModuleList([AMDTransformerBlock(**blk) for _ in range(num_layers)])
        if self._has_temporal_blocks:
            self.spatial_blocks = nn.ModuleList([BasicDiTBlock(**blk) for _ in range(num_layers)])
        self.norm_final = nn.LayerNorm(hidden_dim, norm_eps, norm_elementwise_affine)
        self.norm_out = AdaLayerNorm(embedding_dim=time_embed_dim, output_dim=2 * hidden_dim,
                                     norm_elementwise_affine=norm_elementwise_affine, norm_eps=norm_eps, chunk_dim=1)
        self.proj_out = nn.Linear(hidden_dim, image_patch_size * image_patch_size * out_channels)
        self.gradient_checkpointing = False

    def _prepare(self, ctx):
        P = _Prepared()
        P.ctx = ctx
        n = len(self.spatial_blocks) if self._has_temporal_blocks else len(self.object_transformer_blocks)
        P.obj = [engine.prepare_block(ctx, b, "joint") for b in self.object_transformer_blocks] if self.use_object else [None] * n
        P.cam = [engine.prepare_block(ctx, b, "joint") for b in self.camera_transformer_blocks] if self.use_camera else [None] * n
        P.sp = [engine.prepare_block(ctx, b, "dit") for b in self.spatial_blocks] if self._has_temporal_blocks else [None] * n
        order = [bp for trio in zip(P.obj, P.cam, P.sp) for bp in trio if bp is not None]
        P.ada = AdaPack(ctx, order, extra=[(self.norm_out.linear.weight, self.norm_out.linear.bias)])
        P.pw, P.pb = self._patch_params(ctx, self.image_patch_embed)
        P.pos, P.tpos = ctx.f32(self.pos_embedding[0]), ctx.f32(self.img_temporal_embedding[0])
        P.mpos = ctx.f32(self.motion_pos_embedding[0])
        te = self.time_embedding
        P.tw1, P.tb1 = ctx.weight(te.linear_1.weight), ctx.f32(te.linear_1.bias)
        P.tw2, P.tb2 = ctx.weight(te.linear_2.weight), ctx.f32(te.linear_2.bias)
        if self.use_camera:
            P.cw, P.cb = ctx.f32(self.camera_motion_patch_embed.weight), ctx.f32(self.camera_motion_patch_embed.bias)
        if self.use_object:
            P.ow, P.ob = ctx.f32(self.object_motion_patch_embed.weight), ctx.f32(self.object_motion_patch_embed.bias)
            P.stok, P.ttok = ctx.f32(self.source_token[0]), ctx.f32(self.target_token[0])
        P.nfw, P.nfb = ctx.f32(self.norm_final.weight), ctx.f32(self.norm_final.bias)
        P.now, P.nob = ctx.f32(self.norm_out.norm.weight), ctx.f32(self.norm_out.norm.bias)
        P.hw, P.hb = ctx.f32(self.proj_out.weight), ctx.f32(self.proj_out.bias)
        return P

    # -- step 1 (once per decode): embed the motion tokens ---------------------------------------
    def set_motion(self, N, device, camera=None, obj_src=None, obj_tgt=None):
        """Builds the per-frame initial motion token rows (transformer.py:1947-1980); they do not
        depend on the timestep or on z_t, so a sampling loop computes them once."""
        P = self._prepared(device)
        ctx = P.ctx
        D = P.hw.shape[1]
        M = _Prepared()
        M.Lo = M.Lc = 0
        if obj_src is not None:
            l = obj_tgt.shape[1]
            M.Lo = 2 * l + 2
            M.obj = torch.empty((N, M.Lo, D), dtype=torch.float32, device=device)   # owned by M (every row is written below)
            src = engine.small_linear(ctx, obj_src.reshape(N * l, -1).float().contiguous(), P.ow, P.ob, "dit_osrc")
            ops.rows_scatter(src, l, 1, P.mpos[1:1 + l], Rows(M.obj, 1, l))
            tgt = engine.small_linear(ctx, obj_tgt.reshape(N * l, -1).float().contiguous(), P.ow, P.ob, "dit_otgt")
            ops.rows_scatter(tgt, l, 1, P.mpos[l + 2:2 * l + 2], Rows(M.obj, l + 2, l))
            ops.rows_scatter(P.stok, 0, 0, P.mpos[0:1], Rows(M.obj, 0, 1))
            ops.rows_scatter(P.ttok, 0, 0, P.mpos[l + 1:l + 2], Rows(M.obj, l + 1, 1))
        if camera is not None:
            M.Lc = camera.shape[2]
            cam = camera.reshape(N * M.Lc, -1).float().contiguous()  # 'n t s d -> (n t) s d'
            M.cam = engine.small_linear(ctx, cam, P.cw, P.cb, "dit_minit_cam").clone()   # out of the shared workspace: two M can coexist
        return M

    # -- step 2 (every Euler step) -----------------------------------------------------------------
    def run(self, M, in1, in2, timestep, pre=None, zt=None, dt=0.0):
        """One denoiser evaluation.  in1 (+ in2) are the NCHW channel groups of the image input
        (zi, zt — the torch.cat of model_AMD.py:2178 is folded into the patch-embed kernel);
        writes the prediction to `pre` and/or applies zt += dt * pre in place."""
        N, _, Hi, Wi = in1.shape
        P = self._prepared(in1.device)
        ctx = P.ctx
        T, p = self.target_frame, self.image_patch_size
        n = N // T
        D = P.hw.shape[1]
        S = (Hi // p) * (Wi // p)
        Lo, Lc = M.Lo, M.Lc
        FR = Lo + S + Lc
        res = ctx.buf("dit_res", (N, FR, D), torch.float32)
        ops.patch_embed(in1, in2, p, P.pw, P.pb, P.pos, P.tpos, res, FR, Lo, D)
        if Lo:
            ops.rows_scatter(M.obj.view(N * Lo, D), Lo, 1, None, Rows(res, 0, Lo))
        if Lc:
            ops.rows_scatter(M.cam, Lc, 1, None, Rows(res, Lo + S, Lc))
        # timestep embedding -> temb -> every adaLN table of the step in ONE GEMM
        ts = timestep.to(device=in1.device, dtype=torch.float32).contiguous()
        te = ops.timestep_embed(ts, D, ctx.buf("dit_te", (N, D)))
        h = engine.small_linear(ctx, te, P.tw1, P.tb1, "dit_te_h")
        hs = ops.silu(h, ctx.buf("dit_te_hs", h.shape))
        emb = engine.small_linear(ctx, hs, P.tw2, P.tb2, "dit_emb")
        table = P.ada.table(ctx, emb, "dit_ada")
        ada_f, ada_s = Ada(table, 1, 1), Ada(table, T, T)  # emb per frame / emb of frame 0 of the clip (1943-1945)
        obj_rows = Rows(res, 0, Lo + S, img_start=Lo, img_len=S)
        cam_rows = Rows(res, Lo, S + Lc, img_start=Lo, img_len=S)
        img_rows = Rows(res, Lo, S, img_start=Lo, img_len=S)
        g_obj, g_cam = engine.dense_geom(N, Lo + S), engine.dense_geom(N, S + Lc)
        g_tmp = engine.strided_geom(n, S, T)
        for bo, bc, bs in zip(P.obj, P.cam, P.sp):
            if Lo:
                engine.run_block(ctx, bo, obj_rows, g_obj, ada_f)
            if Lc:
                engine.run_block(ctx, bc, cam_rows, g_cam, ada_f)
            if bs is not None:
                engine.run_block(ctx, bs, img_rows, g_tmp, ada_s)
        off = P.ada.extra_off[0]
        ops.final_head(img_rows, P.nfw, P.nfb, P.now, P.nob, self.norm_final.eps, ada_f, off, off + D,
                       P.hw, P.hb, self.out_channels, p, Wi // p, pre, zt, dt)
        return pre

    @torch.no_grad()
    def forward(self, image_hidden_states, timestep: Union[int, float, torch.Tensor], timestep_cond=None,
                camera_motion_target_hidden_states=None, object_motion_source_hidden_states=None,
                object_motion_target_hidden_states=None, return_dict=True):
        N, Ci, Hi, Wi = image_hidden_states.shape
        x = image_hidden_states.float().contiguous()
        M = self.set_motion(N, x.device, camera_motion_target_hidden_states, object_motion_source_hidden_states,
                            object_motion_target_hidden_states)
        if not torch.is_tensor(timestep):
            timestep = torch.full((N,), float(timestep), device=x.device)
        pre = torch.empty(N, self.out_channels, Hi, Wi, dtype=torch.float32, device=x.device)
        self.run(M, x, None, timestep, pre=pre)
        return pre.to(image_hidden_states.dtype)


class AMDDiffusionTransformerModelTempMotion(AMDDiffusionTransformerModelImgSpatialTempMotion):
    """The diffusion_model_type='default' denoiser of AMDModel_New (transformer.py:1483-1743): object joint blocks
    only (image + object-motion tokens, per-frame timestep embedding), no camera stream and no temporal blocks.
    Same parameter names as the reference, so its checkpoints load unchanged."""

    _has_temporal_blocks = False

    def __init__(self, num_attention_heads=20, attention_head_dim=64, out_channels: Optional[int] = 4, num_layers=12,
                 use_camera=False, use_object=False, **kwargs):
        # the reference class builds the object stream only (use_camera is accepted and ignored, 1612-1630)
        super().__init__(num_attention_heads=num_attention_heads, attention_head_dim=attention_head_dim,
                         out_channels=out_channels, num_layers=num_layers, use_camera=False, use_object=use_object, **kwargs)

    @torch.no_grad()
    def forward(self, image_hidden_states, timestep, timestep_cond=None, object_motion_source_hidden_states=None,
                object_motion_target_hidden_states=None, return_dict=True):
        return super().forward(image_hidden_states, timestep, timestep_cond, None, object_motion_source_hidden_states,
                               object_motion_target_hidden_states, return_dict)


# ================================================================================================
class AMDDiffusionTransformerModelImgSpatial(_HvModule):
    """The single-motion-stream sibling of the default denoiser (transformer.py:2044-2356; AMDModel's
    diffusion_model_type='spatial'): ONE `motion_patch_embed` and ONE stack of joint `transformer_blocks`, each
    followed by a temporal `spatial_blocks[i]`.

    motion_type='plus'     : the motion stream is [source_token | object source | target_token | object target].
    motion_type='decouple' : the stream starts as the camera tokens ([source_token | camera source | target_token |
                             camera target], or [target_token | camera target] without a camera source); when object
                             motions are given, layers 0-5 run on the camera stream, then the camera rows are
                             REPLACED by the object tokens (the evolved source/target tokens stay) and layers >= 6
                             run on that stream (2259-2322).  The reference forms the object TARGET rows from the
                             object SOURCE embedding (2262-2263); that is reproduced.
    All streams live in one fp32 frames buffer (rows [0, 2L+2) motion, then the image rows), so the swap is two
    small row scatters and no block ever concatenates anything."""

    _has_temporal_blocks = True   # spatial_blocks[i] (attention over time) after every joint block
    _camera_layers_end = 6        # 'decouple' with object motions: camera stream for layers [0, 6) ...
    _object_layers_start = 6      # ... object stream for layers [6, n)
    _plus_adds_camera = False     # 'plus' embeds the object tokens only (2231-2239)

    def __init__(self, num_attention_heads=20, attention_head_dim=64, out_channels: Optional[int] = 4, num_layers=12,
                 motion_type="plus", image_width=32, image_height=32, image_patch_size=2,
                 image_in_channels: Optional[int] = 4, motion_token_num=12, motion_in_channels: Optional[int] = 128,
                 motion_target_num_frame=16, flip_sin_to_cos=True, freq_shift=0, time_embed_dim=512, dropout=0.0,
                 attention_bias=True, temporal_compression_ratio=4, activation_fn="gelu-approximate",
                 timestep_activation_fn="silu", norm_elementwise_affine=True, norm_eps=1e-5,
                 spatial_interpolation_scale=1.0, temporal_interpolation_scale=1.0):
        super().__init__()
        hidden_dim = num_attention_heads * attention_head_dim
        iph, ipw = image_height // image_patch_size, image_width // image_patch_size
        self.image_patch_size, self.out_channels = image_patch_size, out_channels
        self.target_frame, self.motion_type = motion_target_num_frame, motion_type
        self.image_patch_embed = PatchEmbed(patch_size=image_patch_size, in_channels=image_in_channels,
                                            embed_dim=hidden_dim, bias=True)
        self.motion_patch_embed = nn.Linear(motion_in_channels, hidden_dim)
        self.embedding_dropout = nn.Dropout(dropout)
        self.register_buffer("pos_embedding", E.table(E.sincos_2d(hidden_dim, iph, ipw)), persistent=False)
        self.register_buffer("motion_pos_embedding", E.table(E.sincos_1d(hidden_dim, np.arange(2 + 2 * motion_token_num))),
                             persistent=False)
        if self._has_temporal_blocks:
            self.register_buffer("img_temporal_embedding", E.table(E.sincos_1d(hidden_dim, np.arange(self.target_frame))),
                                 persistent=False)
        self.source_token = nn.Parameter(torch.zeros(1, 1, hidden_dim), requires_grad=True)
        self.target_token = nn.Parameter(torch.zeros(1, 1, hidden_dim), requires_grad=True)
        self.time_proj = Timesteps(hidden_dim, flip_sin_to_cos, freq_shift)
        self.time_embedding = TimestepEmbedding(hidden_dim, time_embed_dim, timestep_activation_fn)
        blk = dict(dim=hidden_dim, num_attention_heads=num_attention_heads, attention_head_dim=attention_head_dim,
                   time_embed_dim=time_embed_dim, dropout=dropout, activation_fn=activation_fn,
                   attention_bias=attention_bias, norm_elementwise_affine=norm_elementwise_affine, norm_eps=norm_eps)
        self.transformer_blocks = nn.ModuleList([AMDTransformerBlock(**blk) for _ in range(num_layers)])
        if self._has_temporal_blocks:
            self.spatial_blocks = nn.ModuleList([BasicDiTBlock(**blk) for _ in range(num_layers)])
        self.norm_final = nn.LayerNorm(hidden_dim, norm_eps, norm_elementwise_affine)
        self.norm_out = AdaLayerNorm(embedding_dim=time_embed_dim, output_dim=2 * hidden_dim,
                                     norm_elementwise_affine=norm_elementwise_affine, norm_eps=norm_eps, chunk_dim=1)
        self.proj_out = nn.Linear(hidden_dim, image_patch_size * image_patch_size * out_channels)
        self.gradient_checkpointing = False

    def _prepare(self, ctx):
        P = _Prepared()
        P.ctx = ctx
        P.joint = [engine.prepare_block(ctx, b, "joint") for b in self.transformer_blocks]
        P.sp = [engine.prepare_block(ctx, b, "dit") for b in self.spatial_blocks] if self._has_temporal_blocks else [None] * len(P.joint)
        order = [bp for pair in zip(P.joint, P.sp) for bp in pair if bp is not None]
        P.ada = AdaPack(ctx, order, extra=[(self.norm_out.linear.weight, self.norm_out.linear.bias)])
        P.pw, P.pb = self._patch_params(ctx, self.image_patch_embed)
        P.pos = ctx.f32(self.pos_embedding[0])
        P.tpos = ctx.f32(self.img_temporal_embedding[0]) if self._has_temporal_blocks else None
        P.mpos = ctx.f32(self.motion_pos_embedding[0])
        te = self.time_embedding
        P.tw1, P.tb1 = ctx.weight(te.linear_1.weight), ctx.f32(te.linear_1.bias)
        P.tw2, P.tb2 = ctx.weight(te.linear_2.weight), ctx.f32(te.linear_2.bias)
        P.mw, P.mb = ctx.f32(self.motion_patch_embed.weight), ctx.f32(self.motion_patch_embed.bias)
        P.stok, P.ttok = ctx.f32(self.source_token[0]), ctx.f32(self.target_token[0])
        P.nfw, P.nfb = ctx.f32(self.norm_final.weight), ctx.f32(self.norm_final.bias)
        P.now, P.nob = ctx.f32(self.norm_out.norm.weight), ctx.f32(self.norm_out.norm.bias)
        P.hw, P.hb = ctx.f32(self.proj_out.weight), ctx.f32(self.proj_out.bias)
        return P

    @torch.no_grad()
    def forward(self, camera_motion_target_hidden_states, image_hidden_states, timestep, timestep_cond=None,
                camera_motion_source_hidden_states=None, object_motion_source_hidden_states=None,
                object_motion_target_hidden_states=None, return_dict=True):
        image2 = None
        if isinstance(image_hidden_states, (tuple, list)):  # (zi, zt): the channel concat of model_AMD.py:380 is folded into the patch embed
            image_hidden_states, image2 = image_hidden_states
        N, Ci, Hi, Wi = image_hidden_states.shape
        _, L, Cm = camera_motion_target_hidden_states.shape
        P = self._prepared(image_hidden_states.device)
        ctx = P.ctx
        T, p = (self.target_frame if self._has_temporal_blocks else 1), self.image_patch_size
        n, D, S = N // T, P.hw.shape[1], (Hi // p) * (Wi // p)
        LM = 2 * L + 2
        FR = LM + S
        res = ctx.buf("dits_res", (N, FR, D), torch.float32)
        emb_m = lambda x, name: engine.small_linear(ctx, x.reshape(N * L, Cm).float().contiguous(), P.mw, P.mb, name)
        obj = object_motion_source_hidden_states is not None
        cam_src = camera_motion_source_hidden_states is not None

        def put_tokens(src_rows, tgt_rows, pos_src, pos_tgt):
            """rows 1..L <- src_rows + pos_src, rows L+2.. <- tgt_rows + pos_tgt  (either may be None)"""
            if src_rows is not None:
                ops.rows_scatter(src_rows, L, 1, pos_src, Rows(res, 1, L))
            if tgt_rows is not None:
                ops.rows_scatter(tgt_rows, L, 1, pos_tgt, Rows(res, L + 2, L))

        if self.motion_type == "plus":
            ms, mt = object_motion_source_hidden_states, object_motion_target_hidden_states
            if self._plus_adds_camera:  # embed(camera + object) (transformer.py:1171-1175)
                ms, mt = camera_motion_source_hidden_states + ms, camera_motion_target_hidden_states + mt
            put_tokens(emb_m(ms, "dits_s"), emb_m(mt, "dits_t"), P.mpos[1:1 + L], P.mpos[L + 2:LM])
            ops.rows_scatter(P.stok, 0, 0, P.mpos[0:1], Rows(res, 0, 1))
            ops.rows_scatter(P.ttok, 0, 0, P.mpos[L + 1:L + 2], Rows(res, L + 1, 1))
            phases = [(0, 0, len(P.joint))]                              # (first motion row, first layer, last layer)
        elif self.motion_type == "decouple":
            ct = emb_m(camera_motion_target_hidden_states, "dits_ct")
            if cam_src:
                put_tokens(emb_m(camera_motion_source_hidden_states, "dits_cs"), ct, P.mpos[1:1 + L], P.mpos[L + 2:LM])
                ops.rows_scatter(P.stok, 0, 0, P.mpos[0:1], Rows(res, 0, 1))
                ops.rows_scatter(P.ttok, 0, 0, P.mpos[L + 1:L + 2], Rows(res, L + 1, 1))
                off = 0
            else:  # [target_token | camera target] + pos[:L+1], kept right-aligned next to the image rows
                put_tokens(None, ct, None, P.mpos[1:1 + L])
                ops.rows_scatter(P.ttok, 0, 0, P.mpos[0:1], Rows(res, L + 1, 1))
                off = L + 1
            phases = ([(off, 0, self._camera_layers_end), (0, self._object_layers_start, len(P.joint))] if obj
                      else [(off, 0, len(P.joint))])
        else:
            raise UnboundLocalError("motion_hidden_states")  # the reference leaves it undefined for other motion types
        x = image_hidden_states.float().contiguous()
        ops.patch_embed(x, image2.float().contiguous() if image2 is not None else None, p, P.pw, P.pb, P.pos, P.tpos, res, FR, LM, D)
        if not torch.is_tensor(timestep):
            timestep = torch.full((N,), float(timestep), device=x.device)
        ts = timestep.to(device=x.device, dtype=torch.float32).contiguous()
        te = ops.timestep_embed(ts, D, ctx.buf("dits_te", (N, D)))
        h = engine.small_linear(ctx, te, P.tw1, P.tb1, "dits_te_h")
        emb = engine.small_linear(ctx, ops.silu(h, ctx.buf("dits_te_hs", h.shape)), P.tw2, P.tb2, "dits_emb")
        table = P.ada.table(ctx, emb, "dits_ada")
        ada_f, ada_s = Ada(table, 1, 1), Ada(table, T, T)
        img_rows, g_tmp = Rows(res, LM, S, img_start=LM, img_len=S), engine.strided_geom(n, S, T)
        for k, (off, l0, l1) in enumerate(phases):
            if k == 1:  # camera rows -> object tokens; row 0 only when there was no camera source (fresh source_token, no pos)
                so = emb_m(object_motion_source_hidden_states, "dits_os")
                src_pos = ctx.buf("dits_osp", (N * L, D), torch.float32)
                tmp = Rows(src_pos.view(N, L, D), 0, L)
                ops.rows_scatter(so, L, 1, P.mpos[1:1 + L], tmp)                 # object source + pos[1:L+1]
                put_tokens(src_pos, src_pos, None, P.mpos[L + 2:LM])            # target rows = SOURCE + pos[L+2:] (2262-2263)
                if not cam_src:
                    ops.rows_scatter(P.stok, 0, 0, None, Rows(res, 0, 1))
            rows = Rows(res, off, LM - off + S, img_start=LM, img_len=S)
            geom = engine.dense_geom(N, LM - off + S)
            for i in range(l0, min(l1, len(P.joint))):
                engine.run_block(ctx, P.joint[i], rows, geom, ada_f)
                if P.sp[i] is not None:
                    engine.run_block(ctx, P.sp[i], img_rows, g_tmp, ada_s)
        pre = torch.empty(N, self.out_channels, Hi, Wi, dtype=torch.float32, device=x.device)
        o = P.ada.extra_off[0]
        ops.final_head(img_rows, P.nfw, P.nfb, P.now, P.nob, self.norm_final.eps, ada_f, o, o + D,
                       P.hw, P.hb, self.out_channels, p, Wi // p, pre, None, 0.0)
        return pre.to(image_hidden_states.dtype)


class AMDDiffusionTransformerModel(AMDDiffusionTransformerModelImgSpatial):
    """The oldest single-stream denoiser (transformer.py:1010-1259; AMDModel's diffusion_model_type='default'): joint
    blocks only (no temporal blocks, no temporal position table).  'plus' embeds camera + object tokens;
    'decouple' with object motions runs layers [0, 8) on the camera stream and layers [6, n) on the object
    stream — layers 6 and 7 run in BOTH passes (1211, 1231), which is reproduced."""

    _has_temporal_blocks = False
    _camera_layers_end = 8
    _object_layers_start = 6
    _plus_adds_camera = True

    def __init__(self, num_attention_heads=20, attention_head_dim=64, out_channels: Optional[int] = 4, num_layers=12,
                 motion_type="plus", image_width=32, image_height=32, image_patch_size=2,
                 image_in_channels: Optional[int] = 4, motion_token_num=12, motion_in_channels: Optional[int] = 128, **kw):
        kw.pop("motion_target_num_frame", None)
        super().__init__(num_attention_heads=num_attention_heads, attention_head_dim=attention_head_dim,
                         out_channels=out_channels, num_layers=num_layers, motion_type=motion_type, image_width=image_width,
                         image_height=image_height, image_patch_size=image_patch_size, image_in_channels=image_in_channels,
                         motion_token_num=motion_token_num, motion_in_channels=motion_in_channels,
                         motion_target_num_frame=1, **kw)


# ================================================================================================
class AMDReconstructTransformerModel(_HvModule):
    """The reconstruction (non-diffusion) transformer of AMDModel_Rec (transformer.py:703-854): plain pre-LN blocks over
    [image tokens | source_token, source motion, target_token, target motion], norm_final + proj_out on the image rows,
    (c, p, q) unpatchify.  No timestep, no adaLN."""

    def __init__(self, num_attention_heads=20, attention_head_dim=64, out_channels: Optional[int] = 4, num_layers=12, image_width=32,
                 image_height=32, image_patch_size=2, image_in_channels: Optional[int] = 4, motion_token_num=12,
                 motion_in_channels: Optional[int] = 128, flip_sin_to_cos=True, freq_shift=0, time_embed_dim=512, dropout=0.0,
                 attention_bias=True, temporal_compression_ratio=4, activation_fn="gelu-approximate", timestep_activation_fn="silu",
                 norm_elementwise_affine=True, norm_eps=1e-5, spatial_interpolation_scale=1.0, temporal_interpolation_scale=1.0):
        super().__init__()
        hidden_dim = num_attention_heads * attention_head_dim
        iph, ipw = image_height // image_patch_size, image_width // image_patch_size
        self.image_patch_size, self.out_channels = image_patch_size, out_channels
        self.image_patch_embed = PatchEmbed(patch_size=image_patch_size, in_channels=image_in_channels, embed_dim=hidden_dim, bias=True)
        self.motion_patch_embed = nn.Linear(motion_in_channels, hidden_dim)
        self.embedding_dropout = nn.Dropout(dropout)
        self.register_buffer("pos_embedding", E.table(E.sincos_2d(hidden_dim, iph, ipw)), persistent=False)
        self.register_buffer("motion_pos_embedding", E.table(E.sincos_1d(hidden_dim, np.arange(2 + 2 * motion_token_num))), persistent=False)
        self.source_token = nn.Parameter(torch.zeros(1, 1, hidden_dim), requires_grad=True)
        self.target_token = nn.Parameter(torch.zeros(1, 1, hidden_dim), requires_grad=True)
        self.transformer_blocks = nn.ModuleList([
            BasicTransformerBlock(dim=hidden_dim, num_attention_heads=num_attention_heads, attention_head_dim=attention_head_dim,
                                  dropout=dropout, activation_fn=activation_fn, attention_bias=attention_bias,
                                  norm_elementwise_affine=norm_elementwise_affine, norm_eps=norm_eps) for _ in range(num_layers)])
        self.norm_final = nn.LayerNorm(hidden_dim, norm_eps, norm_elementwise_affine)
        self.proj_out = nn.Linear(hidden_dim, image_patch_size * image_patch_size * out_channels)
        self.gradient_checkpointing = False

    def _prepare(self, ctx):
        P = _Prepared()
        P.ctx = ctx
        P.blocks = [engine.prepare_block(ctx, b, "basic") for b in self.transformer_blocks]
        P.pw, P.pb = self._patch_params(ctx, self.image_patch_embed)
        P.pos, P.mpos = ctx.f32(self.pos_embedding[0]), ctx.f32(self.motion_pos_embedding[0])
        P.mw, P.mb = ctx.f32(self.motion_patch_embed.weight), ctx.f32(self.motion_patch_embed.bias)
        P.stok, P.ttok = ctx.f32(self.source_token[0]), ctx.f32(self.target_token[0])
        P.nfw, P.nfb = ctx.f32(self.norm_final.weight), ctx.f32(self.norm_final.bias)
        P.hw, P.hb = ctx.f32(self.proj_out.weight), ctx.f32(self.proj_out.bias)
        return P

    @torch.no_grad()
    def forward(self, motion_source_hidden_states, motion_target_hidden_states, image_hidden_states):
        image2 = None
        if isinstance(image_hidden_states, (tuple, list)):  # (zi, zt): the channel concat is folded into the patch embed
            image_hidden_states, image2 = image_hidden_states
        N, Ci, Hi, Wi = image_hidden_states.shape
        _, L, Cm = motion_source_hidden_states.shape
        P = self._prepared(image_hidden_states.device)
        ctx = P.ctx
        p = self.image_patch_size
        D, S, LM = P.hw.shape[1], (Hi // p) * (Wi // p), 2 * L + 2
        assert LM <= P.mpos.shape[0]
        FR = S + LM
        res = ctx.buf("rec_res", (N, FR, D), torch.float32)  # rows [0, S) image, then the motion stream (transformer.py:838)
        x = image_hidden_states.float().contiguous()
        ops.patch_embed(x, image2.float().contiguous() if image2 is not None else None, p, P.pw, P.pb, P.pos, None, res, FR, 0, D)
        emb = lambda t, name: engine.small_linear(ctx, t.reshape(N * L, Cm).float().contiguous(), P.mw, P.mb, name)
        ops.rows_scatter(P.stok, 0, 0, P.mpos[0:1], Rows(res, S, 1))
        ops.rows_scatter(emb(motion_source_hidden_states, "rec_s"), L, 1, P.mpos[1:1 + L], Rows(res, S + 1, L))
        ops.rows_scatter(P.ttok, 0, 0, P.mpos[L + 1:L + 2], Rows(res, S + L + 1, 1))
        ops.rows_scatter(emb(motion_target_hidden_states, "rec_t"), L, 1, P.mpos[L + 2:LM], Rows(res, S + L + 2, L))
        rows, geom = Rows(res, 0, FR), engine.dense_geom(N, FR)
        for bp in P.blocks:
            engine.run_block(ctx, bp, rows, geom)
        y = ops.ln_modulate(Rows(res, 0, S), P.nfw, P.nfb, self.norm_final.eps, ctx.buf("rec_fin", (N * S, D), torch.float32))
        tok = engine.small_linear(ctx, y, P.hw, P.hb, "rec_tok")
        return ops.unpatchify(tok.view(N, S, -1), p, self.out_channels, order=1).to(image_hidden_states.dtype)
