"""Transformer blocks with the reference's class names, constructor arguments and state-dict keys
(reference: model/modules.py:651-1188 and the diffusers Attention / FeedForward they wrap), whose
forward runs on the hand-written sm_100a kernels through hivae_b200.engine.

nn.Linear / nn.LayerNorm instances below are PARAMETER CONTAINERS only (they give the same
parameter names and default initialisation as the reference); their own forward is never called.
"""
import torch
from torch import nn

from . import engine, ops
from .engine import Ctx
from .ops import Ada, Rows


class Attention(nn.Module):
    """Parameter layout of diffusers.models.attention.Attention as used at modules.py:699-707."""

    def __init__(self, query_dim, cross_attention_dim=None, heads=8, dim_head=64, dropout=0.0, bias=False,
                 qk_norm=None, eps=1e-5, out_bias=True):
        super().__init__()
        if qk_norm != "layer_norm" or dim_head != 64:
            raise NotImplementedError("hivae_b200 attention kernels implement qk_norm='layer_norm' with 64-wide heads")
        inner = heads * dim_head
        self.heads, self.dim_head = heads, dim_head
        cross = cross_attention_dim if cross_attention_dim is not None else query_dim
        self.norm_q = nn.LayerNorm(dim_head, eps=eps)
        self.norm_k = nn.LayerNorm(dim_head, eps=eps)
        self.to_q = nn.Linear(query_dim, inner, bias=bias)
        self.to_k = nn.Linear(cross, inner, bias=bias)
        self.to_v = nn.Linear(cross, inner, bias=bias)
        self.to_out = nn.ModuleList([nn.Linear(inner, query_dim, bias=out_bias), nn.Dropout(dropout)])


class _GELUProj(nn.Module):
    def __init__(self, dim_in, dim_out, bias=True):
        super().__init__()
        self.proj = nn.Linear(dim_in, dim_out, bias=bias)


class FeedForward(nn.Module):
    """Parameter layout of diffusers FeedForward(activation_fn='gelu-approximate') (modules.py:712-719)."""

    def __init__(self, dim, dropout=0.0, activation_fn="gelu-approximate", final_dropout=True, inner_dim=None, bias=True):
        super().__init__()
        if activation_fn != "gelu-approximate":
            raise NotImplementedError("only gelu-approximate is implemented (the reference's only setting)")
        inner_dim = inner_dim or 4 * dim
        self.net = nn.ModuleList([_GELUProj(dim, inner_dim, bias), nn.Dropout(dropout), nn.Linear(inner_dim, dim, bias=bias)])
        if final_dropout:
            self.net.append(nn.Dropout(dropout))


class _AdaNorm(nn.Module):
    def __init__(self, conditioning_dim, embedding_dim, k, elementwise_affine=True, eps=1e-5, bias=True):
        super().__init__()
        self.embed_dim = embedding_dim
        self.silu = nn.SiLU()
        self.linear = nn.Linear(conditioning_dim, k * embedding_dim, bias=bias)
        self.norm = nn.LayerNorm(embedding_dim, eps=eps, elementwise_affine=elementwise_affine)


    @torch.no_grad()
    def _modulate(self, streams, temb):
        """standalone forward of the adaLN-Zero family: table = Linear(SiLU(temb)) by one small GEMM, then one
        LayerNorm + modulate kernel per stream; returns ([modulated streams], [gates (B, 1, D)]).  (Inside the blocks
        the table of ALL norms of a step comes from a single GEMM instead, engine.AdaPack.)"""
        ctx = Ctx(temb.device, torch.float32)
        D = self.embed_dim
        t32 = temb.float().contiguous()
        table = engine.small_linear(ctx, ops.silu(t32, torch.empty_like(t32)), ctx.f32(self.linear.weight),
                                    ctx.f32(self.linear.bias), "ada_standalone").clone()
        ada, outs, gates = Ada(table, 1, 1), [], []
        nw = ctx.f32(self.norm.weight) if self.norm.weight is not None else None
        nb = ctx.f32(self.norm.bias) if self.norm.bias is not None else None
        for i, x in enumerate(streams):
            res = x.float().contiguous()
            B, L, _ = res.shape
            y = torch.empty(B * L, D, dtype=torch.float32, device=res.device)
            o = 3 * i * D  # chunk order per stream: shift, scale, gate
            ops.ln_modulate(Rows(res, 0, L), nw, nb, self.norm.eps, y, ada, shift_mot=o, scale_mot=o + D, shift_img=o, scale_img=o + D)
            outs.append(y.view(B, L, D).to(x.dtype))
            gates.append(table[:, o + 2 * D:o + 3 * D][:, None, :].to(x.dtype))
        return outs, gates


class AMDLayerNormZero(_AdaNorm):
    """modules.py:881-904 (6 chunks: shift, scale, gate, enc_shift, enc_scale, enc_gate)."""

    def __init__(self, conditioning_dim, embedding_dim, elementwise_affine=True, eps=1e-5, bias=True):
        super().__init__(conditioning_dim, embedding_dim, 6, elementwise_affine, eps, bias)

    def forward(self, hidden_states, encoder_hidden_states, temb):
        (h, e), (g, eg) = self._modulate([hidden_states, encoder_hidden_states], temb)
        return h, e, g, eg


class AMDLayerNormZero_OneVariable(_AdaNorm):
    """modules.py:906-928 (3 chunks: shift, scale, gate)."""

    def __init__(self, conditioning_dim, embedding_dim, elementwise_affine=True, eps=1e-5, bias=True):
        super().__init__(conditioning_dim, embedding_dim, 3, elementwise_affine, eps, bias)

    def forward(self, hidden_states, temb):
        (h,), (g,) = self._modulate([hidden_states], temb)
        return h, g


class AMDLayerNormZero2Condition(_AdaNorm):
    """modules.py:930-956 (9 chunks: hidden, condition 1, condition 2, each shift, scale, gate)."""

    def __init__(self, conditioning_dim, embedding_dim, elementwise_affine=True, eps=1e-5, bias=True):
        super().__init__(conditioning_dim, embedding_dim, 9, elementwise_affine, eps, bias)

    def forward(self, hidden_states, condition_states1, condition_states2, temb):
        (h, c1, c2), (g, g1, g2) = self._modulate([hidden_states, condition_states1, condition_states2], temb)
        return h, c1, c2, g, g1, g2


class AdaLayerNorm(nn.Module):
    """modules.py:958-1013 with chunk_dim=1 (order shift, scale); used as the DiT's norm_out."""

    def __init__(self, embedding_dim, num_embeddings=None, output_dim=None, norm_elementwise_affine=False, norm_eps=1e-5,
                 chunk_dim=0):
        super().__init__()
        if num_embeddings is not None or chunk_dim != 1:
            raise NotImplementedError("only the chunk_dim=1 / temb form used by the DiT head is implemented")
        self.chunk_dim = chunk_dim
        output_dim = output_dim or embedding_dim * 2
        self.silu = nn.SiLU()
        self.linear = nn.Linear(embedding_dim, output_dim)
        self.norm = nn.LayerNorm(output_dim // 2, norm_eps, norm_elementwise_affine)

    @torch.no_grad()
    def forward(self, x, timestep=None, temb=None):
        """x (B, L, D), temb (B, E): norm(x) * (1 + scale) + shift with the (shift, scale) chunk order of chunk_dim=1"""
        ctx = Ctx(x.device, torch.float32)
        D = self.norm.normalized_shape[0]
        t32 = temb.float().contiguous()
        table = engine.small_linear(ctx, ops.silu(t32, torch.empty_like(t32)), ctx.f32(self.linear.weight),
                                    ctx.f32(self.linear.bias), "adaln_standalone")
        res = x.float().contiguous()
        B, L, _ = res.shape
        y = torch.empty(B * L, D, dtype=torch.float32, device=res.device)
        nw = ctx.f32(self.norm.weight) if self.norm.weight is not None else None
        nb = ctx.f32(self.norm.bias) if self.norm.bias is not None else None
        ops.ln_modulate(Rows(res, 0, L), nw, nb, self.norm.eps, y, Ada(table, 1, 1), shift_mot=0, scale_mot=D, shift_img=0, scale_img=D)
        return y.view(B, L, D).to(x.dtype)


class PatchEmbed(nn.Module):
    """modules.py:843-879."""

    def __init__(self, patch_size=2, in_channels=16, embed_dim=1920, bias=True):
        super().__init__()
        self.patch_size = patch_size
        self.proj = nn.Conv2d(in_channels, embed_dim, kernel_size=(patch_size, patch_size), stride=patch_size, bias=bias)

    @torch.no_grad()
    def forward(self, image_embeds):
        lead = image_embeds.shape[:-3]
        x = image_embeds.reshape(-1, *image_embeds.shape[-3:]).float().contiguous()
        D = self.proj.weight.shape[0]
        S = (x.shape[2] // self.patch_size) * (x.shape[3] // self.patch_size)
        out = torch.empty(x.shape[0], S, D, dtype=torch.float32, device=x.device)
        ops.patch_embed(x, None, self.patch_size, self.proj.weight.detach().float().contiguous(),
                        None if self.proj.bias is None else self.proj.bias.detach().float().contiguous(),
                        None, None, out, S, 0, D)
        if len(lead) == 2:
            return out.view(lead[0], lead[1] * S, D)
        return out


# ------------------------------------------------------------------------------------------------
class _Block(nn.Module):
    KIND = "basic"

    def _make(self, dim, heads, head_dim, dropout, activation_fn, attention_bias, qk_norm, final_dropout, ff_inner_dim,
              ff_bias, attention_out_bias, cross_dim=None):
        self.attn1 = Attention(query_dim=dim, cross_attention_dim=cross_dim, dim_head=head_dim, heads=heads,
                               qk_norm="layer_norm" if qk_norm else None, eps=1e-6, bias=attention_bias,
                               out_bias=attention_out_bias)
        self.ff = FeedForward(dim, dropout=dropout, activation_fn=activation_fn, final_dropout=final_dropout,
                              inner_dim=ff_inner_dim, bias=ff_bias)

    def _prepared(self, x):
        dtype = engine.compute_dtype_for(self)
        key = (x.device, dtype, engine.params_key(self))
        if getattr(self, "_hv_key", None) != key:
            self._hv_ctx = Ctx(x.device, dtype)
            self._hv_bp = engine.prepare_block(self._hv_ctx, self, self.KIND)
            self._hv_ada = engine.AdaPack(self._hv_ctx, [self._hv_bp]) if self._hv_bp.ada_src else None
            self._hv_key = key
            self._hv_ctx.version = engine.params_version(self)
        elif self._hv_ctx.version != engine.params_version(self):  # parameters updated in place: refresh the packed copies
            self._hv_ctx.refresh()
            self._hv_ctx.version = engine.params_version(self)
        return self._hv_ctx, self._hv_bp, self._hv_ada

    @staticmethod
    def _residual(*streams):
        """concatenate (B, L_i, D) streams into one fp32 residual buffer (frames, sum L_i, D)"""
        return torch.cat([s.float() for s in streams], dim=1).contiguous()


class BasicTransformerBlock(_Block):
    """modules.py:651-743: x += Attn(LN(x)); x += FF(LN(x))."""
    KIND = "basic"

    def __init__(self, dim, num_attention_heads, attention_head_dim, time_embed_dim=None, dropout=0.0,
                 activation_fn="gelu-approximate", attention_bias=False, qk_norm=True, norm_elementwise_affine=True,
                 norm_eps=1e-5, final_dropout=True, ff_inner_dim=None, ff_bias=True, attention_out_bias=True):
        super().__init__()
        self.norm1 = nn.LayerNorm(dim, eps=norm_eps, elementwise_affine=norm_elementwise_affine)
        self.norm2 = nn.LayerNorm(dim, eps=norm_eps, elementwise_affine=norm_elementwise_affine)
        self._make(dim, num_attention_heads, attention_head_dim, dropout, activation_fn, attention_bias, qk_norm,
                   final_dropout, ff_inner_dim, ff_bias, attention_out_bias)

    @torch.no_grad()
    def forward(self, hidden_states):
        ctx, bp, _ = self._prepared(hidden_states)
        res = self._residual(hidden_states)
        B, L, _ = res.shape
        engine.run_block(ctx, bp, Rows(res, 0, L), engine.dense_geom(B, L))
        return res.to(hidden_states.dtype)


class AMDTransformerMotionBlock(BasicTransformerBlock):
    """modules.py:1190-1285 with time_embed_dim=None (the only runnable form: `if temb:` on a tensor raises in the
    reference): plain pre-LN self-attention + feed-forward block."""

    def __init__(self, dim, num_attention_heads, attention_head_dim, time_embed_dim=None, **kw):
        if time_embed_dim:
            raise NotImplementedError("AMDTransformerMotionBlock with time_embed_dim is not runnable in the reference (modules.py:1250)")
        super().__init__(dim, num_attention_heads, attention_head_dim, **kw)

    @torch.no_grad()
    def forward(self, hidden_states, temb=None):
        assert temb is None
        return super().forward(hidden_states)


class BasicCrossTransformerBlock(_Block):
    """modules.py:745-840: queries from hidden_states, K/V from the (un-normalised) encoder_hidden_state."""
    KIND = "cross"

    def __init__(self, dim, cross_dim, num_attention_heads, attention_head_dim, time_embed_dim=None, dropout=0.0,
                 activation_fn="gelu-approximate", attention_bias=False, qk_norm=True, norm_elementwise_affine=True,
                 norm_eps=1e-5, final_dropout=True, ff_inner_dim=None, ff_bias=True, attention_out_bias=True):
        super().__init__()
        self.norm1 = nn.LayerNorm(dim, eps=norm_eps, elementwise_affine=norm_elementwise_affine)
        self.norm2 = nn.LayerNorm(dim, eps=norm_eps, elementwise_affine=norm_elementwise_affine)
        self._make(dim, num_attention_heads, attention_head_dim, dropout, activation_fn, attention_bias, qk_norm,
                   final_dropout, ff_inner_dim, ff_bias, attention_out_bias, cross_dim=cross_dim)

    @torch.no_grad()
    def forward(self, hidden_states, encoder_hidden_state):
        ctx, bp, _ = self._prepared(hidden_states)
        res = self._residual(hidden_states)
        B, Lq, _ = res.shape
        Lk = encoder_hidden_state.shape[1]
        kv = ctx.act(encoder_hidden_state.reshape(B * Lk, -1).float().contiguous(), "enc_dense")
        mq, mk = ops.seq_map(1, Lq, 0, 1), ops.seq_map(1, Lk, 0, 1)
        engine.run_block(ctx, bp, Rows(res, 0, Lq), engine.AttnGeom(B, Lq, Lk, mq, mk), kv_dense=kv)
        return res.to(hidden_states.dtype)


class AMDTransformerBlock(_Block):
    """modules.py:1016-1109: adaLN-Zero two-stream block; self-attention over [image || motion]."""
    KIND = "joint"

    def __init__(self, dim, num_attention_heads, attention_head_dim, time_embed_dim, dropout=0.0,
                 activation_fn="gelu-approximate", attention_bias=False, qk_norm=True, norm_elementwise_affine=True,
                 norm_eps=1e-5, final_dropout=True, ff_inner_dim=None, ff_bias=True, attention_out_bias=True):
        super().__init__()
        self.norm1 = AMDLayerNormZero(time_embed_dim, dim, norm_elementwise_affine, norm_eps, bias=True)
        self.norm2 = AMDLayerNormZero(time_embed_dim, dim, norm_elementwise_affine, norm_eps, bias=True)
        self._make(dim, num_attention_heads, attention_head_dim, dropout, activation_fn, attention_bias, qk_norm,
                   final_dropout, ff_inner_dim, ff_bias, attention_out_bias)

    @torch.no_grad()
    def forward(self, hidden_states, temb, encoder_hidden_states=None):
        ctx, bp, adap = self._prepared(hidden_states)
        Lm = hidden_states.shape[1]
        if encoder_hidden_states is None:
            # reference quirk (modules.py:1076-1077): the stream is used as its own encoder stream
            encoder_hidden_states = hidden_states
        Li = encoder_hidden_states.shape[1]
        res = self._residual(hidden_states, encoder_hidden_states)  # [motion | image]
        B = res.shape[0]
        ada = Ada(adap.table(ctx, temb.float().contiguous()))
        engine.run_block(ctx, bp, Rows(res, 0, Lm + Li, img_start=Lm, img_len=Li), engine.dense_geom(B, Lm + Li), ada)
        return res[:, :Lm].to(hidden_states.dtype), res[:, Lm:].to(hidden_states.dtype)


class BasicDiTBlock(_Block):
    """modules.py:1111-1188: adaLN-Zero single-stream block."""
    KIND = "dit"

    def __init__(self, dim, num_attention_heads, attention_head_dim, time_embed_dim, dropout=0.0,
                 activation_fn="gelu-approximate", attention_bias=False, qk_norm=True, norm_elementwise_affine=True,
                 norm_eps=1e-5, final_dropout=True, ff_inner_dim=None, ff_bias=True, attention_out_bias=True):
        super().__init__()
        self.norm1 = AMDLayerNormZero_OneVariable(time_embed_dim, dim, norm_elementwise_affine, norm_eps, bias=True)
        self.norm2 = AMDLayerNormZero_OneVariable(time_embed_dim, dim, norm_elementwise_affine, norm_eps, bias=True)
        self._make(dim, num_attention_heads, attention_head_dim, dropout, activation_fn, attention_bias, qk_norm,
                   final_dropout, ff_inner_dim, ff_bias, attention_out_bias)

    @torch.no_grad()
    def forward(self, hidden_states, temb):
        ctx, bp, adap = self._prepared(hidden_states)
        res = self._residual(hidden_states)
        B, L, _ = res.shape
        ada = Ada(adap.table(ctx, temb.float().contiguous()))
        engine.run_block(ctx, bp, Rows(res, 0, L), engine.dense_geom(B, L), ada)
        return res.to(hidden_states.dtype)
