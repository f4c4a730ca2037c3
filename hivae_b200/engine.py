"""Execution engine: prepared (packed / down-cast) weights, workspace, and the block runner
that strings the sm_100a kernels together.

Data layout in HBM: token streams are fp32 residual buffers
(frames, frame_rows, D); each transformer block works on a contiguous row range of every
frame; all intermediate activations (LN output, QKV, attention output, MLP hidden) are dense
(M, *) matrices in the compute dtype (bf16 or fp32) that only ever feed the next GEMM.
"""
from dataclasses import dataclass, field
from typing import Optional

import torch

from . import ops
from .ops import Ada, Rows


def compute_dtype_for(module):
    """bf16 when called under torch.autocast(cuda, bf16) (the reference's only low-precision mode,
    train_amd.py:160-165) or when the module was tagged with `hv_compute_dtype`; fp32 otherwise."""
    tagged = getattr(module, "hv_compute_dtype", None)
    if tagged is not None:
        return tagged
    if torch.is_autocast_enabled("cuda") and torch.get_autocast_dtype("cuda") == torch.bfloat16:
        return torch.bfloat16
    return torch.float32


# Bumped whenever parameters are updated behind torch's back (optim.AdamW.step writes the flat arena through raw
# pointers, which never touches `p._version`): prepared weight copies older than this are refreshed in place.
WEIGHT_EPOCH = [0]


def bump_weight_epoch():
    WEIGHT_EPOCH[0] += 1


def params_key(module):
    """identity of a module's parameter storage: changes when parameters are re-homed (FlatArena, .to(device))"""
    return hash(tuple(p.data_ptr() for p in module.parameters()))


def params_version(module):
    return (WEIGHT_EPOCH[0], sum(p._version for p in module.parameters()))


class Ctx:
    """Device, compute dtype and a name-keyed workspace (stable addresses -> TMA descriptor cache
    hits, CUDA-graph friendly).  Every packed / down-cast weight copy is recorded as a binding
    (destination rows <- source parameters) so that `refresh()` re-derives it in place after an optimizer step."""

    def __init__(self, device, dtype):
        assert dtype in (torch.float32, torch.bfloat16)
        self.device, self.dtype = torch.device(device), dtype
        self._ws = {}
        self.bindings = []
        self.version = None
        self.generation = 0

    def _bind(self, dst, srcs):
        self.bindings.append((dst, list(srcs)))
        return dst

    def refresh(self):
        """re-copy every derived weight from its (updated) source parameters; addresses stay the same"""
        self.generation += 1
        for dst, srcs in self.bindings:
            off = 0
            for p in srcs:
                n = p.shape[0]
                src = p.detach()
                if src.device != self.device or src.dtype != torch.float32 or not src.is_contiguous():
                    src = src.to(self.device, torch.float32).contiguous()
                ops.convert(src, dst[off:off + n])
                off += n

    def pack(self, params, gemm=True):
        """the parameters concatenated along dim 0: a GEMM weight (bf16 when the tcgen05 path applies) or an fp32
        vector / matrix; recorded for refresh()"""
        cat = torch.cat([self.f32(p, bind=False) for p in params], 0).contiguous() if len(params) > 1 else self.f32(params[0], bind=False)
        if gemm:
            N, K = cat.shape
            if self.dtype == torch.bfloat16 and K >= 64 and K % 8 == 0 and N >= 64:
                return self._bind(cat.to(torch.bfloat16), params)
        if len(params) == 1 and cat.data_ptr() == params[0].data_ptr():
            return cat  # live alias of the parameter storage
        return self._bind(cat.clone() if cat.data_ptr() == params[0].data_ptr() else cat, params)

    def buf(self, name, shape, dtype=None, zero=False):
        """zero=True: zero-filled at creation (for buffers whose padding is never written afterwards)"""
        dtype = dtype or self.dtype
        key = (name, tuple(int(s) for s in shape), dtype)
        t = self._ws.get(key)
        if t is None:
            t = self._ws[key] = (torch.zeros if zero else torch.empty)(key[1], dtype=dtype, device=self.device)
        return t

    def f32(self, t, bind=True):
        if t is None:
            return None
        out = t.detach().to(self.device, torch.float32).contiguous()
        if bind and out.data_ptr() != t.data_ptr():
            self._bind(out, [t])
        return out

    def weight(self, w):
        """GEMM weight (N,K): bf16 copy for the tcgen05 path when it applies, fp32 otherwise."""
        return self.pack([w])

    def act(self, x, name):
        """activation matrix in the dtype the GEMM with weight dtype == self.dtype expects"""
        if x.dtype == self.dtype and x.is_contiguous():
            return x
        out = self.buf(name, x.shape)
        return ops.convert(x.contiguous(), out)


def _gemm_in(ctx, A, W, name):
    """bf16 weights need bf16 activations; fp32 weights take either."""
    if W.dtype == torch.bfloat16 and A.dtype != torch.bfloat16:
        return ops.convert(A, ctx.buf(name, A.shape, torch.bfloat16))
    return A


# ------------------------------------------------------------------------------------------------
@dataclass
class BlockP:
    """Prepared weights of one transformer block (any of the four reference block classes)."""
    kind: str  # 'basic' | 'cross' | 'joint' | 'dit'
    D: int
    heads: int
    eps: float
    n1w: Optional[torch.Tensor]
    n1b: Optional[torch.Tensor]
    n2w: Optional[torch.Tensor]
    n2b: Optional[torch.Tensor]
    qn_w: torch.Tensor
    qn_b: torch.Tensor
    kn_w: torch.Tensor
    kn_b: torch.Tensor
    wo: torch.Tensor
    bo: Optional[torch.Tensor]
    w1: torch.Tensor
    b1: Optional[torch.Tensor]
    w2: torch.Tensor
    b2: Optional[torch.Tensor]
    wqkv: Optional[torch.Tensor] = None
    bqkv: Optional[torch.Tensor] = None
    wq: Optional[torch.Tensor] = None
    bq: Optional[torch.Tensor] = None
    wkv: Optional[torch.Tensor] = None
    bkv: Optional[torch.Tensor] = None
    ada_off1: int = 0  # column offsets of this block's norm1 / norm2 chunks in the adaLN table
    ada_off2: int = 0
    ada_src: list = field(default_factory=list)  # [(weight, bias)] of norm1.linear, norm2.linear


def _cat_bias(ctx, *bs):
    if any(b is None for b in bs):
        return None
    return ctx.pack(list(bs), gemm=False)


def prepare_block(ctx, blk, kind):
    """blk: a module with the reference's parameter names (norm1, attn1.{to_q,to_k,to_v,norm_q,norm_k,
    to_out.0}, norm2, ff.net.0.proj, ff.net.2)."""
    at, ff = blk.attn1, blk.ff
    ada = kind in ("joint", "dit")
    n1 = blk.norm1.norm if ada else blk.norm1
    n2 = blk.norm2.norm if ada else blk.norm2
    D = at.to_q.weight.shape[0]
    bp = BlockP(
        kind=kind, D=D, heads=at.heads, eps=n1.eps,
        n1w=ctx.f32(n1.weight), n1b=ctx.f32(n1.bias), n2w=ctx.f32(n2.weight), n2b=ctx.f32(n2.bias),
        qn_w=ctx.f32(at.norm_q.weight), qn_b=ctx.f32(at.norm_q.bias),
        kn_w=ctx.f32(at.norm_k.weight), kn_b=ctx.f32(at.norm_k.bias),
        wo=ctx.weight(at.to_out[0].weight), bo=ctx.f32(at.to_out[0].bias),
        w1=ctx.weight(ff.net[0].proj.weight), b1=ctx.f32(ff.net[0].proj.bias),
        w2=ctx.weight(ff.net[2].weight), b2=ctx.f32(ff.net[2].bias),
    )
    if kind == "cross":
        bp.wq, bp.bq = ctx.weight(at.to_q.weight), ctx.f32(at.to_q.bias)
        bp.wkv = ctx.pack([at.to_k.weight, at.to_v.weight])
        bp.bkv = _cat_bias(ctx, at.to_k.bias, at.to_v.bias)
    else:
        bp.wqkv = ctx.pack([at.to_q.weight, at.to_k.weight, at.to_v.weight])
        bp.bqkv = _cat_bias(ctx, at.to_q.bias, at.to_k.bias, at.to_v.bias)
    if ada:
        bp.ada_src = [(blk.norm1.linear.weight, blk.norm1.linear.bias), (blk.norm2.linear.weight, blk.norm2.linear.bias)]
    return bp


class AdaPack:
    """All adaLN Linear(temb_dim -> k*D) layers that share one temb, concatenated into a single GEMM
    (SURVEY K12): table = Linear_cat(SiLU(temb)), fp32, one column range per norm."""

    def __init__(self, ctx, blocks, extra=()):
        ws, bs, off = [], [], 0
        for bp in blocks:
            if not bp.ada_src:
                continue
            (w1, b1), (w2, b2) = bp.ada_src
            bp.ada_off1, off = off, off + w1.shape[0]
            bp.ada_off2, off = off, off + w2.shape[0]
            ws += [w1, w2]
            bs += [b1, b2]
        self.extra_off = []
        for w, b in extra:
            self.extra_off.append(off)
            off += w.shape[0]
            ws.append(w)
            bs.append(b)
        self.N = off
        self.w = ctx.pack(ws)
        self.b = ctx.pack(bs, gemm=False)

    def table(self, ctx, temb, name="ada"):
        """temb (rows, E) fp32 -> (rows, N) fp32"""
        rows = temb.shape[0]
        s = ops.silu(temb, ctx.buf(name + "_silu", temb.shape, self.w.dtype if self.w.dtype == torch.bfloat16 else torch.float32))
        out = ctx.buf(name, (rows, self.N), torch.float32)
        ops.linear(s, self.w, ops.epi_store(out, self.b))
        return out


@dataclass
class AttnGeom:
    seqs: int
    Lq: int
    Lkv: int
    qmap: object
    kvmap: object


def dense_geom(frames, L):
    m = ops.seq_map(1, L, 0, 1)
    return AttnGeom(frames, L, L, m, m)


def strided_geom(outer, inner, T):
    """sequences (o, s) for o < outer, s < inner, tokens t < T at dense row (o*T + t)*inner + s
    (the '(n t) s d -> (n s) t d' regrouping of transformer.py:2022 without moving any data)."""
    m = ops.seq_map(inner, T * inner, 1, inner)
    return AttnGeom(outer * inner, T, T, m, m)


def run_block(ctx, bp, rows, geom, ada=None, kv_dense=None):
    """One transformer block over `rows` (in place on the fp32 residual buffer).
    basic/cross: modules.py:721-743 / 817-840; joint: 1065-1109; dit: 1160-1188."""
    M, D = rows.M, bp.D
    lib_scale = 1.0 / (D // bp.heads) ** 0.5
    hd = D // bp.heads
    if bp.kind == "joint":
        o1, o2 = bp.ada_off1, bp.ada_off2
        c1 = dict(shift_mot=o1, scale_mot=o1 + D, shift_img=o1 + 3 * D, scale_img=o1 + 4 * D)
        c2 = dict(shift_mot=o2, scale_mot=o2 + D, shift_img=o2 + 3 * D, scale_img=o2 + 4 * D)
        g1, g2 = (o1 + 2 * D, o1 + 5 * D), (o2 + 2 * D, o2 + 5 * D)
    elif bp.kind == "dit":
        o1, o2 = bp.ada_off1, bp.ada_off2
        c1 = dict(shift_mot=o1, scale_mot=o1 + D, shift_img=o1, scale_img=o1 + D)
        c2 = dict(shift_mot=o2, scale_mot=o2 + D, shift_img=o2, scale_img=o2 + D)
        g1, g2 = (o1 + 2 * D, o1 + 2 * D), (o2 + 2 * D, o2 + 2 * D)
    else:
        ada, c1, c2, g1, g2 = None, {}, {}, (0, 0), (0, 0)

    xn = ctx.buf("xn", (M, D))
    ops.ln_modulate(rows, bp.n1w, bp.n1b, bp.eps, xn, ada, **c1)
    if bp.kind == "cross":
        q = ctx.buf("q", (M, D))
        ops.linear(_gemm_in(ctx, xn, bp.wq, "xn16"), bp.wq, ops.epi_qknorm(q, bp.bq, bp.qn_w, bp.qn_b, None, None, D, 0))
        kv = ctx.buf("kv", (kv_dense.shape[0], 2 * D))
        ops.linear(_gemm_in(ctx, kv_dense, bp.wkv, "kvd16"), bp.wkv, ops.epi_qknorm(kv, bp.bkv, None, None, bp.kn_w, bp.kn_b, 0, D))
        qv, kk, vv = q, kv[:, :D], kv[:, D:]
    else:
        qkv = ctx.buf("qkv", (M, 3 * D))
        ops.linear(_gemm_in(ctx, xn, bp.wqkv, "xn16"), bp.wqkv,
                   ops.epi_qknorm(qkv, bp.bqkv, bp.qn_w, bp.qn_b, bp.kn_w, bp.kn_b, D, D))
        qv, kk, vv = qkv[:, :D], qkv[:, D:2 * D], qkv[:, 2 * D:]
    att = ctx.buf("att", (M, D))
    ops.attention(qv, kk, vv, att, geom.seqs, bp.heads, hd, geom.Lq, geom.Lkv, geom.qmap, geom.kvmap, lib_scale)
    ops.linear(_gemm_in(ctx, att, bp.wo, "att16"), bp.wo, ops.epi_residual(rows, bp.bo, ada, *g1))
    ops.ln_modulate(rows, bp.n2w, bp.n2b, bp.eps, xn, ada, **c2)
    h = ctx.buf("ffh", (M, bp.w1.shape[0]))
    ops.linear(_gemm_in(ctx, xn, bp.w1, "xn16"), bp.w1, ops.epi_store(h, bp.b1, gelu=True))
    ops.linear(_gemm_in(ctx, h, bp.w2, "ffh16"), bp.w2, ops.epi_residual(rows, bp.b2, ada, *g2))


def small_linear(ctx, x, w, b, name):
    """fp32 output Linear for the skinny / tiny-K layers (x: (M,K) any dtype; w fp32 or bf16)."""
    out = ctx.buf(name, (x.shape[0], w.shape[0]), torch.float32)
    ops.linear(_gemm_in(ctx, x, w, name + "_in"), w, ops.epi_store(out, b))
    return out
