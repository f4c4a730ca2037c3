"""Building blocks of the backward pass of the blocks (`accelerator.backward(loss)`, train_amd.py:489)
on the sm_100a kernels of csrc/backward.cu and the tcgen05 GEMM: gradients of nn.Linear, nn.LayerNorm and the
tanh-GELU of FeedForward, the dense attention and the QK-norm, and on top of them the whole-block backward of
BasicTransformerBlock (the motion encoders' block) behind a torch.autograd.Function.  The adaLN /
gated-residual blocks of the denoiser are the next step.

No CPU fallback: every function raises without the CUDA library."""
import ctypes as C
import os

import torch

from . import _lib, ops
from .ops import _chk_cuda, _dt, _p, _stream


_SPLITK = os.environ.get("HV_WGRAD_SPLITK", "1") != "0"
_TN = os.environ.get("HV_WGRAD_TN", "1") != "0"   # weight gradients on MN-major operands (no transposed copies); 0 = round-1 path


def _timed(name):
    """records the wrapped call in ops.PROFILE (per-launch CUDA events) when the bench's profile leg is on"""
    def deco(fn):
        def wrapper(*a, **k):
            tok = ops._prof(name)
            try:
                return fn(*a, **k)
            finally:
                ops._prof_end(tok)
        wrapper.__name__, wrapper.__doc__ = fn.__name__, fn.__doc__
        return wrapper
    return deco


@_timed("transpose")
def transpose2d(x, out=None):
    """(R, C) -> (C, R), bf16 or fp32; x may be a row-strided view with unit column stride"""
    _lib.require_device()
    assert x.dim() == 2 and x.is_cuda and (x.shape[1] == 1 or x.stride(1) == 1)
    R, Cc = x.shape
    out = torch.empty(Cc, R, dtype=x.dtype, device=x.device) if out is None else out
    _lib.check(_lib.load().hv_transpose2d(_p(x), _p(out), R, Cc, x.stride(0), out.stride(0), _dt(x), _stream()))
    return out


@_timed("transpose_colsum")
def transpose2d_colsum(x):
    """(x^T, column sums of x) in one pass over x when the fused kernel applies (bf16, sizes / strides multiples of 8,
    16-byte aligned), otherwise the two separate kernels"""
    _lib.require_device()
    assert x.dim() == 2 and x.is_cuda and x.stride(1) == 1
    R, Cc = x.shape
    if x.dtype != torch.bfloat16 or (R | Cc | x.stride(0)) & 7 or x.data_ptr() & 15 or R == 0:
        return transpose2d(x), colsum(x)
    out = torch.empty(Cc, R, dtype=x.dtype, device=x.device)
    ws = torch.empty(-(-R // 64) * Cc, dtype=torch.float32, device=x.device)
    cs = torch.empty(Cc, dtype=torch.float32, device=x.device)
    _lib.check(_lib.load().hv_transpose2d_colsum(_p(x), _p(out), R, Cc, x.stride(0), out.stride(0), _dt(x), _p(ws), _p(cs), _stream()))
    return out, cs


@_timed("colsum")
def colsum(x):
    """sum over rows of a (M, N) matrix -> fp32 (N,)  (bias gradient)"""
    _lib.require_device()
    assert x.dim() == 2 and x.is_cuda and x.stride(1) == 1
    M, N = x.shape
    lib = _lib.load()
    ws = torch.empty(lib.hv_colsum_slices() * N, dtype=torch.float32, device=x.device)
    out = torch.empty(N, dtype=torch.float32, device=x.device)
    _lib.check(lib.hv_colsum(_p(x), M, N, x.stride(0), _dt(x), _p(ws), _p(out), _stream()))
    return out


def colsum_into(x, dests, accumulate=True):
    """column sums of x (M, N) written (or added, accumulate=True) straight into up to three contiguous fp32 destinations that
    together cover the N columns in order: the bias gradient(s) of a (packed) Linear go into their gradient-arena slots
    without a temporary and an axpy per destination"""
    _lib.require_device()
    assert x.dim() == 2 and x.is_cuda and x.stride(1) == 1 and 1 <= len(dests) <= 3
    M, N = x.shape
    assert sum(t.numel() for t in dests) == N
    for t in dests:
        assert t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()
    lib = _lib.load()
    ws = torch.empty(lib.hv_colsum_slices() * N, dtype=torch.float32, device=x.device)
    d = list(dests) + [None] * (3 - len(dests))
    n0 = dests[0].numel()
    n1 = dests[1].numel() if len(dests) > 1 else 0
    _lib.check(lib.hv_colsum_acc(_p(x), M, N, x.stride(0), _dt(x), _p(ws), _p(d[0]), n0, _p(d[1]) if d[1] is not None else None, n1,
                                 _p(d[2]) if d[2] is not None else None, 1 if accumulate else 0, _stream()))


def linear_backward(dy, x, w, need_dx=True, dx_dtype=torch.bfloat16, w_t=None):
    """Gradients of y = x @ w.T + b for bf16 operands (fp32 accumulation on the tensor cores):
    dy (M, N), x (M, K), w (N, K)  ->  dx (M, K) in `dx_dtype`, dw (N, K) fp32, db (N,) fp32.
    dX = dY W and dW = dY^T X both run on the tcgen05 GEMM (C = A B^T) with transposed bf16 operand copies."""
    for t in (dy, x, w):
        assert t.dtype == torch.bfloat16 and t.dim() == 2
    _chk_cuda(dy, x, w)
    M, N = dy.shape
    K = x.shape[1]
    assert x.shape[0] == M and w.shape == (N, K)
    dx = None
    if need_dx:
        dx = torch.empty(M, K, dtype=dx_dtype, device=dy.device)
        ops.linear(dy, w_t if w_t is not None else transpose2d(w), ops.epi_store(dx))  # (M, N) x (K, N)^T; W^T is reusable for a whole step
    # dW = dY^T X: (N, M) x (K, M)^T.  The output has only ceil(N/256) * ceil(K/256) tiles while the contraction runs over
    # all M tokens, so the tokens are split over `ks` work items per tile that reduce-add into a zeroed dW (fp32 partials
    # are added in arrival order; HV_WGRAD_SPLITK=0 restores the single-pass, bit-reproducible form)
    dy_t, db = transpose2d_colsum(dy)
    tiles = -(-N // 256) * -(-K // 256)
    ks = max(1, min(16, 74 // tiles, M // 512)) if _SPLITK and N >= 32 and K % 4 == 0 else 1
    if ks > 1:
        dw = torch.zeros(N, K, dtype=torch.float32, device=dy.device)
        e = ops.epi_residual(ops.Rows(dw.view(1, N, K), 0, N))
        e.k_split = ks
        ops.linear(dy_t, transpose2d(x), e)
    else:
        dw = torch.empty(N, K, dtype=torch.float32, device=dy.device)
        ops.linear(dy_t, transpose2d(x), ops.epi_store(dw))
    return dx, dw, db


@_timed("gelu_bwd")
def gelu_tanh_backward(pre, dy, out_dtype=None):
    """dx = dy * gelu_tanh'(pre); pre / dy bf16 or fp32 (same dtype), dx in out_dtype (default: same)"""
    _chk_cuda(pre, dy)
    assert pre.shape == dy.shape and pre.dtype == dy.dtype
    dx = torch.empty(pre.shape, dtype=out_dtype or pre.dtype, device=pre.device)
    _lib.check(_lib.load().hv_gelu_tanh_bwd(_p(pre), _p(dy), _p(dx), pre.numel(), _dt(pre), _dt(dx), _stream()))
    return dx


@_timed("ln_bwd")
def layernorm_backward(x, dy, weight, eps, dx=None, accumulate=False, affine_grads=True):
    """nn.LayerNorm backward over the last dimension: x fp32 (..., D), dy fp32 / bf16 -> (dx fp32, dw, db).
    dx may be an existing fp32 buffer to write or accumulate into (the residual-stream gradient)."""
    _chk_cuda(x, dy, weight, dx)
    assert x.dtype == torch.float32 and x.shape == dy.shape
    D = x.shape[-1]
    rows = x.numel() // D
    lib = _lib.load()
    dx = torch.empty_like(x) if dx is None else dx
    ws = torch.empty((lib.hv_layernorm_bwd_ctas() + 1) * 2 * D, dtype=torch.float32, device=x.device)
    dw = torch.empty(D, dtype=torch.float32, device=x.device) if affine_grads else None
    db = torch.empty(D, dtype=torch.float32, device=x.device) if affine_grads else None
    _lib.check(lib.hv_layernorm_bwd(_p(x), _p(dy), _dt(dy), _p(weight), float(eps), rows, D, _p(dx), int(bool(accumulate)),
                                    _p(ws), _p(dw), _p(db), _stream()))
    return dx, dw, db


@_timed("attention_bwd")
def attention_backward(q, k, v, out, dout, seqs, heads, Lq, Lkv, qmap, kvmap, scale, dq=None, dk=None, dv=None, stats=None,
                       have_lse=False):
    """Gradients of ops.attention (head_dim 64, no mask): q / k / v / out / dout are 2-D (rows, heads*64) views in one
    dtype (bf16 or fp32) with unit column stride -> (dq, dk, dv) of the same dtype, dense (rows, heads*64)."""
    _lib.require_device()
    for t in (q, k, v, out, dout):
        assert t.is_cuda and t.dim() == 2 and t.stride(1) == 1 and t.dtype == q.dtype
    new = lambda t, ref: torch.empty(ref.shape[0], heads * 64, dtype=q.dtype, device=q.device) if t is None else t
    dq, dk, dv = new(dq, q), new(dk, k), new(dv, v)   # callers may pass column slices of a wider gradient matrix
    for t in (dq, dk, dv):
        assert t.is_cuda and t.dim() == 2 and t.stride(1) == 1 and t.dtype == q.dtype and t.shape[1] == heads * 64
    # stats + have_lse: the workspace that ops.attention(..., lse=stats) filled in the forward (it returned True)
    assert not have_lse or stats is not None
    if stats is None:
        stats = torch.empty(seqs * heads * Lq * 2, dtype=torch.float32, device=q.device)
    assert stats.is_cuda and stats.dtype == torch.float32 and stats.is_contiguous() and stats.numel() >= seqs * heads * Lq * 2
    _lib.check(_lib.load().hv_attention_bwd_lse(
        _p(q), q.stride(0), _p(k), k.stride(0), _p(v), v.stride(0), _p(out), out.stride(0), _p(dout), dout.stride(0),
        _p(dq), dq.stride(0), _p(dk), dk.stride(0), _p(dv), dv.stride(0), _dt(q), seqs, heads, 64, Lq, Lkv,
        C.byref(qmap), C.byref(kvmap), float(scale), _p(stats), 1 if have_lse else 0, _stream()))
    return dq, dk, dv


@_timed("headnorm_bwd")
def headnorm64_backward(x, dy, weight, heads, eps=1e-6, out=None, dw_acc=None, db_acc=None):
    """backward of the per-head LayerNorm(64) of q / k (QK-norm): x (pre-norm), dy are (tokens, heads*64) views (bf16 or
    fp32, unit column stride) -> (dx dense, dw (64,), db (64,))"""
    _lib.require_device()
    assert x.is_cuda and x.dim() == 2 and x.shape == dy.shape and x.shape[1] == heads * 64 and x.dtype == dy.dtype
    assert x.stride(1) == 1 and dy.stride(1) == 1
    lib = _lib.load()
    dx = torch.empty(x.shape, dtype=x.dtype, device=x.device) if out is None else out   # `out`: e.g. the q columns of d(qkv)
    assert dx.shape == x.shape and dx.dtype == x.dtype and dx.stride(1) == 1
    ws = torch.empty((lib.hv_layernorm_bwd_ctas() + 1) * 128, dtype=torch.float32, device=x.device)
    if dw_acc is not None or db_acc is not None:
        # dw / db are ADDED to the given fp32 (64,) tensors (gradient-arena slots); a missing one is not computed
        for t in (dw_acc, db_acc):
            assert t is None or (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() and t.numel() == 64)
        _lib.check(lib.hv_headnorm64_bwd_acc(_p(x), x.stride(0), _p(dy), dy.stride(0), _p(weight), float(eps), x.shape[0], heads,
                                             _p(dx), dx.stride(0), _dt(x), _p(ws), _p(dw_acc) if dw_acc is not None else None,
                                             _p(db_acc) if db_acc is not None else None, 1, _stream()))
        return dx, dw_acc, db_acc
    dw, db = (torch.empty(64, dtype=torch.float32, device=x.device) for _ in range(2))
    _lib.check(lib.hv_headnorm64_bwd(_p(x), x.stride(0), _p(dy), dy.stride(0), _p(weight), float(eps), x.shape[0], heads,
                                     _p(dx), dx.stride(0), _dt(x), _p(ws), _p(dw), _p(db), _stream()))
    return dx, dw, db


# ------------------------------------------------------------------------------------------------
# kernels of csrc/train_ops.cu: backward of the adaLN-Zero blocks and of the small layers around them

def _ws_floats(frames, L, D, per=1, groups=0):
    return (frames * (L // _lib.load().hv_train_chunk_rows() + 3) * per + groups * 2) * D


@_timed("gate_bwd")
def gate_bwd(drows, y, ada, gate_mot, gate_img, dY, dtable=None):
    """dY (dense M x D, fp32 / bf16) = gate * dres rows; dtable[cond row, gate column] += sum_rows dres * y.
    ada None: plain residual branch -> dY = the gradient rows, gathered and converted."""
    _chk_cuda(drows.res, y, dY)
    t = ada.table if ada is not None else None
    ws = torch.empty(_ws_floats(drows.frames, drows.L, drows.D), dtype=torch.float32, device=dY.device) if t is not None else None
    _lib.check(_lib.load().hv_gate_bwd(
        _p(drows.res), drows.frames, drows.frame_rows, drows.row_off, drows.L, drows.D, _p(y), _p(t),
        t.shape[1] if t is not None else 0, gate_mot, gate_img, drows.img_start, drows.img_len,
        ada.cond_div if ada else 1, ada.cond_mul if ada else 1, _p(dY), _dt(dY), _p(dtable), _p(ws), ws.numel() if ws is not None else 0,
        _stream()))
    return dY


@_timed("ln_modulate_bwd")
def ln_modulate_bwd(x, dy, ln_w, ln_b, eps, drows, ada=None, dtable=None, dw=None, db=None, shift_mot=0, scale_mot=0, shift_img=0,
                    scale_img=0):
    """backward of ops.ln_modulate: x (M, D) fp32 = the saved input rows, dy (M, D) fp32 / bf16; the input gradient is
    accumulated into `drows`, dshift / dscale into dtable, dw / db (fp32, D) are accumulated"""
    _chk_cuda(x, dy, ln_w, ln_b, drows.res, dtable, dw, db)
    assert x.dtype == torch.float32 and x.shape == dy.shape and x.shape[1] == drows.D and x.shape[0] == drows.M
    t = ada.table if ada is not None else None
    cond_div = ada.cond_div if ada is not None else 8
    groups = -(-drows.frames // cond_div)
    ws = torch.empty(_ws_floats(drows.frames, drows.L, drows.D, 2, groups), dtype=torch.float32, device=x.device)
    _lib.check(_lib.load().hv_ln_modulate_bwd(
        _p(x), _p(dy), _dt(dy), _p(ln_w), _p(ln_b), float(eps), _p(drows.res), drows.frames, drows.frame_rows, drows.row_off,
        drows.L, drows.D, _p(t), t.shape[1] if t is not None else 0, shift_mot, scale_mot, shift_img, scale_img,
        drows.img_start, drows.img_len, cond_div, ada.cond_mul if ada else 1, _p(dtable), _p(dw), _p(db), _p(ws), ws.numel(),
        _stream()))


@_timed("gate_residual")
def gate_residual(rows, y, ada=None, gate_mot=0, gate_img=0):
    """res rows += gate * y (training forward: the branch output y is kept for the gate gradient)"""
    _chk_cuda(rows.res, y)
    t = ada.table if ada is not None else None
    _lib.check(_lib.load().hv_gate_residual(
        _p(rows.res), rows.frames, rows.frame_rows, rows.row_off, rows.L, rows.D, _p(y), _dt(y), _p(t),
        t.shape[1] if t is not None else 0, gate_mot, gate_img, rows.img_start, rows.img_len,
        ada.cond_div if ada else 1, ada.cond_mul if ada else 1, _stream()))


def silu_backward(x, dy, out):
    _chk_cuda(x, dy, out)
    assert x.dtype == torch.float32 and x.shape == dy.shape == out.shape
    _lib.check(_lib.load().hv_silu_bwd(_p(x), _p(dy), _dt(dy), _p(out), _dt(out), x.numel(), _stream()))
    return out


def mse_backward(pre, gt, gout=None):
    """d mse_loss(pre, gt) / d pre * gout (gout: 0-dim / 1-element fp32 device tensor or None)"""
    _chk_cuda(pre, gt, gout)
    assert pre.dtype == gt.dtype == torch.float32 and pre.shape == gt.shape
    out = torch.empty_like(pre)
    _lib.check(_lib.load().hv_mse_bwd(_p(pre), _p(gt), pre.numel(), _p(gout), _p(out), _stream()))
    return out


def rows_reduce(src, D, n_out, base, stride_out, A, stride_a, B=1, stride_b=0, out=None, accumulate=False):
    """out[o] (+)= sum_{a, b} src row (base + o*stride_out + a*stride_a + b*stride_b); rows of D floats"""
    _chk_cuda(src, out)
    out = torch.empty(n_out, D, dtype=torch.float32, device=src.device) if out is None else out
    _lib.check(_lib.load().hv_rows_reduce(_p(src), D, n_out, base, stride_out, A, stride_a, B, stride_b, _p(out), int(accumulate), _stream()))
    return out


def rows_index(src, out, idx, frames, S, keep, D, in_frame_rows, out_frame_rows, scatter=False):
    """src / out may be row-range views of a frames buffer (their frame strides are passed explicitly)"""
    _chk_cuda(idx)
    for t in (src, out):
        assert t.is_cuda and t.dtype == torch.float32 and t.stride(-1) == 1 and (t.dim() < 2 or t.stride(-2) == D)
    assert idx.dtype == torch.int32
    _lib.check(_lib.load().hv_rows_index(_p(src), _p(out), _p(idx), frames, S, keep, D, in_frame_rows, out_frame_rows, int(scatter), _stream()))
    return out


def conv3x3_backward(x, w, dy, stride, dw=None, db=None, need_dx=True):
    _chk_cuda(x, w, dy, dw, db)
    B_, Cin, H, W = x.shape
    dx = torch.empty_like(x) if need_dx else None
    _lib.check(_lib.load().hv_conv3x3_bwd(_p(x), _p(w), _p(dy), _p(dx), _p(dw), _p(db), B_, Cin, w.shape[0], H, W, stride, _stream()))
    return dx


def maxpool2_backward(x, dy):
    _chk_cuda(x, dy)
    B_, Cc, H, W = x.shape
    dx = torch.empty_like(x)
    _lib.check(_lib.load().hv_maxpool2_bwd(_p(x), _p(dy), _p(dx), B_ * Cc, H, W, _stream()))
    return dx


def weight_t(ctx, w):
    """W^T of a prepared (packed) weight of `ctx`, cached in the context (same lifetime as the weights it was derived
    from) and recomputed in place when the context's weights were refreshed"""
    from . import engine
    cache = ctx.__dict__.setdefault("_wt", {})
    key = (w.data_ptr(), tuple(w.shape), w.dtype)
    ent = cache.get(key)
    gen = (engine.WEIGHT_EPOCH[0], ctx.generation)
    if ent is None or ent[0] != gen:
        t = ent[1] if ent is not None else torch.empty(w.shape[1], w.shape[0], dtype=w.dtype, device=w.device)
        transpose2d(w, out=t)
        cache[key] = ent = (gen, t)
    return ent[1]


def linear_grads(ctx, dy, x, w, dw=None, db=None, need_dx=True, dx_accum=None, tag="lg", splitk=None):
    """Gradients of y = x @ w.T + b in the compute dtype of `w` (bf16 -> tcgen05 GEMM, fp32 -> SIMT GEMM, the parity mode):
    dy (M, N), x (M, K), w (N, K).  dw: an fp32 (N, K) tensor or a list of row blocks of it (the parameters behind a fused
    weight) and db likewise are ACCUMULATED into (autograd semantics; they may be views of the gradient arena).
    Returns dx (M, K) in the compute dtype, or None; dx_accum: an fp32 (M, K) buffer to add dx into instead.
    A token count that is not a multiple of 8 is handled by zero-padding the contraction dimension of the weight-gradient
    GEMM's transposed operands."""
    cd = w.dtype
    M, N = dy.shape
    K = x.shape[1]
    assert x.shape[0] == M and tuple(w.shape) == (N, K), (dy.shape, x.shape, w.shape)
    if dy.dtype != cd:
        dy = ops.convert(dy.contiguous(), ctx.buf(tag + "_dyc", (M, N), cd))
    if x.dtype != cd:
        x = ops.convert(x.contiguous(), ctx.buf(tag + "_xc", (M, K), cd))
    dx = None
    if need_dx:
        wt = weight_t(ctx, w)
        if dx_accum is not None:
            assert dx_accum.dtype == torch.float32 and tuple(dx_accum.shape) == (M, K) and dx_accum.is_contiguous()
            e = ops.epi_residual(ops.Rows(dx_accum.view(1, M, K), 0, M))
            if cd == torch.bfloat16 and N >= 16384:  # few output tiles, very long contraction (the packed adaLN Linear)
                e.k_split = min(64, N // 4096)
            ops.linear(dy, wt, e)
        else:
            dx = ctx.buf(tag + "_dx", (M, K), cd)
            ops.linear(dy, wt, ops.epi_store(dx))
    if dw is None and db is None:
        return dx
    # weight-gradient operands: the compute dtype of the layer — except for the small fp32 layers (patch embeds, heads) inside
    # a bf16 context with many tokens, whose dW = dY^T X contracts over ALL tokens with a handful of output tiles: there the
    # operands are rounded to bf16 and the contraction runs on the tensor cores (the reference's autocast runs these layers in
    # bf16 as well)
    wd = torch.bfloat16 if (cd == torch.float32 and ctx.dtype == torch.bfloat16 and M >= 2048 and dw is not None) else cd
    if wd != cd:
        dy_w = ops.convert(dy.contiguous(), ctx.buf(tag + "_dy16", (M, N), wd))
        x_w = ops.convert(x.contiguous(), ctx.buf(tag + "_x16", (M, K), wd))
    else:
        dy_w, x_w = dy, x
    dws = None if dw is None else (dw if isinstance(dw, (list, tuple)) else [dw])
    if wd == torch.bfloat16 and _TN and ops.tn_ok(dy_w, x_w) and (dws is None or all(t.shape[0] % 8 == 0 for t in dws)):
        # dW = dY^T X straight from dY and X as they lie in memory (MN-major tcgen05 operands): no transposed copies, any
        # token count (TMA zero-fills the tail of the contraction)
        if db is not None:
            dbs = list(db) if isinstance(db, (list, tuple)) else [db]
            if len(dbs) <= 3 and all(t.is_contiguous() and t.dtype == torch.float32 for t in dbs):
                colsum_into(dy, [t.view(-1) for t in dbs])   # += straight into the gradient slots
            else:
                cs, o = colsum(dy), 0
                for t in dbs:
                    ops.axpy(1.0, cs[o:o + t.numel()], t.view(-1))
                    o += t.numel()
        o = 0
        for t in dws or ():
            n = t.shape[0]
            t2 = t.view(n, K)
            assert t2.dtype == torch.float32 and t2.is_contiguous()
            e = ops.epi_residual(ops.Rows(t2.view(1, n, K), 0, n))
            tiles = -(-n // 256) * -(-K // 256)
            e.k_split = splitk if splitk is not None else max(1, min(16, 74 // tiles, M // 512))
            ops.linear_tn(dy_w[:, o:o + n], x_w, e)
            o += n
        return dx
    Mp = -(-M // 8) * 8 if wd == torch.bfloat16 else M
    dyT = ctx.buf(tag + "_dyT", (N, Mp), wd)
    if Mp != M:
        dyT[:, M:].zero_()
    fused = wd == torch.bfloat16 and not ((M | N | dy_w.stride(0)) & 7) and not (dy_w.data_ptr() & 15) and db is not None and wd == cd
    if fused:
        ws = ctx.buf(tag + "_csw", (-(-M // 64) * N,), torch.float32)
        cs = ctx.buf(tag + "_cs", (N,), torch.float32)
        _lib.check(_lib.load().hv_transpose2d_colsum(_p(dy_w), _p(dyT), M, N, dy_w.stride(0), Mp, _dt(dy_w), _p(ws), _p(cs), _stream()))
    else:
        transpose2d(dy_w, out=dyT)
        cs = colsum(dy) if db is not None else None
    if db is not None:
        dbs, o = (db if isinstance(db, (list, tuple)) else [db]), 0
        for t in dbs:
            ops.axpy(1.0, cs[o:o + t.numel()], t.view(-1))
            o += t.numel()
    if dw is not None:
        xT = ctx.buf(tag + "_xT", (K, Mp), wd)
        if Mp != M:
            xT[:, M:].zero_()
        transpose2d(x_w, out=xT)
        o = 0
        for t in dws:
            n = t.shape[0]
            t2 = t.view(n, K)
            assert t2.dtype == torch.float32 and t2.is_contiguous()
            e = ops.epi_residual(ops.Rows(t2.view(1, n, K), 0, n))
            if wd == torch.bfloat16:
                tiles = -(-n // 256) * -(-K // 256)
                e.k_split = splitk if splitk is not None else max(1, min(16, 74 // tiles, Mp // 512))
            ops.linear(dyT[o:o + n], xT, e)
            o += n
    return dx


# ------------------------------------------------------------------------------------------------
# Whole-block backward: BasicTransformerBlock (modules.py:721-743, the object motion encoder's block, a10 / a14) and
# BasicCrossTransformerBlock (modules.py:817-840, the camera motion encoder's block, a15).
# bf16 compute mode = the reference's `--mp bf16` autocast semantics (train_amd.py:160-165): bf16 GEMM / attention
# operands, fp32 accumulation, fp32 residual stream, fp32 LayerNorm statistics, fp32 parameter gradients.

def block_forward_train(ctx, bp, x, enc=None):
    """x (B, L, D) [, enc (B, Lk, cross_dim) for the cross block] -> (out (B, L, D) fp32, saved activations).
    Same kernels as engine.run_block; the q/k/v and the first feed-forward GEMM run a second time with the plain store
    epilogue because the backward needs the pre-norm q / k and the pre-GELU activations (first version: a dual-output
    epilogue removes the extra GEMMs)."""
    from . import engine
    assert bp.kind in ("basic", "cross") and (enc is not None) == (bp.kind == "cross")
    assert ctx.dtype == torch.bfloat16, "the block backward is built for the bf16 compute mode"
    B, L, D = x.shape
    M, bf, dev = B * L, torch.bfloat16, x.device
    new = lambda *shape: torch.empty(*shape, dtype=bf, device=dev)
    x0 = x.detach().float().contiguous()
    res = x0.clone()
    rows = ops.Rows(res, 0, L)
    scale = 1.0 / (D // bp.heads) ** 0.5
    xn1 = ops.ln_modulate(rows, bp.n1w, bp.n1b, bp.eps, new(M, D))
    s = dict(x0=x0, xn1=xn1, scale=scale)
    if bp.kind == "cross":
        Lk = enc.shape[1]
        geom = engine.AttnGeom(B, L, Lk, ops.seq_map(1, L, 0, 1), ops.seq_map(1, Lk, 0, 1))
        enc32 = enc.detach().reshape(B * Lk, -1).float().contiguous()
        enc16 = ops.convert(enc32, new(*enc32.shape))
        q_pre, q, kv_pre, kv = new(M, D), new(M, D), new(B * Lk, 2 * D), new(B * Lk, 2 * D)
        ops.linear(xn1, bp.wq, ops.epi_store(q_pre, bp.bq))
        ops.linear(xn1, bp.wq, ops.epi_qknorm(q, bp.bq, bp.qn_w, bp.qn_b, None, None, D, 0))
        ops.linear(enc16, bp.wkv, ops.epi_store(kv_pre, bp.bkv))
        ops.linear(enc16, bp.wkv, ops.epi_qknorm(kv, bp.bkv, None, None, bp.kn_w, bp.kn_b, 0, D))
        s.update(enc16=enc16, q=q, k=kv[:, :D], v=kv[:, D:], q_pre=q_pre, k_pre=kv_pre[:, :D])
    else:
        geom = engine.dense_geom(B, L)
        qkv_pre, qkv = new(M, 3 * D), new(M, 3 * D)
        ops.linear(xn1, bp.wqkv, ops.epi_store(qkv_pre, bp.bqkv))
        ops.linear(xn1, bp.wqkv, ops.epi_qknorm(qkv, bp.bqkv, bp.qn_w, bp.qn_b, bp.kn_w, bp.kn_b, D, D))
        s.update(q=qkv[:, :D], k=qkv[:, D:2 * D], v=qkv[:, 2 * D:], q_pre=qkv_pre[:, :D], k_pre=qkv_pre[:, D:2 * D])
    att = new(M, D)
    ops.attention(s["q"], s["k"], s["v"], att, geom.seqs, bp.heads, D // bp.heads, geom.Lq, geom.Lkv, geom.qmap, geom.kvmap, scale)
    ops.linear(att, bp.wo, ops.epi_residual(rows, bp.bo))
    x1 = res.clone()
    xn2 = ops.ln_modulate(rows, bp.n2w, bp.n2b, bp.eps, new(M, D))
    F = bp.w1.shape[0]
    hpre, h = new(M, F), new(M, F)
    ops.linear(xn2, bp.w1, ops.epi_store(hpre, bp.b1))
    ops.linear(xn2, bp.w1, ops.epi_store(h, bp.b1, gelu=True))
    ops.linear(h, bp.w2, ops.epi_residual(rows, bp.b2))
    s.update(att=att, x1=x1, xn2=xn2, hpre=hpre, h=h, geom=geom)
    return res, s


def block_backward(bp, saved, dy):
    """dy (B, L, D) = d loss / d out  ->  (dx (B, L, D) fp32, d enc (B*Lk, cross_dim) fp32 or None,
    {reference parameter name: fp32 gradient}).
    Chain: W2 -> GELU -> W1 -> LayerNorm2 (+ residual) -> Wo -> attention -> QK-norm -> Wq/Wk/Wv -> LayerNorm1 (+ residual)."""
    s = saved
    B, L, D = dy.shape
    M, bf, dev = B * L, torch.bfloat16, dy.device
    g = {}
    d2 = dy.detach().reshape(M, D).float().contiguous()
    dx = d2.clone()                                                 # running residual-stream gradient
    d16 = ops.convert(d2, torch.empty(M, D, dtype=bf, device=dev))
    dh, g["ff.net.2.weight"], g["ff.net.2.bias"] = linear_backward(d16, s["h"], bp.w2)
    dhpre = gelu_tanh_backward(s["hpre"], dh)
    dxn2, g["ff.net.0.proj.weight"], g["ff.net.0.proj.bias"] = linear_backward(dhpre, s["xn2"], bp.w1)
    _, g["norm2.weight"], g["norm2.bias"] = layernorm_backward(s["x1"].view(M, D), dxn2, bp.n2w, bp.eps, dx=dx, accumulate=True)
    ops.convert(dx, d16)
    datt, g["attn1.to_out.0.weight"], g["attn1.to_out.0.bias"] = linear_backward(d16, s["att"], bp.wo)
    geom = s["geom"]
    # d(q_pre) | d(k_pre) | dv are written straight into the column slices of the gradient of the fused projection output
    cross = bp.kind == "cross"
    Mk = s["k"].shape[0]
    dproj = torch.empty(Mk, 2 * D if cross else 3 * D, dtype=bf, device=dev)
    dqp = torch.empty(M, D, dtype=bf, device=dev) if cross else dproj[:, :D]
    dkp, dvv = (dproj[:, :D], dproj[:, D:]) if cross else (dproj[:, D:2 * D], dproj[:, 2 * D:])
    dq, dk, _ = attention_backward(s["q"], s["k"], s["v"], s["att"], datt, geom.seqs, bp.heads, geom.Lq, geom.Lkv,
                                   geom.qmap, geom.kvmap, s["scale"], dv=dvv)
    _, g["attn1.norm_q.weight"], g["attn1.norm_q.bias"] = headnorm64_backward(s["q_pre"], dq, bp.qn_w, bp.heads, out=dqp)
    _, g["attn1.norm_k.weight"], g["attn1.norm_k.bias"] = headnorm64_backward(s["k_pre"], dk, bp.kn_w, bp.heads, out=dkp)
    denc = None
    if cross:
        dxn1, g["attn1.to_q.weight"], g["attn1.to_q.bias"] = linear_backward(dqp, s["xn1"], bp.wq)
        denc, dwkv, dbkv = linear_backward(dproj, s["enc16"], bp.wkv, dx_dtype=torch.float32)
        for i, n in enumerate(("to_k", "to_v")):
            g["attn1.%s.weight" % n], g["attn1.%s.bias" % n] = dwkv[i * D:(i + 1) * D], dbkv[i * D:(i + 1) * D]
    else:
        dxn1, dwqkv, dbqkv = linear_backward(dproj, s["xn1"], bp.wqkv)
        for i, n in enumerate(("to_q", "to_k", "to_v")):
            g["attn1.%s.weight" % n], g["attn1.%s.bias" % n] = dwqkv[i * D:(i + 1) * D], dbqkv[i * D:(i + 1) * D]
    _, g["norm1.weight"], g["norm1.bias"] = layernorm_backward(s["x0"].view(M, D), dxn1, bp.n1w, bp.eps, dx=dx, accumulate=True)
    return dx.view(B, L, D), denc, g


class BlockFn(torch.autograd.Function):
    """autograd wrapper: forward(blk, x, enc | None, *blk.parameters()) -> out; backward returns dx, d enc and
    every parameter gradient from the sm_100a kernels.  `blk` is a hivae_b200.modules.BasicTransformerBlock or
    BasicCrossTransformerBlock."""

    @staticmethod
    def forward(fctx, blk, x, enc, *params):
        ctx, bp, _ = blk._prepared(x)
        out, saved = block_forward_train(ctx, bp, x, enc)
        fctx.bp, fctx.saved, fctx.names = bp, saved, [n for n, _ in blk.named_parameters()]
        fctx.x_dtype, fctx.enc_meta = x.dtype, (None if enc is None else (enc.shape, enc.dtype))
        return out.to(x.dtype)

    @staticmethod
    def backward(fctx, dy):
        dx, denc, g = block_backward(fctx.bp, fctx.saved, dy)
        fctx.saved = None
        if denc is not None:
            denc = denc.view(fctx.enc_meta[0]).to(fctx.enc_meta[1])
        return (None, dx.to(fctx.x_dtype), denc) + tuple(g[n] for n in fctx.names)


def basic_block_apply(blk, x):
    """differentiable BasicTransformerBlock.forward: gradients flow to x and to blk's parameters"""
    return BlockFn.apply(blk, x, None, *blk.parameters())


def cross_block_apply(blk, x, encoder_hidden_state):
    """differentiable BasicCrossTransformerBlock.forward: gradients flow to x, encoder_hidden_state and the parameters"""
    return BlockFn.apply(blk, x, encoder_hidden_state, *blk.parameters())
