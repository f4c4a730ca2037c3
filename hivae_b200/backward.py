"""Building blocks of the backward pass of the blocks (`accelerator.backward(loss)`, train_amd.py:489)
on the sm_100a kernels of csrc/backward.cu and the tcgen05 GEMM: gradients of nn.Linear, nn.LayerNorm and the
tanh-GELU of FeedForward.  These are the operator-level pieces; the autograd wiring of whole blocks (attention
backward, adaLN / gated-residual backward, torch.autograd.Function wrappers) is the next round's work.

No CPU fallback: every function raises without the CUDA library."""
import ctypes as C

import torch

from . import _lib, ops
from .ops import _chk_cuda, _dt, _p, _stream


def transpose2d(x, out=None):
    """(R, C) -> (C, R), bf16 or fp32; x may be a row-strided view with unit column stride"""
    _lib.require_device()
    assert x.dim() == 2 and x.is_cuda and (x.shape[1] == 1 or x.stride(1) == 1)
    R, Cc = x.shape
    out = torch.empty(Cc, R, dtype=x.dtype, device=x.device) if out is None else out
    _lib.check(_lib.load().hv_transpose2d(_p(x), _p(out), R, Cc, x.stride(0), out.stride(0), _dt(x), _stream()))
    return out


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
    dw = torch.empty(N, K, dtype=torch.float32, device=dy.device)
    ops.linear(transpose2d(dy), transpose2d(x), ops.epi_store(dw))     # (N, M) x (K, M)^T
    return dx, dw, colsum(dy)


def gelu_tanh_backward(pre, dy, out_dtype=None):
    """dx = dy * gelu_tanh'(pre); pre / dy bf16 or fp32 (same dtype), dx in out_dtype (default: same)"""
    _chk_cuda(pre, dy)
    assert pre.shape == dy.shape and pre.dtype == dy.dtype
    dx = torch.empty(pre.shape, dtype=out_dtype or pre.dtype, device=pre.device)
    _lib.check(_lib.load().hv_gelu_tanh_bwd(_p(pre), _p(dy), _p(dx), pre.numel(), _dt(pre), _dt(dx), _stream()))
    return dx


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


def attention_backward(q, k, v, out, dout, seqs, heads, Lq, Lkv, qmap, kvmap, scale):
    """Gradients of ops.attention (head_dim 64, no mask): q / k / v / out / dout are 2-D (rows, heads*64) views in one
    dtype (bf16 or fp32) with unit column stride -> (dq, dk, dv) of the same dtype, dense (rows, heads*64)."""
    _lib.require_device()
    for t in (q, k, v, out, dout):
        assert t.is_cuda and t.dim() == 2 and t.stride(1) == 1 and t.dtype == q.dtype
    dq = torch.zeros(q.shape[0], heads * 64, dtype=q.dtype, device=q.device)
    dk = torch.zeros(k.shape[0], heads * 64, dtype=q.dtype, device=q.device)
    dv = torch.zeros(v.shape[0], heads * 64, dtype=q.dtype, device=q.device)
    stats = torch.empty(seqs * heads * Lq * 2, dtype=torch.float32, device=q.device)
    _lib.check(_lib.load().hv_attention_bwd(
        _p(q), q.stride(0), _p(k), k.stride(0), _p(v), v.stride(0), _p(out), out.stride(0), _p(dout), dout.stride(0),
        _p(dq), dq.stride(0), _p(dk), dk.stride(0), _p(dv), dv.stride(0), _dt(q), seqs, heads, 64, Lq, Lkv,
        C.byref(qmap), C.byref(kvmap), float(scale), _p(stats), _stream()))
    return dq, dk, dv


def headnorm64_backward(x, dy, weight, heads, eps=1e-6):
    """backward of the per-head LayerNorm(64) of q / k (QK-norm): x (pre-norm), dy are (tokens, heads*64) views (bf16 or
    fp32, unit column stride) -> (dx dense, dw (64,), db (64,))"""
    _lib.require_device()
    assert x.is_cuda and x.dim() == 2 and x.shape == dy.shape and x.shape[1] == heads * 64 and x.dtype == dy.dtype
    assert x.stride(1) == 1 and dy.stride(1) == 1
    lib = _lib.load()
    dx = torch.empty(x.shape, dtype=x.dtype, device=x.device)
    ws = torch.empty((lib.hv_layernorm_bwd_ctas() + 1) * 128, dtype=torch.float32, device=x.device)
    dw, db = (torch.empty(64, dtype=torch.float32, device=x.device) for _ in range(2))
    _lib.check(lib.hv_headnorm64_bwd(_p(x), x.stride(0), _p(dy), dy.stride(0), _p(weight), float(eps), x.shape[0], heads,
                                     _p(dx), dx.stride(0), _dt(x), _p(ws), _p(dw), _p(db), _stream()))
    return dx, dw, db
