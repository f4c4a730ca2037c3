"""Torch-tensor facing wrappers over the C ABI (pointer + size marshalling only).

PyTorch is used for device memory and streams; every computation below is one of the
hand-written sm_100a kernels behind include/hivae_b200.h.
"""
import ctypes as C
from dataclasses import dataclass

import torch

from . import _lib
from ._lib import HV_BF16, HV_F32, EPI_GELU, EPI_QKNORM, EPI_RESIDUAL, EPI_STORE, HvEpilogue, HvSeqMap  # noqa: F401


def _dt(t):
    if t.dtype == torch.float32:
        return HV_F32
    if t.dtype == torch.bfloat16:
        return HV_BF16
    raise TypeError("hivae_b200 supports float32 / bfloat16 tensors, got %s" % t.dtype)


def _dtc(dtype):
    return HV_F32 if dtype == torch.float32 else HV_BF16


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _chk_cuda(*ts):
    _lib.require_device()
    for t in ts:
        if t is not None:
            if not t.is_cuda:
                raise RuntimeError("hivae_b200 ops need CUDA tensors (no CPU fallback)")
            if not t.is_contiguous():
                raise RuntimeError("hivae_b200 ops need contiguous tensors")


def launch_count():
    return int(_lib.load().hv_launch_count())


# ------------------------------------------------------------------ per-launch timing (bench.py roofline leg)
PROFILE = None  # set to a list to record (kernel, flops, bytes, start_event, end_event) per launch


def _prof(name, flops=0.0, nbytes=0.0):
    if PROFILE is None:
        return None
    e0 = torch.cuda.Event(enable_timing=True)
    e0.record()
    return (name, float(flops), float(nbytes), e0)


def _prof_end(tok):
    if tok is not None:
        e1 = torch.cuda.Event(enable_timing=True)
        e1.record()
        PROFILE.append(tok + (e1,))


def profile_summary(records):
    """{kernel: {launches, ms, flops, bytes}} from a PROFILE list (call after a device synchronize)."""
    out = {}
    for name, fl, nb, e0, e1 in records:
        d = out.setdefault(name, {"launches": 0, "ms": 0.0, "flops": 0.0, "bytes": 0.0})
        d["launches"] += 1
        d["ms"] += e0.elapsed_time(e1)
        d["flops"] += fl
        d["bytes"] += nb
    return out


# ------------------------------------------------------------------ wavelet / frequency
def dwt2_haar(x):
    """(B,C,H,W) -> LL, HL, LH, HH each (B,C,H/2,W/2), dtype preserved (wavelet.py:13-27)."""
    _chk_cuda(x)
    B, Cc, H, W = x.shape
    outs = [torch.empty(B, Cc, H // 2, W // 2, dtype=x.dtype, device=x.device) for _ in range(4)]
    tok = _prof("dwt2_haar", 0, 2 * x.numel() * x.element_size())
    _lib.check(_lib.load().hv_dwt2_haar(_p(x), *[_p(o) for o in outs], B * Cc, H, W, _dt(x), _stream()))
    _prof_end(tok)
    return tuple(outs)


def idwt2_haar(ll, hl, lh, hh):
    """four (B,C,h,w) sub-bands -> (B,C,2h,2w) fp32 (wavelet.py:31-48)."""
    _chk_cuda(ll, hl, lh, hh)
    if not (hl.dtype == lh.dtype == hh.dtype) or not (ll.shape == hl.shape == lh.shape == hh.shape):
        raise ValueError("idwt2_haar: the detail bands must share one dtype and all four bands one shape")
    B, Cc, h, w = ll.shape
    out = torch.empty(B, Cc, 2 * h, 2 * w, dtype=torch.float32, device=ll.device)
    code = _dt(ll) if ll.dtype == hl.dtype else (_dt(ll) | ((_dt(hl) + 1) << 8))  # mixed: torch.cat promotion
    tok = _prof("idwt2_haar", 0, ll.numel() * (ll.element_size() + 3 * hl.element_size()) + out.numel() * 4)
    _lib.check(_lib.load().hv_idwt2_haar(_p(ll), _p(hl), _p(lh), _p(hh), _p(out), B * Cc, h, w, code, _stream()))
    _prof_end(tok)
    return out


def _ptr_array(ts):
    return (C.c_void_p * len(ts))(*[t.data_ptr() for t in ts])


def pyramid_fusable(shape, levels):
    """the fused pyramid kernels need 1..3 levels, H divisible by 2^levels and W by 8"""
    H, W = shape[-2], shape[-1]
    return 1 <= levels <= 3 and H % (1 << levels) == 0 and W % 8 == 0 and H > 0 and W > 0


def dwt2_haar_pyramid(x, levels):
    """fused `levels`-level analysis in one launch -> ([(HL, LH, HH) per level], LL of the last level)"""
    _chk_cuda(x)
    B, Cc, H, W = x.shape
    mk = lambda k: torch.empty(B, Cc, H >> (k + 1), W >> (k + 1), dtype=x.dtype, device=x.device)
    hl, lh, hh = [mk(k) for k in range(levels)], [mk(k) for k in range(levels)], [mk(k) for k in range(levels)]
    ll = mk(levels - 1)
    # algorithmic bytes of the per-level formulation: read + write N * sum_k 4^-k
    tok = _prof("dwt2_haar_pyramid", 0, 2 * x.numel() * x.element_size() * sum(4.0 ** -k for k in range(levels)))
    _lib.check(_lib.load().hv_dwt2_haar_pyramid(_p(x), levels, _ptr_array(hl), _ptr_array(lh), _ptr_array(hh), _p(ll),
                                                B * Cc, H, W, _dt(x), _stream()))
    _prof_end(tok)
    return list(zip(hl, lh, hh)), ll


def idwt2_haar_pyramid(bands, ll):
    """fused synthesis of a same-dtype pyramid (bands[k] = (HL, LH, HH) of level k, finest first) -> fp32 image"""
    flat = [b for lv in bands for b in lv]
    _chk_cuda(ll, *flat)
    levels = len(bands)
    B, Cc, h, w = ll.shape
    H, W = h << levels, w << levels
    out = torch.empty(B, Cc, H, W, dtype=torch.float32, device=ll.device)
    tok = _prof("idwt2_haar_pyramid", 0, out.numel() * (ll.element_size() + 4) * sum(4.0 ** -k for k in range(levels)))
    _lib.check(_lib.load().hv_idwt2_haar_pyramid(_p(ll), levels, _ptr_array([b[0] for b in bands]), _ptr_array([b[1] for b in bands]),
                                                 _ptr_array([b[2] for b in bands]), _p(out), B * Cc, H, W, _dt(ll), _stream()))
    _prof_end(tok)
    return out


def freq_split3d(x, ct, ch, cw):
    """x (..., T, H, W) fp32 -> (low, high) via the three circulant matrices."""
    _chk_cuda(x, ct, ch, cw)
    T, H, W = x.shape[-3:]
    vols = x.numel() // (T * H * W) if x.numel() else 0
    lo, hi = torch.empty_like(x), torch.empty_like(x)
    tok = _prof("freq_split3d", 0, 3 * x.numel() * 4)
    _lib.check(_lib.load().hv_freq_split3d(_p(x), _p(ct), _p(ch), _p(cw), _p(lo), _p(hi), vols, T, H, W, _stream()))
    _prof_end(tok)
    return lo, hi


# ------------------------------------------------------------------ row addressing
@dataclass
class Rows:
    """Row range [row_off, row_off+L) of every frame of an fp32 residual buffer
    res: (frames, frame_rows, D).  Rows in [img_start, img_start+img_len) are image rows."""
    res: torch.Tensor
    row_off: int
    L: int
    img_start: int = 0
    img_len: int = 0

    @property
    def frames(self):
        return self.res.shape[0]

    @property
    def frame_rows(self):
        return self.res.shape[1]

    @property
    def D(self):
        return self.res.shape[2]

    @property
    def M(self):
        return self.frames * self.L


@dataclass
class Ada:
    """adaLN table (rows, ld) fp32 = Linear(SiLU(temb)); the conditioning row of frame f is
    (f // cond_div) * cond_mul."""
    table: torch.Tensor
    cond_div: int = 1
    cond_mul: int = 1


def ln_modulate(rows, ln_w, ln_b, eps, out, ada=None, shift_mot=0, scale_mot=0, shift_img=0, scale_img=0, save=None):
    """save: optional fp32 (M, D) buffer that receives the rows entering the norm (training forward)"""
    _chk_cuda(rows.res, ln_w, ln_b, out, save)
    t = ada.table if ada is not None else None
    tok = _prof("ln_modulate", 0, rows.M * rows.D * (4 + out.element_size() + (4 if save is not None else 0)))
    args = (_p(rows.res), rows.frames, rows.frame_rows, rows.row_off, rows.L, rows.D, _p(ln_w), _p(ln_b), float(eps),
            _p(t), t.shape[1] if t is not None else 0, shift_mot, scale_mot, shift_img, scale_img,
            rows.img_start, rows.img_len, ada.cond_div if ada else 1, ada.cond_mul if ada else 1, _p(out), _dt(out))
    if save is None:
        _lib.check(_lib.load().hv_ln_modulate(*args, _stream()))
    else:
        assert save.dtype == torch.float32 and tuple(save.shape) == (rows.M, rows.D)
        _lib.check(_lib.load().hv_ln_modulate_save(*args, _p(save), _stream()))
    _prof_end(tok)
    return out


def patch_embed(in1, in2, p, w, bias, pos, tpos, out, frame_rows, row_off, D):
    _chk_cuda(in1, in2, w, bias, pos, tpos, out)
    frames, C1, H, W = in1.shape
    C2 = in2.shape[1] if in2 is not None else 0
    _lib.check(_lib.load().hv_patch_embed(_p(in1), C1, _p(in2), C2, frames, H, W, p, _p(w), _p(bias), _p(pos), _p(tpos),
                                          tpos.shape[0] if tpos is not None else 1, _p(out), _dt(out), frame_rows, row_off, D,
                                          _stream()))
    return out


def final_head(rows, n1w, n1b, n2w, n2b, eps, ada, shift_col, scale_col, w, bias, C_out, p, Wp, pre, zt, dt):
    _chk_cuda(rows.res, w, bias, pre, zt)
    t = ada.table if ada is not None else None
    _lib.check(_lib.load().hv_final_head(_p(rows.res), rows.frames, rows.frame_rows, rows.row_off, rows.L, rows.D,
                                         _p(n1w), _p(n1b), _p(n2w), _p(n2b), float(eps), _p(t),
                                         t.shape[1] if t is not None else 0, shift_col, scale_col, _p(w), _p(bias),
                                         C_out, p, Wp, _p(pre), _p(zt), float(dt), _stream()))


def timestep_embed(t, dim, out):
    _chk_cuda(t, out)
    _lib.check(_lib.load().hv_timestep_embed(_p(t), t.numel(), dim, _p(out), _dt(out), _stream()))
    return out


def silu(x, out):
    _chk_cuda(x, out)
    _lib.check(_lib.load().hv_silu(_p(x), _dt(x), _p(out), _dt(out), x.numel(), _stream()))
    return out


def convert(x, out):
    _chk_cuda(x, out)
    tok = _prof("convert", 0, x.numel() * (x.element_size() + out.element_size()))
    _lib.check(_lib.load().hv_convert(_p(x), _dt(x), _p(out), _dt(out), x.numel(), _stream()))
    _prof_end(tok)
    return out


def axpy(a, x, y):
    _chk_cuda(x, y)
    _lib.check(_lib.load().hv_axpy(float(a), _p(x), _p(y), x.numel(), _stream()))
    return y


def rows_scatter(src, stride_f, stride_i, add, rows):
    _chk_cuda(src, add, rows.res)
    _lib.check(_lib.load().hv_rows_scatter(_p(src), stride_f, stride_i, _p(add), _p(rows.res), rows.frames, rows.frame_rows,
                                           rows.row_off, rows.L, rows.D, _stream()))


def rows_gather(rows, out):
    _chk_cuda(rows.res, out)
    tok = _prof("rows_gather", 0, rows.M * rows.D * 8)
    _lib.check(_lib.load().hv_rows_gather(_p(rows.res), rows.frames, rows.frame_rows, rows.row_off, rows.L, rows.D, _p(out),
                                          _stream()))
    _prof_end(tok)
    return out


def unpatchify(x, p, c=None, order=0):
    """tokens (N, S, p*p*c) -> (N, c, H, W); order 0 = (p,q,c) features (model_AMD.py:2816-2830), 1 = (c,p,q)"""
    _chk_cuda(x)
    N, S, F = x.shape
    h = w = int(S ** 0.5)
    assert h * w == S
    c = F // (p * p) if c is None else c
    x = x.float().contiguous()
    out = torch.empty(N, c, h * p, w * p, dtype=torch.float32, device=x.device)
    _lib.check(_lib.load().hv_unpatchify(_p(x), _p(out), N, c, h, w, p, order, 0, _stream()))
    return out


def patchify(img, p, order=0):
    """(N, c, H, W) -> tokens (N, (H/p)*(W/p), p*p*c) (model_MAE.py:116-129)"""
    _chk_cuda(img)
    N, c, H, W = img.shape
    assert H % p == 0 and W % p == 0
    img = img.float().contiguous()
    out = torch.empty(N, (H // p) * (W // p), p * p * c, dtype=torch.float32, device=img.device)
    _lib.check(_lib.load().hv_unpatchify(_p(img), _p(out), N, c, H // p, W // p, p, order, 1, _stream()))
    return out


def conv3x3(x, w, bias, stride):
    _chk_cuda(x, w, bias)
    B, Cin, H, W = x.shape
    Cout = w.shape[0]
    Ho, Wo = (H + 2 - 3) // stride + 1, (W + 2 - 3) // stride + 1
    y = torch.empty(B, Cout, Ho, Wo, dtype=torch.float32, device=x.device)
    _lib.check(_lib.load().hv_conv3x3(_p(x), _p(w), _p(bias), _p(y), B, Cin, Cout, H, W, stride, _stream()))
    return y


def maxpool2(x):
    _chk_cuda(x)
    B, Cc, H, W = x.shape
    y = torch.empty(B, Cc, H // 2, W // 2, dtype=torch.float32, device=x.device)
    _lib.check(_lib.load().hv_maxpool2(_p(x), _p(y), B * Cc, H, W, _stream()))
    return y


def rope_(q, k, batch, L, heads, head_dim, start_pos=0, base=10000.0):
    _chk_cuda(q, k)
    _lib.check(_lib.load().hv_rope(_p(q), _p(k), batch, L, heads, head_dim, heads * head_dim, start_pos, float(base), _stream()))


# ------------------------------------------------------------------ GEMM
def epi_store(out, bias=None, gelu=False, pre=None):
    """pre: optional second output (GELU only) that receives acc + bias BEFORE the activation (training forward)"""
    e = HvEpilogue()
    e.mode = EPI_GELU if gelu else EPI_STORE
    e.out_dtype, e.out, e.ldo, e.bias = _dt(out), out.data_ptr(), out.stride(0), (bias.data_ptr() if bias is not None else None)
    if pre is not None:
        assert pre.dtype == out.dtype and pre.shape == out.shape and pre.stride(1) == 1
        e.out2, e.ldo2 = pre.data_ptr(), pre.stride(0)
    return e


def epi_qknorm(out, bias, qn_w, qn_b, kn_w, kn_b, q_cols, k_cols, eps=1e-6, pre=None):
    """pre: optional second output receiving the projections BEFORE the per-head LayerNorm (training forward)"""
    e = epi_store(out, bias, pre=pre)
    e.mode = EPI_QKNORM
    e.qn_w = qn_w.data_ptr() if qn_w is not None else None
    e.qn_b = qn_b.data_ptr() if qn_b is not None else None
    e.kn_w = kn_w.data_ptr() if kn_w is not None else None
    e.kn_b = kn_b.data_ptr() if kn_b is not None else None
    e.q_cols, e.k_cols, e.qk_eps = q_cols, k_cols, eps
    return e


def epi_residual(rows, bias=None, ada=None, gate_mot=0, gate_img=0):
    e = HvEpilogue()
    e.mode = EPI_RESIDUAL
    e.bias = bias.data_ptr() if bias is not None else None
    e.res, e.L, e.frame_rows, e.row_off = rows.res.data_ptr(), rows.L, rows.frame_rows, rows.row_off
    e.img_start, e.img_len = rows.img_start, rows.img_len
    if ada is not None:
        e.gate, e.gate_ld = ada.table.data_ptr(), ada.table.shape[1]
        e.gate_col_mot, e.gate_col_img, e.cond_div, e.cond_mul = gate_mot, gate_img, ada.cond_div, ada.cond_mul
    else:
        e.cond_div = e.cond_mul = 1
    return e


def linear(A, W, epi):
    """epilogue(A[M,K] @ W[N,K]^T).  bf16 x bf16 -> tcgen05 GEMM; anything with fp32 W -> fp32 SIMT GEMM."""
    _lib.require_device()
    for t in (A, W):  # row-strided 2-D views (column slices of a wider matrix) are fine: ld is passed explicitly
        if not t.is_cuda or t.dim() != 2 or (t.shape[1] > 1 and t.stride(1) != 1):
            raise RuntimeError("hivae_b200.linear needs 2-D CUDA tensors with unit column stride (no CPU fallback)")
    M, K = A.shape
    N = W.shape[0]
    assert W.shape[1] == K, (A.shape, W.shape)
    lib = _lib.load()
    if W.dtype == torch.bfloat16:
        if A.dtype != torch.bfloat16:
            raise TypeError("bf16 weights need bf16 activations")
        tok = _prof("gemm_tc", 2.0 * M * N * K, 2.0 * (M * K + N * K + M * N))
        _lib.check(lib.hv_linear_bf16(_p(A), A.stride(0), _p(W), W.stride(0), M, N, K, C.byref(epi), _stream()))
    else:
        tok = _prof("gemm_simt", 2.0 * M * N * K, 4.0 * (M * K + N * K + M * N))
        _lib.check(lib.hv_linear_f32(_p(A), _dt(A), A.stride(0), _p(W), W.stride(0), M, N, K, C.byref(epi), _stream()))
    _prof_end(tok)


def linear_tn(At, Bt, epi):
    """epilogue(At^T @ Bt) for row-major bf16 At (K, M), Bt (K, N): the weight gradient dY^T X without transposed copies"""
    _lib.require_device()
    for t in (At, Bt):
        if not t.is_cuda or t.dim() != 2 or t.stride(1) != 1 or t.dtype != torch.bfloat16:
            raise RuntimeError("hivae_b200.linear_tn needs 2-D bf16 CUDA tensors with unit column stride")
    K, M = At.shape
    N = Bt.shape[1]
    assert Bt.shape[0] == K
    tok = _prof("gemm_tc", 2.0 * M * N * K, 2.0 * K * (M + N) + 4.0 * M * N)
    _lib.check(_lib.load().hv_linear_bf16_tn(_p(At), At.stride(0), _p(Bt), Bt.stride(0), M, N, K, C.byref(epi), _stream()))
    _prof_end(tok)


def tn_ok(At, Bt):
    return (At.dtype == Bt.dtype == torch.bfloat16 and At.stride(1) == 1 and Bt.stride(1) == 1 and At.shape[1] % 8 == 0
            and Bt.shape[1] % 8 == 0 and At.stride(0) % 8 == 0 and Bt.stride(0) % 8 == 0 and At.data_ptr() % 16 == 0
            and Bt.data_ptr() % 16 == 0)


# ------------------------------------------------------------------ attention
def seq_map(sdiv, sa, sb, sc):
    m = HvSeqMap()
    m.sdiv, m.sa, m.sb, m.sc = sdiv, sa, sb, sc
    return m


def attention(q, k, v, out, seqs, heads, hd, Lq, Lkv, qmap, kvmap, scale, mask=None, lse=None):
    """q/k/v/out are 2-D (rows, cols) views whose column 0 is head 0; row strides are taken from the views.
    lse: optional fp32 workspace of seqs * heads * Lq * 2 floats (the `stats` of backward.attention_backward); the call then
    returns True when the kernel that ran left the per-row log-sum-exp there (hv_attention_lse), which lets the backward skip
    its own maximum / sum passes, else False (the output is complete either way).  Without `lse` it returns `out`."""
    _lib.require_device()
    tok = _prof("attention_L%d" % Lkv, 4.0 * seqs * heads * Lq * Lkv * hd, q.element_size() * heads * hd * seqs * (2.0 * Lq + 2.0 * Lkv))
    if lse is not None:
        assert lse.is_cuda and lse.dtype == torch.float32 and lse.is_contiguous() and lse.numel() >= seqs * heads * Lq * 2
        rc = _lib.load().hv_attention_lse(_p(q), q.stride(0), _p(k), k.stride(0), _p(v), v.stride(0), _p(out), out.stride(0),
                                          _dt(q), seqs, heads, hd, Lq, Lkv, C.byref(qmap), C.byref(kvmap), float(scale),
                                          _p(mask), _p(lse), _stream())
        if rc < 0:
            _lib.check(rc)
        _prof_end(tok)
        return rc == 1
    _lib.check(_lib.load().hv_attention(_p(q), q.stride(0), _p(k), k.stride(0), _p(v), v.stride(0), _p(out), out.stride(0),
                                        _dt(q), seqs, heads, hd, Lq, Lkv, C.byref(qmap), C.byref(kvmap), float(scale),
                                        _p(mask), _stream()))
    _prof_end(tok)
    return out
