"""Haar DWT / IDWT with the reference's API (model/wavelet.py) on the sm_100a kernels."""
import torch
from torch import nn

from . import ops


def dwt_init(x):
    """(B,C,H,W) -> (x_LL, x_HL, x_LH, x_HH), dtype preserved (wavelet.py:13-27)."""
    return ops.dwt2_haar(x.contiguous())


def iwt_init(x):
    """x = cat(LL,HL,LH,HH) on dim 0 -> (B,C,2h,2w), always fp32 (wavelet.py:31-48)."""
    x = x.contiguous()
    B = x.shape[0] // 4
    return ops.idwt2_haar(x[0:B], x[B:2 * B], x[2 * B:3 * B], x[3 * B:4 * B])


class DWT(nn.Module):
    def __init__(self):
        super().__init__()
        self.requires_grad = False

    def forward(self, x):
        return dwt_init(x)


class IWT(nn.Module):
    def __init__(self):
        super().__init__()
        self.requires_grad = False

    def forward(self, x):
        return iwt_init(x)


def dwt_pyramid(x, levels):
    """Multi-level analysis (the caller-side recursion on LL of frenquence_wavelet.py:86-89).
    Returns ([(HL, LH, HH) per level], LL of the last level).  One fused launch when the shape allows
    (<= 3 levels, H % 2^levels == 0, W % 8 == 0), bit-identical to the per-level recursion."""
    x = x.contiguous()
    if x.dim() == 4 and ops.pyramid_fusable(x.shape, levels) and x.data_ptr() % 16 == 0:
        return ops.dwt2_haar_pyramid(x, levels)
    bands, ll = [], x
    for _ in range(levels):
        ll, hl, lh, hh = dwt_init(ll)
        bands.append((hl, lh, hh))
    return bands, ll


def iwt_pyramid(bands, ll):
    """Multi-level synthesis.  Every level returns fp32 (wavelet.py:40-41), so from the second level on a
    bf16 pyramid mixes an fp32 LL with bf16 detail bands — promoted to fp32 like the reference's torch.cat."""
    levels = len(bands)
    same = all(b.dtype == ll.dtype and b.is_contiguous() for lv in bands for b in lv)
    if levels and same and ll.dim() == 4 and ll.is_contiguous() and \
            ops.pyramid_fusable((ll.shape[-2] << levels, ll.shape[-1] << levels), levels) and \
            all(tuple(lv[0].shape[-2:]) == (ll.shape[-2] << (levels - 1 - k), ll.shape[-1] << (levels - 1 - k)) for k, lv in enumerate(bands)):
        return ops.idwt2_haar_pyramid(bands, ll)
    for hl, lh, hh in reversed(bands):
        ll = ops.idwt2_haar(ll.contiguous(), hl, lh, hh)
    return ll
