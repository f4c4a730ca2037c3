"""Gaussian low-pass mask and 3-D frequency split with the reference's API
(model/frequency_utils.py:45-108).

`ifftn(fftshift-mask * fftn(x)).real` with a real, even, separable mask is exactly three 1-D
circular convolutions, so the device kernel applies one small circulant matrix per axis instead
of running three complex FFTs.  The matrices are derived on the host in float64.
"""
import numpy as np
import torch

from . import ops

_CIRC_CACHE = {}


def gaussian_low_pass_filter(shape, d_s=0.25, d_t=0.25):
    """Same values / shape / CPU placement as the reference (frequency_utils.py:45-81); the result is
    tagged with its cut-offs so freq_3d_filter can use the separable kernel directly."""
    T, H, W = shape[-3], shape[-2], shape[-1]
    if d_s == 0 or d_t == 0:
        mask = torch.zeros(shape)
    else:
        t = torch.arange(T).float() * 2 / T - 1
        h = torch.arange(H).float() * 2 / H - 1
        w = torch.arange(W).float() * 2 / W - 1
        gt, gh, gw = torch.meshgrid(t, h, w, indexing="ij")
        mask = torch.exp(-0.5 * ((gt * (1 / d_t)).pow(2) + (gh * (1 / d_s)).pow(2) + (gw * (1 / d_s)).pow(2)))
        if len(shape) > 3:
            mask = mask.unsqueeze(0).unsqueeze(0).repeat(shape[0], shape[1], 1, 1, 1)
    mask._hv_gauss = (float(d_s), float(d_t))
    return mask


def _gauss_1d(N, d):
    x = np.arange(N, dtype=np.float32).astype(np.float32) * np.float32(2) / np.float32(N) - np.float32(1)
    return np.exp(-0.5 * (x.astype(np.float64) / d) ** 2) if d != 0 else np.zeros(N)


def _circulant(g):
    """shifted-spectrum 1-D mask g (DC at index N//2) -> N x N real circulant matrix of the filter."""
    N = g.shape[0]
    if N % 2:
        raise NotImplementedError("freq_3d_filter: odd axis length %d (the shifted mask is not even-symmetric)" % N)
    m = np.fft.ifftshift(g.astype(np.float64))
    if not np.allclose(m[1:], m[1:][::-1], rtol=0, atol=1e-5 * max(1.0, np.abs(m).max())):
        raise NotImplementedError("freq_3d_filter: mask is not even-symmetric along an axis")
    # (an fp32-rounded mask is symmetric only to ~1e-7; .real of the reference keeps the even part)
    h = np.fft.ifft(m).real
    idx = (np.arange(N)[:, None] - np.arange(N)[None, :]) % N
    return h[idx]


def circulants(T, H, W, d_s, d_t, device):
    key = (T, H, W, float(d_s), float(d_t), str(device))
    if key not in _CIRC_CACHE:
        mats = [_circulant(_gauss_1d(n, d)) for n, d in ((T, d_t), (H, d_s), (W, d_s))]
        _CIRC_CACHE[key] = tuple(torch.from_numpy(m).float().contiguous().to(device) for m in mats)
    return _CIRC_CACHE[key]


def _factor_mask(LPF):
    """rank-1 factorisation mask[t,h,w] = a[t] b[h] c[w] of an arbitrary (untagged) mask, or raise."""
    m = LPF.detach().double().cpu()
    while m.dim() > 3:
        if not torch.equal(m[0], m[-1]):
            raise NotImplementedError("freq_3d_filter: per-volume masks are not supported")
        m = m[0]
    m = m.numpy()
    i = np.unravel_index(np.argmax(np.abs(m)), m.shape)
    if m[i] == 0:
        return [np.zeros(n) for n in m.shape]
    a, b, c = m[:, i[1], i[2]], m[i[0], :, i[2]] / m[i], m[i[0], i[1], :] / m[i]
    if not np.allclose(a[:, None, None] * b[None, :, None] * c[None, None, :], m, rtol=1e-5, atol=1e-7):
        raise NotImplementedError("freq_3d_filter: only separable masks are supported by the sm_100a kernel")
    return [a, b, c]


def freq_3d_filter(x, LPF):
    """(x_low, x_high) of frequency_utils.py:84-108 for x (..., T, H, W) on the GPU."""
    T, H, W = x.shape[-3:]
    tag = getattr(LPF, "_hv_gauss", None)
    if tag is not None:
        ct, ch, cw = circulants(T, H, W, tag[0], tag[1], x.device)
    else:
        ct, ch, cw = (torch.from_numpy(_circulant(g)).float().contiguous().to(x.device) for g in _factor_mask(LPF))
    return freq_split_gauss(x, None, None, mats=(ct, ch, cw))


def freq_split_gauss(x, d_s, d_t, mats=None):
    """device-resident form used by the model: no mask tensor, no H2D copy per call."""
    T, H, W = x.shape[-3:]
    ct, ch, cw = mats if mats is not None else circulants(T, H, W, d_s, d_t, x.device)
    return ops.freq_split3d(x.float().contiguous(), ct, ch, cw)
