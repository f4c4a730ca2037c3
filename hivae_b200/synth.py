"""Deterministic, name-keyed synthetic weights and clips (no datasets/checkpoints offline).

Every tensor is drawn from its own generator seeded by crc32(name) ^ seed, so the same
values are produced regardless of module construction order — the reference model, the
CPU oracle and the B200 modules can all be loaded with identical weights by name.
"""
import zlib

import torch


def synth_tensor(name, shape, seed=0):
    g = torch.Generator().manual_seed((zlib.crc32(name.encode()) ^ (seed * 0x9E3779B1)) & 0x7FFFFFFF)
    shape = tuple(shape)
    leaf = name.rsplit(".", 1)[-1]
    if "token" in name and leaf != "weight" and leaf != "bias":
        return torch.randn(shape, generator=g) * 0.02
    is_norm = any(s in name for s in (".norm", "norm_q", "norm_k", "norm_final")) and ".linear." not in name
    if len(shape) >= 2:
        fan_in = 1
        for s in shape[1:]:
            fan_in *= s
        return (torch.rand(shape, generator=g) * 2 - 1) / fan_in ** 0.5
    if is_norm and leaf == "weight":
        return 1.0 + 0.1 * torch.randn(shape, generator=g)
    if is_norm and leaf == "bias":
        return 0.1 * torch.randn(shape, generator=g)
    return 0.05 * torch.randn(shape, generator=g)


def synth_state_dict(named_shapes, seed=0):
    """named_shapes: iterable of (name, shape) — e.g. ((k, v.shape) for k, v in model.state_dict().items())."""
    return {k: synth_tensor(k, s, seed) for k, s in named_shapes}


def synth_clip_latents(n, frames=17, c=4, h=32, w=32, seed=1):
    """VAE-latent-like synthetic clips: randn * 0.18215 * 4."""
    g = torch.Generator().manual_seed(seed)
    return torch.randn(n, frames, c, h, w, generator=g) * (0.18215 * 4)


def split_clip(z):
    """(n,17,c,h,w) -> video (n,16,...), ref_img = frame 0 repeated 16x (pipeline.py:685-708)."""
    t = z.shape[1] - 1
    return z[:, 1:].contiguous(), z[:, :1].repeat(1, t, 1, 1, 1).contiguous()
