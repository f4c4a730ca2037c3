"""Sin-cos position tables (non-persistent buffers of the reference, rebuilt at construction).

Semantics of diffusers 0.30.0 `get_2d_sincos_pos_embed` / `get_1d_sincos_pos_embed_from_grid`
as called at model/transformer.py:84-88, 462-465, 1812-1828: float64 numpy, [sin | cos] halves,
base_size 16 grid scaling and the (w, h) meshgrid order.
"""
import numpy as np
import torch


def sincos_1d(dim, pos):
    if dim % 2:
        raise ValueError("embed_dim must be even")
    omega = 1.0 / 10000 ** (np.arange(dim // 2, dtype=np.float64) / (dim / 2.0))
    ang = np.einsum("m,d->md", np.asarray(pos, dtype=np.float64).reshape(-1), omega)
    return np.concatenate([np.sin(ang), np.cos(ang)], axis=1)


def sincos_2d(dim, grid_h, grid_w, base_size=16, interpolation_scale=1.0):
    gh = np.arange(grid_h, dtype=np.float32) / (grid_h / base_size) / interpolation_scale
    gw = np.arange(grid_w, dtype=np.float32) / (grid_w / base_size) / interpolation_scale
    grid = np.stack(np.meshgrid(gw, gh), axis=0).reshape([2, 1, grid_w, grid_h])
    return np.concatenate([sincos_1d(dim // 2, grid[0]), sincos_1d(dim // 2, grid[1])], axis=1)


def table(arr):
    """(L, D) float64 numpy -> (1, L, D) fp32 tensor, as the reference's register_buffer copies do."""
    t = torch.zeros(1, *arr.shape)
    t.data.copy_(torch.from_numpy(arr))
    return t
