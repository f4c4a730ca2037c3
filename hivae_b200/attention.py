"""Stand-alone RoPE multi-head attention with the reference's API (model/attention.py:64-117)."""
import math

import torch
from torch import nn

from . import ops


class Attention(nn.Module):
    """wq/wk/wv (+ an `out` Linear that the reference defines but never applies, attention.py:77)."""

    def __init__(self, dim, max_seq_len, n_heads):
        super().__init__()
        self.wq = nn.Linear(dim, dim)
        self.wk = nn.Linear(dim, dim)
        self.wv = nn.Linear(dim, dim)
        self.out = nn.Linear(dim, dim)
        self.dim, self.max_seq_len, self.n_heads = dim, max_seq_len, n_heads
        self.head_dim = dim // n_heads

    @torch.no_grad()
    def forward(self, x, start_pos=0, inference=True, mask=None):
        bs, L, dim = x.shape
        xf = x.float().contiguous().view(bs * L, dim)
        q, k, v = (torch.empty(bs * L, dim, dtype=torch.float32, device=x.device) for _ in range(3))
        for lin, o in ((self.wq, q), (self.wk, k), (self.wv, v)):
            ops.linear(xf, lin.weight.detach().float().contiguous(), ops.epi_store(o, lin.bias.detach().float().contiguous()))
        if inference:
            if start_pos + L > 2 * self.max_seq_len:
                raise ValueError("sequence exceeds the 2*max_seq_len RoPE table of the reference")
            ops.rope_(q, k, bs, L, self.n_heads, self.head_dim, start_pos)
        m = None
        if mask is not None:
            m = mask.to(device=x.device, dtype=torch.float32).expand(L, L).contiguous()
        out = torch.empty_like(q)
        sm = ops.seq_map(1, L, 0, 1)
        ops.attention(q, k, v, out, bs, self.n_heads, self.head_dim, L, L, sm, sm, 1.0 / math.sqrt(self.head_dim), m)
        return out.view(bs, L, dim).to(x.dtype)
