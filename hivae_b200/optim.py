"""Train-step tail of the reference's train_amd.py:489-496 on the sm_100a kernels of csrc/optim.cu:

    accelerator.backward(loss)                       -> (autograd: not part of this package yet)
    accelerator.clip_grad_norm_(params, max_norm)    -> clip_grad_norm_(arena, max_norm)
    optimizer.step(); optimizer.zero_grad()          -> AdamW(arena, lr).step() / .zero_grad()
    DDP gradient all-reduce (mean)                   -> GradAllReducer(arena).all_reduce()   (NCCL, bucketed)
    model/loss.py l2                                 -> mse_loss(pre, gt)

B200 layout: every parameter of the module is a view into ONE flat fp32 arena (same for .grad and the two
AdamW moments), so the optimizer is a single 28 B/parameter streaming kernel, the gradient norm a single
4 B/parameter reduction, and the all-reduce runs over a few large contiguous buckets.  Nothing here
synchronises with the host: loss, norm and clip coefficient stay on the device.
"""
import ctypes as C
import math

import torch

from . import _lib

_ALIGN = 4  # elements: 16-byte aligned tensor starts for the 128-bit kernels


def _p(t):
    return C.c_void_p(t.data_ptr())


def _stream(t):
    return C.c_void_p(torch.cuda.current_stream(t.device).cuda_stream) if t.is_cuda else None


class FlatArena:
    """Re-homes the parameters (and gradients) of `params` into flat fp32 buffers.  `p.data` / `p.grad`
    become views, so module code, state_dict() and autograd keep working unchanged."""

    def __init__(self, params):
        self.params = [p for p in params if p.requires_grad]
        if not self.params:
            raise ValueError("FlatArena: no trainable parameters")
        dev = self.params[0].device
        if any(p.dtype != torch.float32 or p.device != dev for p in self.params):
            raise TypeError("FlatArena needs fp32 parameters on one device (fp32 master weights, SURVEY 8a a21)")
        self.offsets, off = [], 0
        for p in self.params:
            self.offsets.append(off)
            off += (p.numel() + _ALIGN - 1) // _ALIGN * _ALIGN
        self.numel = off
        self.flat = torch.zeros(off, dtype=torch.float32, device=dev)
        self.grad = torch.zeros(off, dtype=torch.float32, device=dev)
        for p, o in zip(self.params, self.offsets):
            self.flat[o:o + p.numel()].view_as(p).copy_(p.data)
            p.data = self.flat[o:o + p.numel()].view_as(p)
            p.grad = self.grad[o:o + p.numel()].view_as(p)
        self._ws = None
        self.clip = None  # device tensor [total_norm, clip_coef] of the last clip_grad_norm_

    def workspace(self):
        if self._ws is None:
            n = _lib.load().hv_reduce_workspace_doubles() if self.flat.is_cuda else 1
            self._ws = torch.empty(n, dtype=torch.float64, device=self.flat.device)
        return self._ws

    def zero_grad(self):
        self.grad.zero_()
        self.clip = None


def mse_loss(pre, gt, out=None):
    """model/loss.py:13-14 l2 -> 0-dim device tensor (no host sync)."""
    _lib.require_device()
    pre, gt = pre.contiguous().float(), gt.contiguous().float()
    assert pre.shape == gt.shape and pre.is_cuda
    lib = _lib.load()
    ws = torch.empty(lib.hv_reduce_workspace_doubles(), dtype=torch.float64, device=pre.device)
    out = torch.empty(1, dtype=torch.float32, device=pre.device) if out is None else out
    _lib.check(lib.hv_mse_loss(_p(pre), _p(gt), pre.numel(), _p(ws), _p(out), _stream(pre)))
    return out[0]


def clip_grad_norm_(arena, max_norm, grad_scale=1.0, materialize=False):
    """torch.nn.utils.clip_grad_norm_ semantics over the whole arena; returns the total norm (0-dim device
    tensor).  The scaling itself is deferred into the next AdamW.step() (no extra pass over the gradients);
    materialize=True also scales arena.grad in place, which is what torch makes observable."""
    _lib.require_device()
    arena.clip = torch.empty(2, dtype=torch.float32, device=arena.grad.device)
    _lib.check(_lib.load().hv_grad_norm_clip_coef(_p(arena.grad), arena.numel, float(grad_scale), float(max_norm),
                                                  _p(arena.workspace()), _p(arena.clip), _stream(arena.grad)))
    norm = arena.clip[0]
    if materialize:
        _lib.check(_lib.load().hv_scale_by_device(_p(arena.grad), arena.numel, C.c_void_p(arena.clip.data_ptr() + 4),
                                                  float(grad_scale), _stream(arena.grad)))
        arena.clip = None
    return norm


class AdamW:
    """torch.optim.AdamW (defaults of train_amd.py:255: betas (0.9, 0.999), eps 1e-8, weight_decay 0.01) as one
    fused clip + update kernel over a FlatArena."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2):
        self.arena = params if isinstance(params, FlatArena) else FlatArena(list(params))
        self.lr, self.betas, self.eps, self.weight_decay = lr, betas, eps, weight_decay
        self.exp_avg = torch.zeros_like(self.arena.flat)
        self.exp_avg_sq = torch.zeros_like(self.arena.flat)
        self.step_count = 0
        self.param_groups = [{"lr": lr, "betas": betas, "eps": eps, "weight_decay": weight_decay, "params": self.arena.params}]

    def step(self, grad_scale=1.0):
        _lib.require_device()
        a = self.arena
        self.step_count += 1
        g = self.param_groups[0]
        clip = C.c_void_p(a.clip.data_ptr() + 4) if a.clip is not None else None
        _lib.check(_lib.load().hv_adamw_clip(_p(a.flat), _p(a.grad), _p(self.exp_avg), _p(self.exp_avg_sq), a.numel,
                                             float(g["lr"]), float(g["betas"][0]), float(g["betas"][1]), float(g["eps"]),
                                             float(g["weight_decay"]), self.step_count, float(grad_scale), clip, _stream(a.flat)))
        a.clip = None

    def zero_grad(self, set_to_none=False):
        self.arena.zero_grad()

    def state_dict(self):
        return {"step": self.step_count, "exp_avg": self.exp_avg, "exp_avg_sq": self.exp_avg_sq, "param_groups": [
            {k: v for k, v in self.param_groups[0].items() if k != "params"}]}


class GradAllReducer:
    """DDP-equivalent gradient averaging (train_amd.py:489 through accelerate/DDP): the flat gradient arena is
    all-reduced in `bucket_mb` contiguous buckets on a side stream so that the collective of bucket i overlaps
    whatever still runs on the compute stream; NCCL's AVG does the 1/world (gloo: SUM, 1/world folded into
    the clip / AdamW kernels through `grad_scale`)."""

    def __init__(self, arena, bucket_mb=256, group=None):
        import torch.distributed as dist
        self.dist, self.arena, self.group = dist, arena, group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        n = max(1, int(bucket_mb * (1 << 20) // 4))
        self.buckets = [(s, min(s + n, arena.numel)) for s in range(0, arena.numel, n)]
        self.stream = torch.cuda.Stream(arena.grad.device) if arena.grad.is_cuda else None
        self.backend = dist.get_backend(group) if dist.is_initialized() else None

    def all_reduce(self):
        """returns the grad_scale the following clip_grad_norm_ / AdamW.step must apply (1.0 when NCCL averaged)"""
        if self.world == 1:
            return 1.0
        avg = self.backend == "nccl"
        op = self.dist.ReduceOp.AVG if avg else self.dist.ReduceOp.SUM
        if self.stream is not None:
            self.stream.wait_stream(torch.cuda.current_stream(self.arena.grad.device))
            with torch.cuda.stream(self.stream):
                works = [self.dist.all_reduce(self.arena.grad[s:e], op=op, group=self.group, async_op=True) for s, e in self.buckets]
                for w in works:
                    w.wait()
            torch.cuda.current_stream(self.arena.grad.device).wait_stream(self.stream)
        else:
            for s, e in self.buckets:
                self.dist.all_reduce(self.arena.grad[s:e], op=op, group=self.group)
        return 1.0 if avg else 1.0 / self.world
