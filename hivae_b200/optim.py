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
        self.sync_grads(copy_stray=False)
        self.grad.zero_()
        self.clip = None

    def sync_grads(self, copy_stray=True):
        """`p.grad` is a view of the gradient arena only while nobody rebinds it: `module.zero_grad()` (set_to_none=True by
        default in torch 2.x) or `p.grad = ...` leave the arena stale and the step would silently see zeros.  Every
        parameter is checked; a stray gradient is copied into its arena slice (or treated as zero when it is None, as
        torch's optimizers skip such parameters... which here means a zero update direction) and `p.grad` re-bound."""
        for p, o in zip(self.params, self.offsets):
            view = self.grad[o:o + p.numel()].view_as(p)
            g = p.grad
            if g is not None and g.data_ptr() == view.data_ptr():
                continue
            if copy_stray:
                if g is None:
                    view.zero_()
                else:
                    view.copy_(g)
            p.grad = view


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
    arena.sync_grads()
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
        self._state = None      # device-resident {steps done, lr} (fp64) once enable_device_state() was called
        self._derived = None

    def enable_device_state(self):
        """Keep the step count and the learning rate in device memory (hv_adamw_clip_devstate): needed for a step that is
        replayed from a CUDA graph, where host scalars would be frozen at their capture-time values.  Change the learning
        rate with set_lr() afterwards (a scheduler that writes param_groups[0]['lr'] is picked up by step() in eager mode
        only)."""
        dev = self.arena.flat.device
        self._state = torch.tensor([float(self.step_count), float(self.param_groups[0]["lr"])], dtype=torch.float64, device=dev)
        self._derived = torch.zeros(4, dtype=torch.float32, device=dev)
        self._state_lr = float(self.param_groups[0]["lr"])

    def set_lr(self, lr):
        self.param_groups[0]["lr"] = lr
        self.lr = lr
        if self._state is not None and float(lr) != self._state_lr:
            self._state[1:2].fill_(float(lr))
            self._state_lr = float(lr)

    def step(self, grad_scale=1.0):
        _lib.require_device()
        a = self.arena
        if a.clip is None:
            a.sync_grads()  # clip_grad_norm_ already did otherwise
        self.step_count += 1
        g = self.param_groups[0]
        clip = C.c_void_p(a.clip.data_ptr() + 4) if a.clip is not None else None
        if self._state is not None:
            if not torch.cuda.is_current_stream_capturing():
                self.set_lr(g["lr"])
            _lib.check(_lib.load().hv_adamw_clip_devstate(_p(a.flat), _p(a.grad), _p(self.exp_avg), _p(self.exp_avg_sq), a.numel,
                                                          _p(self._state), _p(self._derived), float(g["betas"][0]), float(g["betas"][1]),
                                                          float(g["eps"]), float(g["weight_decay"]), float(grad_scale), clip, _stream(a.flat)))
            a.clip = None
            from . import engine
            engine.bump_weight_epoch()
            return
        _lib.check(_lib.load().hv_adamw_clip(_p(a.flat), _p(a.grad), _p(self.exp_avg), _p(self.exp_avg_sq), a.numel,
                                             float(g["lr"]), float(g["betas"][0]), float(g["betas"][1]), float(g["eps"]),
                                             float(g["weight_decay"]), self.step_count, float(grad_scale), clip, _stream(a.flat)))
        a.clip = None
        # the arena was written through raw pointers (no p._version bump): prepared bf16 / packed weight copies are stale
        from . import engine
        engine.bump_weight_epoch()

    def zero_grad(self, set_to_none=False):
        """gradients stay views of the arena (set_to_none is accepted for API compatibility and ignored: a None gradient
        would detach the parameter from the flat arena)"""
        self.arena.zero_grad()

    def state_dict(self):
        return {"step": self.step_count, "exp_avg": self.exp_avg, "exp_avg_sq": self.exp_avg_sq, "param_groups": [
            {k: v for k, v in self.param_groups[0].items() if k != "params"}]}


class GradAllReducer:
    """DDP-equivalent gradient averaging (train_amd.py:489 through accelerate/DDP): the flat gradient arena is
    all-reduced in `bucket_mb` contiguous buckets on a side stream so that the collective of bucket i overlaps
    whatever still runs on the compute stream; NCCL's AVG does the 1/world (gloo: SUM, 1/world folded into
    the clip / AdamW kernels through `grad_scale`)."""

    def __init__(self, arena, bucket_mb=256, group=None, issue_after=0.0):
        """issue_after: fraction of the buckets that must be complete before the first one is issued during the backward
        (0 = every bucket as soon as it completes, DDP style; 1 = all at the end).  The compute kernels of this package are
        persistent, one CTA (pair) per SM: an NCCL kernel that holds SMs while they run delays whole waves of tiles, so
        on B200 a coarse schedule — everything of the denoiser (93 % of the bytes) once its backward is done, overlapped with
        the encoders' backward — exposes less than the fine-grained one (measured: profiles/r02_*_bench_n2.json)."""
        import torch.distributed as dist
        self.dist, self.arena, self.group = dist, arena, group
        self.issue_after = issue_after
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        n = max(1, int(bucket_mb * (1 << 20) // 4))
        self._bucket_elems = n
        self._issued, self._works = [], []
        self.buckets = [(s, min(s + n, arena.numel)) for s in range(0, arena.numel, n)]
        self.stream = torch.cuda.Stream(arena.grad.device) if arena.grad.is_cuda else None
        self.backend = dist.get_backend(group) if dist.is_initialized() else None

    # ---- overlap with the backward pass: the backward reports finished parameters, a bucket is issued on the side stream
    # as soon as the last of its parameters is final (DDP's bucketed all-reduce, train_amd.py:489 through accelerate)
    def begin(self, name_to_param):
        """arm the per-bucket counters for one backward pass; name_to_param: {parameter name: parameter}"""
        a = self.arena
        off = {id(p): o for p, o in zip(a.params, a.offsets)}
        self._bucket_of, self._pending, self._works = {}, [0] * len(self.buckets), []
        for name, p in name_to_param.items():
            o = off.get(id(p))
            if o is None:
                continue
            b0, b1 = o // self._bucket_elems, (o + max(p.numel(), 1) - 1) // self._bucket_elems
            self._bucket_of[name] = range(b0, b1 + 1)
            for b in range(b0, b1 + 1):
                self._pending[b] += 1
        self._issued = [False] * len(self.buckets)

    def done(self, names):
        """parameters whose gradient is final; issues every bucket that became complete"""
        if self.world == 1:
            return
        for name in names:
            for b in self._bucket_of.pop(name, ()):
                self._pending[b] -= 1
        ready = [b for b in range(len(self.buckets)) if self._pending[b] == 0 and not self._issued[b]]
        done = sum(1 for b in range(len(self.buckets)) if self._pending[b] == 0)
        if ready and done >= self.issue_after * len(self.buckets):
            for b in ready:
                self._issue(b)

    def _issue(self, b):
        self._issued[b] = True
        s, e = self.buckets[b]
        avg = self.backend == "nccl"
        op = self.dist.ReduceOp.AVG if avg else self.dist.ReduceOp.SUM
        if self.stream is not None:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(self.arena.grad.device))
            self.stream.wait_event(ev)
            with torch.cuda.stream(self.stream):
                self._works.append(self.dist.all_reduce(self.arena.grad[s:e], op=op, group=self.group, async_op=True))
        else:
            self.dist.all_reduce(self.arena.grad[s:e], op=op, group=self.group)

    def finish(self):
        """issue what is left, wait for every bucket; returns the grad_scale for clip / AdamW (1.0 when NCCL averaged)"""
        if self.world == 1:
            return 1.0
        for b in range(len(self.buckets)):
            if not self._issued[b]:
                self._issue(b)
        if self.stream is not None:
            with torch.cuda.stream(self.stream):
                for w in self._works:
                    w.wait()
            torch.cuda.current_stream(self.arena.grad.device).wait_stream(self.stream)
        self._works = []
        return 1.0 if self.backend == "nccl" else 1.0 / self.world

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


def backward_order(model):
    """parameters in the order their gradients become final in hivae_b200.train's backward (denoiser head, denoiser
    layers last to first, denoiser embeddings + adaLN Linears, object encoder, camera encoder, camera_down), so that the
    flat arena fills front to back and buckets complete contiguously"""
    named = dict(model.named_parameters())
    order, seen = [], set()

    def take(pred):
        for k in named:
            if k not in seen and pred(k):
                seen.add(k)
                order.append(k)

    dt = "diffusion_transformer."
    take(lambda k: k.startswith(dt) and any(k.startswith(dt + h) for h in ("proj_out.", "norm_out.norm.", "norm_final.")))
    nl = 0
    for k in named:
        if k.startswith(dt) and "_blocks." in k:
            nl = max(nl, int(k.split("_blocks.")[1].split(".")[0]) + 1)
    for i in reversed(range(nl)):
        for fam in ("spatial_blocks", "camera_transformer_blocks", "object_transformer_blocks"):
            take(lambda k, fam=fam, i=i: k.startswith("%s%s.%d." % (dt, fam, i)) and ".linear." not in k)
    take(lambda k: k.startswith(dt))
    take(lambda k: k.startswith("object_motion_encoder."))
    take(lambda k: k.startswith("camera_motion_encoder."))
    take(lambda k: True)
    return [(k, named[k]) for k in order]


class TrainStep:
    """One optimisation step of train_amd.py:459-496 without an autograd graph:
        forward (activations kept) -> loss -> hand-chained backward writing into the flat gradient arena, gradient buckets
        all-reduced on a side stream as they complete -> clip_grad_norm_(max_grad_norm) -> AdamW -> zero_grad.
    `model` is a hivae_b200.AMDModel_New on the device; parameters with requires_grad=False stay frozen
    (`--frozen_name`, train_amd.py:226-233)."""

    def __init__(self, model, lr=1e-4, max_grad_norm=1.0, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2, bucket_mb=128,
                 group=None, distributed=None, issue_after=0.9):
        import torch.distributed as dist
        self.model = model
        named = [(k, p) for k, p in backward_order(model) if p.requires_grad]
        self.names = [k for k, _ in named]
        self.arena = FlatArena([p for _, p in named])
        self.opt = AdamW(self.arena, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay)
        self.max_grad_norm = max_grad_norm
        self.sink = {k: p.grad for k, p in named}
        self.params = dict(named)
        use_dist = dist.is_initialized() if distributed is None else distributed
        self.reducer = GradAllReducer(self.arena, bucket_mb=bucket_mb, group=group, issue_after=issue_after) if use_dist else None
        self.last = {}
        self._graph, self._graph_key, self._gv, self._gr, self._gt, self._gout = None, None, None, None, None, None

    def _draw_time_step(self, n, t, dev):
        """the reference's draws, in its order (model_AMD.py:1961-1965): on the HOST generator, then copied to the device"""
        m = self.model
        time_step = m.prepare_timestep(batch_size=n * t, device=dev)
        if m.diffusion_model_type != "default":
            time_step = m.prepare_timestep(batch_size=n, device=dev).repeat_interleave(t)
        return time_step

    def forward_backward(self, video, ref_img, time_step=None, noise=None):
        """forward + backward + gradient all-reduce; leaves the (averaged, unclipped) gradients in the arena and returns
        (loss, rec_loss, grad_scale) — grad_scale is what clip / AdamW still have to apply (1.0 unless gloo summed)"""
        from . import train
        from .backward import mse_backward
        m = self.model
        n, t = video.shape[:2]
        dev = video.device
        zj = video.flatten(0, 1).float().contiguous()
        if time_step is None:
            time_step = self._draw_time_step(n, t, dev)
        zt, vel = m.scheduler.get_train_tuple(z0=noise, z1=zj, time_step=time_step)
        self.arena.zero_grad()
        run = train.TrainRun(m)
        pre = run.forward(video, ref_img, time_step.float(), zt.contiguous(), self.sink)
        vel = vel.contiguous()
        loss = mse_loss(pre, vel)
        rec = mse_loss(m.scheduler.get_target_with_zt_vel(zt, pre, time_step), zj)
        if self.reducer is not None:
            self.reducer.begin(self.params)
        run.backward(mse_backward(pre, vel), self.reducer.done if self.reducer is not None else None)
        scale = self.reducer.finish() if self.reducer is not None else 1.0
        return loss, rec, scale

    def tune_schedule(self, video, ref_img, candidates=(0.0, 0.9, 1.0), steps=2):
        """Pick the all-reduce issue point (GradAllReducer.issue_after) that gives the shortest forward+backward on THIS
        box: each candidate is run `steps` times after one untimed pass, timed with CUDA events, max over ranks.  Nothing is
        applied to the weights (no clip / AdamW), so it can run at any point of a job.  Which schedule wins depends on how
        many SMs NCCL takes for the ring/NVLS kernel at this world size, hence measured instead of assumed.
        Returns {candidate: ms}."""
        import torch.distributed as dist
        if self.reducer is None or self.reducer.world == 1 or not video.is_cuda:
            return {}
        out = {}
        for c in candidates:
            self.reducer.issue_after = c
            self.forward_backward(video, ref_img)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            dist.barrier(self.reducer.group)
            torch.cuda.synchronize()
            e0.record()
            for _ in range(steps):
                self.forward_backward(video, ref_img)
            e1.record()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1) / steps], device=video.device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.reducer.group)
            out[c] = float(t)
        self.reducer.issue_after = min(out, key=out.get)   # the same on every rank: t is the max over ranks
        self.arena.zero_grad()
        return out

    # ---- the whole step as ONE CUDA graph.  A step is ~3900 kernel launches through ctypes; the host needs ~110 ms to enqueue
    # them, about what the GPU needs to run them, so the eager step is launch-bound.  Replaying a captured graph removes the
    # host from the loop.  What makes the step capturable: no host synchronisation anywhere in it, time steps / noise / masks
    # (time steps drawn on the host before the replay into a static buffer, noise / masks by torch's graph-safe CUDA generator), the AdamW step count and learning rate in device memory.
    def capture(self, video, ref_img):
        """Capture step() for batches of this shape.  Leaves parameters, AdamW moments, step count and the RNG stream as they
        were (they are snapshotted around the warm-up step that the capture needs)."""
        dev = video.device
        if self.opt._state is None:
            self.opt.enable_device_state()
        self._gv, self._gr = video.clone(), ref_img.clone()
        cpu_rng = torch.get_rng_state()
        self._gt = self._draw_time_step(video.shape[0], video.shape[1], dev).clone()  # static input: drawn on the host per replay
        torch.set_rng_state(cpu_rng)
        snap = [t.clone() for t in (self.arena.flat, self.opt.exp_avg, self.opt.exp_avg_sq, self.opt._state)]
        steps, rng = self.opt.step_count, torch.cuda.get_rng_state(dev)
        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):   # eager warm-up: every buffer, weight copy and kernel attribute exists before the capture
            self.step(self._gv, self._gr, time_step=self._gt)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self._graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self._graph):
            self._gout = self.step(self._gv, self._gr, time_step=self._gt)
        for dst, src in zip((self.arena.flat, self.opt.exp_avg, self.opt.exp_avg_sq, self.opt._state), snap):
            dst.copy_(src)
        self.opt.step_count = steps
        torch.cuda.set_rng_state(rng, dev)
        from . import engine
        engine.bump_weight_epoch()
        self._graph_key = (tuple(video.shape), tuple(ref_img.shape), video.dtype, ref_img.dtype)

    def tune_schedule_graphed(self, video, ref_img, candidates=(0.0, 0.9, 1.0), steps=3):
        """tune_schedule for the graphed step: the issue order of the gradient buckets is frozen into the captured graph, and
        which order exposes least differs between the eager step (host-bound, gaps everywhere) and the replayed one.  Every
        candidate is captured, replayed `steps` times (CUDA events, max over ranks), the best one is kept.  Parameters, AdamW
        moments, step count and the RNG streams are restored afterwards.  Returns {candidate: ms}."""
        import torch.distributed as dist
        if self.reducer is None or self.reducer.world == 1 or not video.is_cuda:
            return {}
        dev = video.device
        if self.opt._state is None:
            self.opt.enable_device_state()
        snap = [t.clone() for t in (self.arena.flat, self.opt.exp_avg, self.opt.exp_avg_sq, self.opt._state)]
        count, rng, cpu_rng = self.opt.step_count, torch.cuda.get_rng_state(dev), torch.get_rng_state()
        out = {}
        for c in candidates:
            self.reducer.issue_after, self._graph = c, None
            self.step_graphed(video, ref_img)   # capture + first replay
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            dist.barrier(self.reducer.group)
            torch.cuda.synchronize(dev)
            e0.record()
            for _ in range(steps):
                self.step_graphed(video, ref_img)
            e1.record()
            torch.cuda.synchronize(dev)
            t = torch.tensor([e0.elapsed_time(e1) / steps], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.reducer.group)
            out[c] = float(t)
        best = min(out, key=out.get)              # the same on every rank: t is the max over ranks
        if best != self.reducer.issue_after:
            self.reducer.issue_after, self._graph = best, None   # re-captured on the next step_graphed
        for dst, src in zip((self.arena.flat, self.opt.exp_avg, self.opt.exp_avg_sq, self.opt._state), snap):
            dst.copy_(src)
        self.opt.step_count = count
        torch.cuda.set_rng_state(rng, dev)
        torch.set_rng_state(cpu_rng)
        from . import engine
        engine.bump_weight_epoch()
        return out

    def step_graphed(self, video, ref_img):
        """step() replayed from a CUDA graph (captured on first use per batch shape); same return value, the tensors are
        overwritten by the next call"""
        key = (tuple(video.shape), tuple(ref_img.shape), video.dtype, ref_img.dtype)
        if self._graph is None or self._graph_key != key:
            self.capture(video, ref_img)
        self.opt.set_lr(self.opt.param_groups[0]["lr"])
        self._gt.copy_(self._draw_time_step(video.shape[0], video.shape[1], video.device))  # host draw, as in the eager step
        self._gv.copy_(video, non_blocking=True)
        self._gr.copy_(ref_img, non_blocking=True)
        self._graph.replay()
        self.opt.step_count += 1
        from . import engine
        engine.bump_weight_epoch()   # eager users of the model (sample, encode) must refresh their bf16 weight copies
        self.last = self._gout
        return self._gout

    def step(self, video, ref_img, time_step=None, noise=None):
        """returns {'loss', 'diff_loss', 'rec_loss', 'grad_norm'} as 0-dim device tensors (no host synchronisation)"""
        loss, rec, scale = self.forward_backward(video, ref_img, time_step, noise)
        norm = clip_grad_norm_(self.arena, self.max_grad_norm, grad_scale=scale)
        self.opt.step(grad_scale=scale)
        self.last = {"loss": loss, "diff_loss": loss, "rec_loss": rec, "grad_norm": norm}
        return self.last
