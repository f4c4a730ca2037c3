"""CPU tier of the train-step tail: flat-arena aliasing, bucket layout, and the world_size-2 gloo path of the
gradient all-reduce (host logic only; the kernels themselves are covered by the gpu tier)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from hivae_b200 import optim


def _params(seed):
    g = torch.Generator().manual_seed(seed)
    return [torch.nn.Parameter(torch.randn(*s, generator=g)) for s in [(5, 3), (7,), (2, 2, 2), (1,)]]


def test_flat_arena_views_alias_and_keep_values():
    ps = _params(0)
    before = [p.detach().clone() for p in ps]
    arena = optim.FlatArena(ps)
    assert arena.numel == 16 + 8 + 8 + 4 and all(o % 4 == 0 for o in arena.offsets)  # 16-byte aligned starts
    for p, b, o in zip(ps, before, arena.offsets):
        assert torch.equal(p.data, b)
        assert p.data.data_ptr() == arena.flat.data_ptr() + 4 * o and p.grad.data_ptr() == arena.grad.data_ptr() + 4 * o
    arena.flat.add_(1.0)
    assert all(torch.equal(p.data, b + 1.0) for p, b in zip(ps, before))
    with torch.enable_grad():  # other test modules switch grad mode off process-wide
        (ps[0].sum() * 2.0 + ps[1].sum()).backward()  # autograd accumulates into the arena views in place
    assert float(arena.grad[:15].min()) == 2.0 and float(arena.grad[16:23].max()) == 1.0 and float(arena.grad[15]) == 0.0
    arena.zero_grad()
    assert float(arena.grad.abs().max()) == 0.0


def test_bucket_layout_covers_arena_once():
    arena = optim.FlatArena([torch.nn.Parameter(torch.zeros(1000)), torch.nn.Parameter(torch.zeros(37))])
    r = optim.GradAllReducer(arena, bucket_mb=1e-3)  # ~262 elements per bucket
    assert r.buckets[0][0] == 0 and r.buckets[-1][1] == arena.numel
    assert all(a[1] == b[0] for a, b in zip(r.buckets, r.buckets[1:])) and len(r.buckets) > 2
    assert r.all_reduce() == 1.0  # no process group: single replica


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    arena = optim.FlatArena(_params(1))  # same init on every rank (set_seed is rank-independent, train_amd.py:154)
    for i, p in enumerate(arena.params):
        p.grad.fill_(float(rank + 1) * (i + 1))
    scale = optim.GradAllReducer(arena, bucket_mb=4e-5).all_reduce()  # ~10 elements per bucket
    mean = [float((p.grad * scale).flatten()[0]) for p in arena.params]
    q.put((rank, scale, mean))
    dist.barrier()
    dist.destroy_process_group()


def test_gloo_world2_gradient_all_reduce_mean():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for _rank, scale, mean in res:
        assert scale == 0.5                      # gloo: SUM, the 1/world is folded into the clip / AdamW kernels
        assert mean == [1.5, 3.0, 4.5, 6.0]      # mean over ranks of (rank+1)*(i+1)


def _worker_overlap(rank, world, port, q):
    """the backward reports finished parameters out of arena order; buckets go out as soon as their last parameter is done"""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    ps = _params(1)
    names = {"p%d" % i: p for i, p in enumerate(ps)}
    arena = optim.FlatArena(ps)
    red = optim.GradAllReducer(arena, bucket_mb=4e-5, issue_after=0.0)  # ~10 elements per bucket: parameters straddle buckets
    issued_after = []
    red.begin(names)
    for name in ("p3", "p1", "p0", "p2"):        # completion order != arena order
        i = int(name[1])
        names[name].grad.fill_(float(rank + 1) * (i + 1))
        red.done([name])
        issued_after.append(sum(red._issued))
    scale = red.finish()
    mean = [float((p.grad * scale).flatten()[0]) for p in arena.params]
    q.put((rank, scale, mean, issued_after, len(red.buckets), all(red._issued)))
    dist.barrier()
    dist.destroy_process_group()


def test_gloo_world2_bucketed_all_reduce_overlapped_with_backward():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker_overlap, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for _rank, scale, mean, issued_after, nb, all_issued in res:
        assert scale == 0.5 and mean == [1.5, 3.0, 4.5, 6.0] and all_issued
        assert issued_after == sorted(issued_after) and issued_after[-1] == nb   # every bucket exactly once, none before its parameters
        assert issued_after[0] < nb                                              # ... and not all at the end


def test_backward_order_covers_every_parameter_once():
    import hivae_b200 as hv
    from hivae_b200 import configs
    m = hv.AMDModel_New(**configs.SMALL)
    order = optim.backward_order(m)
    assert sorted(k for k, _ in order) == sorted(k for k, _ in m.named_parameters())
    names = [k for k, _ in order]
    # the denoiser head completes first, layer i + 1 before layer i, the encoders after the denoiser
    assert names[0].startswith("diffusion_transformer.") and ("proj_out" in names[0] or "norm_" in names[0])
    assert names.index("diffusion_transformer.spatial_blocks.1.ff.net.2.weight") < names.index("diffusion_transformer.spatial_blocks.0.ff.net.2.weight")
    assert max(i for i, k in enumerate(names) if k.startswith("diffusion_transformer.")) < min(i for i, k in enumerate(names) if "motion_encoder." in k)
