"""world_size-2 gloo test (CPU) of the host-side sharding logic used by the multi-GPU path."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from hivae_b200 import sharding, synth


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_clips, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    z = synth.synth_clip_latents(n_clips, seed=3)
    mine = sharding.shard_clips(z, rank, world)
    # stand-in for the per-clip hot path (clips are independent): any per-clip function
    local = mine.flatten(1).sum(1, keepdim=True) * (1.0 + mine.flatten(1)[:, :4])
    full = sharding.gather_clips(local, n_clips, rank, world)
    ref = z.flatten(1).sum(1, keepdim=True) * (1.0 + z.flatten(1)[:, :4])
    ok = torch.equal(full, ref)
    slow = sharding.max_over_ranks(10.0 + rank, torch.device("cpu"), world)
    q.put((rank, ok, slow, mine.shape[0]))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_indices_partition():
    for n in (0, 1, 7, 8, 32):
        for world in (1, 2, 4, 8):
            parts = [sharding.shard_indices(n, r, world) for r in range(world)]
            assert sorted(i for p in parts for i in p) == list(range(n))
            assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1


def test_gloo_world2_shard_gather_roundtrip():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port, n = _free_port(), 5
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert [r[1] for r in res] == [True, True]
    assert [r[2] for r in res] == [11.0, 11.0]  # max over ranks
    assert [r[3] for r in res] == [3, 2]
