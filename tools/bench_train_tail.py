"""Train-step tail of train_amd.py:489-496 at the default model's size on N GPUs (one process per GPU, NCCL):
bucketed gradient all-reduce (mean) of the 2.78 GB fp32 arena -> global-norm clip -> fused AdamW.
Launch:  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/bench_train_tail.py
Times on the device with CUDA events, max over ranks.  The backward pass that would overlap the all-reduce is
not part of this package yet, so the all-reduce time reported here is fully exposed (worst case)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from hivae_b200 import optim

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
torch.cuda.set_device(dev)
if world > 1:
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    dist.init_process_group("nccl", device_id=dev)
N = int(float(os.environ.get("NPARAM", 695e6)))
p = torch.nn.Parameter(torch.randn(N, device=dev) * 0.02)
arena = optim.FlatArena([p])
arena.grad.normal_()
opt = optim.AdamW(arena, lr=1e-4)
red = optim.GradAllReducer(arena, bucket_mb=int(os.environ.get("BUCKET_MB", 256))) if world > 1 else None


def timed(fn, reps=5, warm=2):
    for _ in range(warm):
        fn()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / reps], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def tail():
    s = red.all_reduce() if red else 1.0
    optim.clip_grad_norm_(arena, 1.0, grad_scale=s)
    opt.step(grad_scale=s)


res = {"n_gpus": world, "params": N, "bucket_mb": int(os.environ.get("BUCKET_MB", 256))}
if red:
    res["allreduce_ms"] = round(timed(red.all_reduce), 3)
    res["allreduce_busbw_GBps"] = round(2.0 * (world - 1) / world * 4.0 * N / res["allreduce_ms"] / 1e6, 1)
res["clip_adamw_ms"] = round(timed(lambda: (optim.clip_grad_norm_(arena, 1.0), opt.step())), 3)
res["tail_ms"] = round(timed(tail), 3)
if rank == 0:
    print(json.dumps(res), flush=True)
if world > 1:
    dist.destroy_process_group()
