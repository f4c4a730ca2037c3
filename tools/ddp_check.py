"""World-size-N check of the data-parallel train step (run under torchrun on N GPUs):
    torchrun --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/ddp_check.py
Every rank runs forward + backward on ITS clip, the gradient buckets are all-reduced (NCCL AVG) as they complete; rank 0
then recomputes the gradient of the CONCATENATED batch in one process and compares the two gradient arenas
(DDP semantics: mean over ranks of the per-rank mean loss == mean loss of the concatenated batch for equal shards)."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hivae_b200 as hv  # noqa: E402
from hivae_b200 import configs, optim, synth  # noqa: E402


def build(dev, dtype):
    m = hv.AMDModel_New(**configs.SMALL)
    m.load_state_dict(synth.synth_state_dict(((k, v.shape) for k, v in m.state_dict().items()), 0))
    return m.to(dev).set_compute_dtype(dtype)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    sys.stdout.flush()
    saved = os.dup(1)
    os.dup2(2, 1)
    dist.init_process_group("nccl", device_id=dev)
    dist.all_reduce(torch.zeros(1, device=dev))
    torch.cuda.synchronize()
    os.dup2(saved, 1)
    out = {}
    for dtype in (torch.float32, torch.bfloat16):
        z = synth.synth_clip_latents(world, seed=1)
        video, ref_img = synth.split_clip(z)
        steps = torch.tensor([437.0, 81.0, 903.0, 655.0, 12.0, 777.0, 301.0, 590.0][:world])
        z0 = torch.randn(world * 16, 4, 32, 32, generator=torch.Generator().manual_seed(2))
        m = build(dev, dtype)
        step = optim.TrainStep(m, bucket_mb=4, distributed=True)   # small buckets: several all-reduces in flight during backward
        loss, _, scale = step.forward_backward(video[rank:rank + 1].to(dev), ref_img[rank:rank + 1].to(dev),
                                               time_step=steps[rank:rank + 1].repeat_interleave(16).to(dev),
                                               noise=z0[16 * rank:16 * rank + 16].to(dev))
        torch.cuda.synchronize()
        g_ddp = (step.arena.grad * scale).clone()
        if rank == 0:
            m1 = build(dev, dtype)
            s1 = optim.TrainStep(m1, distributed=False)
            s1.forward_backward(video.to(dev), ref_img.to(dev), time_step=steps.repeat_interleave(16).to(dev), noise=z0.to(dev))
            torch.cuda.synchronize()
            g1 = s1.arena.grad
            err = float((g_ddp - g1).abs().max() / g1.abs().max())
            out[str(dtype)] = {"max_abs_rel_allreduced_vs_concatenated": err, "buckets": len(step.reducer.buckets),
                               "grad_norm_single": float(g1.norm()), "grad_norm_ddp": float(g_ddp.norm())}
            assert err <= (1e-5 if dtype == torch.float32 else 2e-2), err
        dist.barrier()
    if rank == 0:
        print(json.dumps({"world": world, "check": "all-reduced gradients == gradients of the concatenated batch", **out}))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
