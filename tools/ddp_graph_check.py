"""World-size-N check of the GRAPHED data-parallel train step (torchrun, N GPUs): the step with its NCCL bucket all-reduces
captured in one CUDA graph per rank (TrainStep.step_graphed) against the eager step: parameter deltas after two steps from the
same initial weights (small config, fp32), then timing of both on the default 695 M model (bf16, 4 clips per GPU)."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hivae_b200 as hv  # noqa: E402
from hivae_b200 import configs, optim, synth  # noqa: E402


def build(dev, dtype, cfg):
    m = hv.AMDModel_New(**cfg)
    m.load_state_dict(synth.synth_state_dict(((k, v.shape) for k, v in m.state_dict().items()), 0))
    return m.to(dev).set_compute_dtype(dtype)


def timed(fn, steps, dev):
    dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / steps], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


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
    out = {"world": world}
    # ---- equality: two steps, eager vs graphed, same host RNG stream for the time steps, same CUDA seed for the noise
    res = {}
    for mode in ("eager", "graph"):
        m = build(dev, torch.float32, configs.SMALL)
        p0 = {k: p.detach().clone() for k, p in m.named_parameters()}
        step = optim.TrainStep(m, bucket_mb=4, distributed=True, issue_after=0.0)
        torch.manual_seed(5)
        fn = step.step_graphed if mode == "graph" else step.step
        losses = []
        for s in (1, 2):
            v, r = synth.split_clip(synth.synth_clip_latents(1, seed=10 * s + rank))
            losses.append(float(fn(v.to(dev), r.to(dev))["loss"]))
        res[mode] = (losses, torch.cat([(p.detach() - p0[k]).reshape(-1) for k, p in m.named_parameters()]))
    d = float((res["graph"][1] - res["eager"][1]).norm() / res["eager"][1].norm())
    out["small_fp32"] = {"loss_eager": res["eager"][0], "loss_graph": res["graph"][0], "param_delta_rms_rel_graph_vs_eager": d}
    assert d <= 2e-3, d
    # ---- timing on the default model
    m = build(dev, torch.bfloat16, configs.DEFAULT)
    step = optim.TrainStep(m, lr=1e-4, distributed=True)
    v, r = synth.split_clip(synth.synth_clip_latents(4, seed=101 + rank))
    v, r = v.to(dev), r.to(dev)
    for _ in range(2):
        step.step(v, r)
    tuned = step.tune_schedule(v, r)
    ms_eager = timed(lambda: step.step(v, r), 3, dev)
    step.step_graphed(v, r)
    ms_graph = timed(lambda: step.step_graphed(v, r), 3, dev)
    out["default_bf16_4clips"] = {"ms_eager": round(ms_eager, 2), "ms_graph": round(ms_graph, 2), "issue_after": step.reducer.issue_after,
                                  "schedule_candidates_fwd_bwd_ms": {str(k): round(x, 2) for k, x in tuned.items()}}
    if rank == 0:
        print(json.dumps(out))
    dist.barrier()


if __name__ == "__main__":
    main()
