"""Train-step micro-bench (BASELINE config 5): default 695 M model, bf16 compute, n clips per GPU, synthetic latents.
   python tools/bench_train.py [--clips 4] [--steps 5] [--config default]
Prints per-step ms (CUDA events), clips/s, TFLOP/s at 13.5 TFLOP per clip, peak memory."""
import argparse
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hivae_b200 as hv  # noqa: E402
from hivae_b200 import configs, optim, synth  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--clips", type=int, default=4)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--config", default="default")
    ap.add_argument("--dtype", default="bf16")
    ap.add_argument("--profile", action="store_true")
    ap.add_argument("--graph", action="store_true", help="replay the step from one CUDA graph (TrainStep.step_graphed)")
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    cfg = configs.CONFIGS[a.config]
    m = hv.AMDModel_New(**cfg)
    m.load_state_dict(synth.synth_state_dict(((k, v.shape) for k, v in m.state_dict().items()), 0))
    m = m.to(dev).set_compute_dtype(torch.bfloat16 if a.dtype == "bf16" else torch.float32)
    step = optim.TrainStep(m, lr=1e-4, distributed=False)
    video, ref_img = synth.split_clip(synth.synth_clip_latents(a.clips, seed=1))
    video, ref_img = video.to(dev), ref_img.to(dev)
    torch.manual_seed(0)
    losses = []
    run = step.step_graphed if a.graph else step.step
    for i in range(a.warmup):
        losses.append(float(run(video, ref_img)["loss"]))
    torch.cuda.synchronize()
    from hivae_b200 import ops
    if a.profile:
        ops.PROFILE = []
    n0 = ops.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.time()
    e0.record()
    for i in range(a.steps):
        out = run(video, ref_img)
    e1.record()
    host = (time.time() - t0) / a.steps * 1e3   # time the host needs to ENQUEUE a step (no synchronisation inside the loop)
    torch.cuda.synchronize()
    wall = (time.time() - t0) / a.steps * 1e3
    ms = e0.elapsed_time(e1) / a.steps
    rec = {"config": a.config, "dtype": a.dtype, "graph": bool(a.graph), "clips": a.clips, "ms_per_step": round(ms, 2), "wall_ms_per_step": round(wall, 2), "host_enqueue_ms_per_step": round(host, 2),
           "clips_per_s": round(a.clips / ms * 1e3, 3), "tflops_at_13.5_per_clip": round(13.5 * a.clips / ms * 1e3, 1),
           "launches_per_step": (ops.launch_count() - n0) // a.steps, "loss": [round(x, 5) for x in losses] + [round(float(out["loss"]), 5)],
           "grad_norm": round(float(out["grad_norm"]), 4), "peak_mem_gb": round(torch.cuda.max_memory_allocated() / 2 ** 30, 2)}
    if a.profile:
        summ = ops.profile_summary(ops.PROFILE)
        rec["kernels"] = {k: {"launches": v["launches"] // a.steps, "ms": round(v["ms"] / a.steps, 3),
                              "tflops": round(v["flops"] / v["ms"] / 1e9, 1) if v["ms"] else 0} for k, v in summ.items()}
    print(json.dumps(rec))


if __name__ == "__main__":
    main()
