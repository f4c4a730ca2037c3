#!/usr/bin/env python
"""bench.py — clips/sec of encode + rectified-flow decode of 17-frame 256x256 clips (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W            # the sm_100a path (this repo)
  python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path (oracle port) on host cores

A "step" is one pass of the hot path over one batch of synthetic clips: AMDModel_New.sample =
frequency split -> camera / object motion encoders -> `sample_step` Euler steps of the DiT denoiser, on
BASELINE.json configs[1]: the default Hi-VAE autoencoder (AMD_N, 695 M parameters, random-init
name-seeded weights), bf16 autocast semantics, batch 8 of synthetic 17x4x32x32 VAE latents per GPU
(= 17-frame 256x256 clips at the VAE's 8x down-sampling; the VAE itself is outside the path,
SURVEY.md §8f).  Multi-GPU = batch sharding: every rank runs its own 8 clips, no collective in the
data path ("weak" scaling); times are CUDA-event device times, max over ranks.

One JSON line on stdout (rank 0).  `value` = clips/s with inputs resident in HBM; `e2e` = the same through
the public API with pinned HOST buffers (H2D of the latents and D2H of the reconstruction inside the timed
region); `roofline` = the dominant kernel (tcgen05 GEMM) timed per launch with CUDA events in one extra
instrumented pass; `cpu_baseline` = the CPU oracle timed on this box's host cores on a bounded sample.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

SAMPLE_STEPS = 20          # scripts/inference.sh:30 of the reference (the shipped inference recipe)
CLIPS_PER_GPU = 8          # BASELINE.json configs[1]
TFLOP_PER_CLIP = 0.475 + SAMPLE_STEPS * 4.02  # SURVEY.md §8d algorithmic work


def _peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return {"hbm": float(p["hbm_gbs"]), "tf_burst": float(p["bf16_tflops"]),
                "tf_sustained": float(p.get("bf16_tflops_sustained", p["bf16_tflops"])), "src": "measured"}
    except Exception:
        return {"hbm": 6650.0, "tf_burst": 1590.0, "tf_sustained": 1400.0, "src": "fallback"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx, self.proc, self.path = gpu_index, None, None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 8:
                    continue
                try:
                    sm.append(float(f[1])); mx.append(float(f[2]))
                except ValueError:
                    continue
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            busy = sorted(s for s in sm if s >= 0.5 * max(sm)) or sorted(sm)
            out.update(sm_mhz=busy[len(busy) // 2], sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm))
        return out


# ------------------------------------------------------------------------------------------------
def _oracle_clip_time(threads, euler_steps_timed=1):
    """CPU oracle (restatement of the reference's torch path, oracle/amd_oracle.py) on one clip of the
    default model: encode once + `euler_steps_timed` DiT evaluations, extrapolated to SAMPLE_STEPS."""
    import torch
    from oracle import amd_oracle as O
    from hivae_b200 import configs, synth
    import hivae_b200 as hv
    torch.set_num_threads(threads)
    torch.set_grad_enabled(False)
    cfg = configs.DEFAULT
    if not hasattr(_oracle_clip_time, "orc"):
        m = hv.AMDModel_New(**cfg)
        sd = synth.synth_state_dict(((k, v.shape) for k, v in m.state_dict().items()), 0)
        del m
        _oracle_clip_time.orc = O.AMDOracle(sd, cfg["enc_nhead"], cfg["diffusion_attn_num_heads"], 16, 2, cfg["motion_token_num"])
        _oracle_clip_time.clip = synth.split_clip(synth.synth_clip_latents(1, seed=1))
    orc = _oracle_clip_time.orc
    video, ref_img = _oracle_clip_time.clip
    t0 = time.perf_counter()
    cam, osrc, otgt = orc.encode(video, ref_img)
    t1 = time.perf_counter()
    zi, zj = ref_img.flatten(0, 1), video.flatten(0, 1)
    zt = torch.randn_like(zj)
    for i in range(euler_steps_timed):
        pre = orc.dit(zi, zt, torch.full((16,), 1000.0 - 50.0 * i), cam, osrc, otgt)
        zt = zt + pre / SAMPLE_STEPS
    t2 = time.perf_counter()
    t_enc, t_step = t1 - t0, (t2 - t1) / euler_steps_timed
    return t_enc, t_step, t_enc + SAMPLE_STEPS * t_step


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on this box's host cores.
    The reference is pure Python over torch/diffusers (not installable here: diffusers/timm absent, no
    network), so the arm runs the oracle port of it (kind "port")."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    budget_s = 240.0
    t_enc, t_step, t_clip = _oracle_clip_time(cores)          # warm-up sample (also sizes the run)
    warm = 1
    while warm < args.warmup and (warm + 1) * (t_enc + t_step) < 0.25 * budget_s:
        _oracle_clip_time(cores)
        warm += 1
    steps = max(1, min(args.steps, int(0.75 * budget_s / max(t_enc + t_step, 1e-3))))
    tot_enc = tot_step = 0.0
    t0 = time.perf_counter()
    for _ in range(steps):
        e, s, _c = _oracle_clip_time(cores)
        tot_enc += e
        tot_step += s
    wall = time.perf_counter() - t0
    t_clip = tot_enc / steps + SAMPLE_STEPS * tot_step / steps
    value = 1.0 / t_clip
    sample = ("1 clip per step: encode + 1 of %d Euler steps of the default model (fp32, torch CPU, %d threads), "
              "clip time = t_encode + %d * t_euler_step (Euler steps are identical work)" % (SAMPLE_STEPS, cores, SAMPLE_STEPS))
    line = {
        "impl": "reference", "metric": "clips_per_sec_encode_decode_17x256x256", "value": value, "unit": "clips/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": warm, "ms_per_step": 1e3 * wall / steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": dict(_config(args.gpus), reference_arm="CPU arm: 1 clip per step, encode + 1 Euler step timed, x%d extrapolated "
                       "(identical work per Euler step); uses no GPU and does not grow with --gpus" % SAMPLE_STEPS),
        "gpus_used": 0,
        "cpu_baseline": {"value": value, "unit": "clips/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "clips/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def _config(n_gpus):
    name = "configs[1]" if (SAMPLE_STEPS, CLIPS_PER_GPU) == (20, 8) else "configs[2] sampling sweep"
    return {"workload": name + ": default Hi-VAE autoencoder (AMD_N 695M, random-init), bf16 inference, "
                        "batch %d synthetic 17-frame 256x256 clips (17x4x32x32 VAE latents) per GPU, encode + %d-step "
                        "rectified-flow decode" % (CLIPS_PER_GPU, SAMPLE_STEPS),
            "clips_per_gpu": CLIPS_PER_GPU, "global_clips": CLIPS_PER_GPU * n_gpus, "sample_steps": SAMPLE_STEPS,
            "parallelism": "batch-sharded x%d (no data-path collective)" % n_gpus,
            "l2": "per-step working set (1.39 GB bf16 weights + activations) exceeds the 126 MB L2; no explicit flush"}


# ------------------------------------------------------------------------------------------------
def _secondary(torch, dev, share):
    """The two secondary figures BASELINE.json's metric names.
    attn_tensor_pct: in-step TFLOP/s of the spatial attention launches (algorithmic 4 S^2 64 per (sequence, head)) over the
    measured sustained bf16 peak.  dwt_hbm_pct: the fused Haar pyramid kernels on configs[3] sizes, timed HBM-clean: one
    CUDA graph walks over >= 5 GB of distinct inputs + outputs (40x the 126 MB L2), so no launch can be served from L2.
    `gbs` = the bytes the fused kernel has to move (every input element read once, every output element written once:
    2 N sizeof for the analysis, N (sizeof + 4) for the synthesis) / time — the HBM roofline of the operation;
    `gbs_per_level_formula` = SURVEY 8d's per-level count (read N + write N on every level, sum_k 4^-k more), which
    a fused pyramid beats by construction and is therefore NOT compared with the HBM peak.  A `gbs` above 1.2x the
    measured HBM copy peak would be rejected as cache-served."""
    from hivae_b200 import ops
    pk = _peaks()
    sec = {"peaks": {"hbm_gbs": pk["hbm"], "bf16_tflops_sustained": pk["tf_sustained"], "source": pk["src"]}}
    att = {k: v for k, v in share.items() if k.startswith("attention_L") and v["tflops"] > 0 and int(k.split("L")[1]) >= 64}
    if att:
        ms = sum(v["ms"] for v in att.values())
        tf = sum(v["tflops"] * v["ms"] for v in att.values()) / ms
        sec["attn_tensor_pct"] = round(100.0 * tf / pk["tf_sustained"], 2)
        sec["attn_tflops_in_step"] = round(tf, 1)
        sec["attn_share_of_step"] = round(sum(v["share"] for v in att.values()), 4)
    cases = {}
    for name, T, H, L, dt, inverse in (("fwd_f32_33x3x512x512_L3", 33, 512, 3, torch.float32, False),
                                       ("inv_f32_33x3x512x512_L3", 33, 512, 3, torch.float32, True),
                                       ("fwd_bf16_33x3x512x512_L3", 33, 512, 3, torch.bfloat16, False),
                                       ("fwd_f32_17x3x256x256_L2", 17, 256, 2, torch.float32, False)):
        es = 4 if dt == torch.float32 else 2
        numel = T * 3 * H * H
        geo = sum(4.0 ** -k for k in range(L))
        alg = numel * geo * (es + 4) if inverse else 2 * numel * es * geo
        per_inst = numel * es * 2 if not inverse else numel * (es + 4)   # bytes the fused kernel moves
        n_inst = max(4, min(512, int(5.0e9 / per_inst) + 1))
        xs = [torch.randn(T, 3, H, H, device=dev).to(dt) for _ in range(n_inst)]
        if inverse:
            pyr = [ops.dwt2_haar_pyramid(x, L) for x in xs]
            del xs
            keep = []
            run = lambda: keep.append([ops.idwt2_haar_pyramid(*p) for p in pyr])
        else:
            keep = []
            run = lambda: keep.append([ops.dwt2_haar_pyramid(x, L) for x in xs])
        run(); keep.clear()            # warms the allocator: the graph's outputs are then distinct live blocks
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            run()
        g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        gbs = per_inst * n_inst * 3 / (e0.elapsed_time(e1) * 1e6)
        cases[name] = {"gbs": round(gbs, 1), "pct": round(100.0 * gbs / pk["hbm"], 2), "instances": n_inst,
                       "gbs_per_level_formula": round(alg * n_inst * 3 / (e0.elapsed_time(e1) * 1e6), 1),
                       "rejected": bool(gbs > 1.2 * pk["hbm"])}
        del g, keep
        torch.cuda.empty_cache()
    sec["dwt"] = cases
    head = cases["fwd_f32_33x3x512x512_L3"]
    sec["dwt_hbm_pct"] = None if head["rejected"] else head["pct"]
    return sec


TRAIN_CLIPS_PER_GPU = 4        # scripts/amd_new.sh:40 (--batch_size 4), BASELINE.json configs[4]
TRAIN_TFLOP_PER_CLIP = 13.5    # 3 x (one DiT forward + the encoders)


def _train_leg(torch, dist, dev, m, rank, world, timed, steps=3, warmup=2):
    """BASELINE.json configs[4]: one train step of train_amd.py:459-496 on the default model (bf16 compute, fp32 master
    weights and AdamW moments, random init, synthetic latents, 4 clips per GPU): forward with kept activations ->
    hand-chained backward into the flat gradient arena, gradient buckets all-reduced over NCCL as they complete ->
    clip_grad_norm_(1.0) -> AdamW.  Also timed: the same step without the collective (difference = exposed all-reduce)
    and the all-reduce of the whole arena alone."""
    from hivae_b200 import optim, synth
    m.__dict__.pop("_sample_graphs", None)  # the inference graphs (and their memory pools) are stale once the weights train
    torch.cuda.empty_cache()
    step = optim.TrainStep(m, lr=1e-4, max_grad_norm=1.0, distributed=world > 1)
    video, ref_img = synth.split_clip(synth.synth_clip_latents(TRAIN_CLIPS_PER_GPU, seed=101 + rank).to(dev))
    torch.manual_seed(1234)  # every rank draws the same time steps / noise stream, like the reference's set_seed (train_amd.py:154)
    eager = lambda: step.step(video, ref_img)
    for _ in range(warmup):
        out = eager()
    tuned = step.tune_schedule(video, ref_img)     # untimed: picks when the gradient buckets go out (N > 1 only)
    ms_eager = timed(eager, steps) / steps
    # the step replayed from ONE CUDA graph per rank (TrainStep.step_graphed; the NCCL bucket all-reduces are captured with it):
    # ~3900 ctypes launches per step make the eager step launch-bound
    fn = lambda: step.step_graphed(video, ref_img)
    tuned_graph = step.tune_schedule_graphed(video, ref_img)   # untimed; the order that is best for the REPLAYED step (N > 1 only)
    for _ in range(2):
        out = fn()
    ms = timed(fn, steps) / steps
    rec = {"workload": "configs[4]: train step, default model 695M, bf16 compute / fp32 master + AdamW, %d clips per GPU, "
                       "synthetic latents, random init" % TRAIN_CLIPS_PER_GPU,
           "ms_per_step": round(ms, 2), "cuda_graph": True, "ms_per_step_eager": round(ms_eager, 2),
           "clips_per_s": round(TRAIN_CLIPS_PER_GPU * world / ms * 1e3, 3),
           "tflops_per_gpu": round(TRAIN_TFLOP_PER_CLIP * TRAIN_CLIPS_PER_GPU / ms * 1e3, 1),
           "frac_of_sustained_peak": round(TRAIN_TFLOP_PER_CLIP * TRAIN_CLIPS_PER_GPU / ms * 1e3 / _peaks()["tf_sustained"], 4),
           "loss": round(float(out["loss"]), 5), "grad_norm": round(float(out["grad_norm"]), 4),
           "grad_bytes": step.arena.numel * 4, "allreduce_ms": None, "allreduce_exposed_ms": None,
           "peak_mem_gb": round(torch.cuda.max_memory_allocated() / 2 ** 30, 2)}
    if world > 1:
        red = step.reducer
        step.reducer, step._graph = None, None
        fn()                                          # re-captured without the collective
        ms_local = timed(fn, steps) / steps           # same graphed step, gradients left un-reduced
        step.reducer, step._graph = red, None

        def only_allreduce():
            red.begin({})
            red.finish()
        only_allreduce()
        ms_ar = timed(only_allreduce, steps) / steps
        chosen = red.issue_after
        names = {0.0: "fine_grained_ddp_style", 0.9: "coarse_denoiser_then_encoders", 1.0: "all_at_the_end"}
        busbw = 2.0 * (world - 1) / world * step.arena.numel * 4 / (ms_ar * 1e6)
        rec.update(allreduce_ms=round(ms_ar, 3), allreduce_exposed_ms=round(max(ms - ms_local, 0.0), 3),
                   ms_per_step_without_collective=round(ms_local, 2), allreduce_busbw_gbs=round(busbw, 1),
                   buckets=len(red.buckets),
                   schedule="%s (issue_after=%.1f), picked by TrainStep.tune_schedule_graphed in the warm-up: every candidate order "
                            "captured and replayed, ms per step, max over ranks" % (names.get(chosen, "custom"), chosen),
                   schedule_candidates_graph_ms={names.get(c, str(c)): round(v, 3) for c, v in tuned_graph.items()},
                   schedule_candidates_eager_fwd_bwd_ms={names.get(c, str(c)): round(v, 3) for c, v in tuned.items()})
    return rec if rank == 0 else None


# ------------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import hivae_b200 as hv
    from hivae_b200 import configs, ops, synth

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.gpus != world:
        raise SystemExit("bench.py --gpus %d needs %d ranks: launch it as `python -m torch.distributed.run --nnodes=1 --nproc-per-node %d "
                         "--master-addr 127.0.0.1 --master-port P bench.py --gpus %d ...` (WORLD_SIZE is %d)"
                         % (args.gpus, args.gpus, args.gpus, args.gpus, world))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        # NCCL prints its version banner with printf on stdout when the communicator is created; stdout must carry
        # exactly one JSON line, so fd 1 points at stderr until the first collective has gone through
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            dist.all_reduce(torch.zeros(1, device=dev))
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)
    torch.set_grad_enabled(False)

    m = hv.AMDModel_New(**configs.DEFAULT)
    m.load_state_dict(synth.synth_state_dict(((k, v.shape) for k, v in m.state_dict().items()), 0))
    m.to(dev).set_compute_dtype(torch.bfloat16)
    # every rank gets its own clips (seeded by rank): batch sharding
    z_host = synth.synth_clip_latents(CLIPS_PER_GPU, seed=1 + rank).pin_memory()
    out_host = torch.empty(CLIPS_PER_GPU, 16, 4, 32, 32).pin_memory()
    z_dev = z_host.to(dev)

    # --graph: encode + the 20 Euler steps replayed from one CUDA graph (AMDModel_New.sample_graphed) instead of ~5400 ctypes
    # launches per call; same kernels, same arithmetic
    sample = m.sample_graphed if args.graph else m.sample

    def step_resident():
        video, ref_img = synth.split_clip(z_dev)
        return sample(video, ref_img, sample_step=SAMPLE_STEPS)[1]

    def step_e2e():
        z = z_host.to(dev, non_blocking=True)
        video, ref_img = synth.split_clip(z)
        zt = sample(video, ref_img, sample_step=SAMPLE_STEPS)[1]
        out_host.copy_(zt, non_blocking=True)
        return zt

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if dist is not None:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms)

    for _ in range(max(args.warmup, 3)):
        step_resident()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    n0 = ops.launch_count()
    if args.ncu_range:  # `ncu --profile-from-start off`: the launch list covers exactly the timed steps
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
    ms = timed(step_resident, args.steps)
    if args.ncu_range:
        torch.cuda.profiler.stop()
    launches = ops.launch_count() - n0
    clk = clocks.stop() if rank == 0 else None
    if args.graph:  # the replayed graph launches exactly the kernels of one eager call: count those
        n0 = ops.launch_count()
        m.sample(*synth.split_clip(z_dev), sample_step=SAMPLE_STEPS)
        launches = (ops.launch_count() - n0) * args.steps
    step_e2e()
    ms_e2e = timed(step_e2e, args.steps)

    # ---- roofline leg: one extra pass with per-launch CUDA events (not part of any reported throughput)
    roof = None
    share = {}
    if rank == 0:
        torch.cuda.synchronize()
        ops.PROFILE = []
        m.sample(*synth.split_clip(z_dev), sample_step=SAMPLE_STEPS)   # eager: per-launch CUDA events need the ctypes launches
        torch.cuda.synchronize()
        summ = ops.profile_summary(ops.PROFILE)
        ops.PROFILE = None
        tot = sum(d["ms"] for d in summ.values()) or 1.0
        share = {k: {"launches": d["launches"], "ms": round(d["ms"], 3), "share": round(d["ms"] / tot, 4),
                     "tflops": round(d["flops"] / d["ms"] / 1e9, 1) if d["ms"] > 0 else 0.0,
                     "gbs": round(d["bytes"] / d["ms"] / 1e6, 1) if d["ms"] > 0 else 0.0} for k, d in sorted(summ.items())}
        g = summ.get("gemm_tc")
        pk = _peaks()
        if g and g["ms"] > 0:
            ach = g["flops"] / g["ms"] / 1e9  # TFLOP/s
            roof = {"kernel": "gemm_tc_kernel (tcgen05 bf16 GEMM + fused epilogues)", "bound": "tensor",
                    "achieved": round(ach, 1), "peak": pk["tf_sustained"], "unit": "TFLOP/s", "frac": round(ach / pk["tf_sustained"], 4),
                    "peak_source": pk["src"] + " sustained bf16 (kernel timed inside a long step)",
                    "launches": g["launches"], "avg_launch_ms": round(g["ms"] / g["launches"], 4),
                    "share_of_step": round(g["ms"] / tot, 4), "traffic": None}
            try:  # DRAM bytes of one launch from the committed ncu --set full capture (not measurable live)
                with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "ncu_traffic.json")) as f:
                    t = json.load(f)["gemm_tc_kernel"]
                roof.update(traffic=t["dram_bytes_per_launch"], traffic_launch=t["launch"], traffic_algorithmic=t["algorithmic_bytes"])
            except (OSError, KeyError, ValueError):
                pass

    secondary = None
    if rank == 0 and not args.no_secondary:
        secondary = _secondary(torch, dev, share)
    train = None if args.no_train else _train_leg(torch, dist, dev, m, rank, world, timed)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        t_enc, t_step, t_clip = _oracle_clip_time(cores)
        cpu = {"value": 1.0 / t_clip, "unit": "clips/s", "cores": cores, "kind": "port",
               "sample": "1 clip: encode (%.2f s) + 1 of %d Euler steps (%.2f s) of the default model on the CPU oracle "
                         "(fp32, %d torch threads), extrapolated to %d identical steps" % (t_enc, SAMPLE_STEPS, t_step, cores, SAMPLE_STEPS)}

    if rank == 0:
        clips = CLIPS_PER_GPU * world * args.steps
        value = clips / (ms / 1e3)
        line = {
            "metric": "clips_per_sec_encode_decode_17x256x256", "value": value, "unit": "clips/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "config": dict(_config(world), cuda_graph=bool(args.graph)),
            "tflops_per_gpu": round(TFLOP_PER_CLIP * CLIPS_PER_GPU * args.steps / (ms / 1e3) , 1),
            "e2e": {"value": clips / (ms_e2e / 1e3), "unit": "clips/s",
                    "h2d_bytes_per_step": z_host.numel() * 4, "d2h_bytes_per_step": out_host.numel() * 4},
            "gpu_launches": int(launches), "clocks": clk, "roofline": roof, "cpu_baseline": cpu, "secondary": secondary,
            "train": train, "kernels": share,
        }
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-train", action="store_true", help="skip the train-step sub-record (BASELINE configs[4])")
    ap.add_argument("--no-secondary", action="store_true", help="skip the attention tensor-pipe %% / DWT HBM %% legs")
    ap.add_argument("--ncu-range", action="store_true", help="cudaProfilerStart/Stop around the timed steps")
    ap.add_argument("--no-graph", dest="graph", action="store_false",
                    help="eager sample() (~5400 ctypes launches per call) instead of the default replay of encode + Euler loop from one CUDA graph "
                         "(AMDModel_New.sample_graphed: same kernels, bit-identical output)")
    ap.add_argument("--sample-steps", type=int, default=None, help="Euler steps per clip (default 20 = configs[1]; 10/25/50 = configs[2] sweep)")
    ap.add_argument("--clips-per-gpu", type=int, default=None, help="clips per GPU (default 8 = configs[1]; configs[2] uses 32 / n_gpus)")
    args = ap.parse_args()
    global SAMPLE_STEPS, CLIPS_PER_GPU, TFLOP_PER_CLIP
    if args.sample_steps:
        SAMPLE_STEPS = args.sample_steps
    if args.clips_per_gpu:
        CLIPS_PER_GPU = args.clips_per_gpu
    TFLOP_PER_CLIP = 0.475 + SAMPLE_STEPS * 4.02
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
