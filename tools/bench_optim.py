"""train-step tail at the default model's size (695 M fp32 parameters): GB/s of the fused clip + AdamW kernel
(28 B/parameter), the gradient-norm reduction (4 B/parameter) and the mse loss (8 B/element)."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hivae_b200 import optim
dev = torch.device("cuda:0")
N = int(float(os.environ.get("NPARAM", 695e6)))
p = torch.nn.Parameter(torch.randn(N, device=dev) * 0.02)
arena = optim.FlatArena([p])
arena.grad.normal_()
opt = optim.AdamW(arena, lr=1e-4)
res = {}
def timeit(fn, nbytes, name, reps=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    res[name] = round(nbytes / ms / 1e6, 1)
    print("%-34s %8.3f ms %8.1f GB/s" % (name, ms, nbytes / ms / 1e6), flush=True)
timeit(lambda: optim.clip_grad_norm_(arena, 1.0), 4.0 * N, "grad_norm_clip_coef 695M")
def step():
    optim.clip_grad_norm_(arena, 1.0); opt.step()
timeit(opt.step, 28.0 * N, "adamw_clip 695M (28 B/param)")
timeit(step, 32.0 * N, "clip_grad_norm_ + AdamW.step")
a, b = torch.randn(8 * 16 * 4 * 32 * 32, device=dev), torch.randn(8 * 16 * 4 * 32 * 32, device=dev)
timeit(lambda: optim.mse_loss(a, b), 8.0 * a.numel(), "mse_loss 8 clips", reps=20)
print(json.dumps(res))
