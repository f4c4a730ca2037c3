"""freq_split3d alone, HBM-clean: `copies` distinct (n, 4, 32, 32, 32) inputs totalling several times the L2 are processed back
to back inside one CUDA graph; GB/s = algorithmic bytes (read x, write low and high = 12 B per element) / time per call.
The tensor-core kernel (default) and the SIMT kernels (HV_FREQ_TC=0) are timed in the same process on the same inputs."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hivae_b200.frequency_utils import freq_split_gauss

torch.set_grad_enabled(False)
dev = torch.device("cuda:0")
peak = 6550.4
try:
    peak = float(json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass


def time_graph(xs, reps=5):
    for x in xs[:2]:
        freq_split_gauss(x, 0.6, 0.6)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for x in xs:
            freq_split_gauss(x, 0.6, 0.6)
    g.replay(); torch.cuda.synchronize()
    tot = 0.0
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / reps / len(xs)


res = {}
for shape in ((8, 4, 32, 32, 32), (32, 4, 32, 32, 32), (128, 4, 32, 32, 32), (32, 4, 16, 32, 32), (128, 4, 8, 16, 16)):
    numel = 1
    for d in shape:
        numel *= d
    copies = min(64, max(4, int(4 * 126e6 / (12 * numel)) + 1))
    xs = [torch.randn(*shape, device=dev) for _ in range(copies)]
    row = {}
    for name, env in (("tensor", "1"), ("simt", "0"), ("default", None)):
        if env is None:
            os.environ.pop("HV_FREQ_TC", None)
        else:
            os.environ["HV_FREQ_TC"] = env
        ms = time_graph(xs)
        row[name] = round(12 * numel / ms / 1e6, 1)
    os.environ["HV_FREQ_TC"] = "1"
    lo_t, hi_t = freq_split_gauss(xs[0], 0.6, 0.6)
    os.environ["HV_FREQ_TC"] = "0"
    lo_s, hi_s = freq_split_gauss(xs[0], 0.6, 0.6)
    os.environ.pop("HV_FREQ_TC")
    row["max_abs_diff_low"] = float((lo_t - lo_s).abs().max())
    row["pct_of_hbm_tensor"] = round(100 * row["tensor"] / peak, 1)
    row["pct_of_hbm_default"] = round(100 * row["default"] / peak, 1)
    res["x".join(map(str, shape))] = row
    print("freq_split3d %-16s tensor %7.1f GB/s (%4.1f %% of HBM)   simt %7.1f GB/s   default dispatch %7.1f GB/s   |low_t - low_s| %.2e  (%d instances per graph)"
          % ("x".join(map(str, shape)), row["tensor"], row["pct_of_hbm_tensor"], row["simt"], row["default"], row["max_abs_diff_low"], copies), flush=True)
    del xs
print(json.dumps(res))
