"""micro-benchmark of hv_linear_bf16 at the DiT's GEMM shapes (batch 8): TFLOP/s per epilogue, cuBLAS beside it"""
import sys, os, math, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hivae_b200 import ops
torch.set_grad_enabled(False)
dev = torch.device("cuda:0")
M = int(os.environ.get("GM", 35072))
D = 1024
g = torch.Generator().manual_seed(0)
def mk(r, c): return (torch.randn(r, c, generator=g) / math.sqrt(c)).bfloat16().to(dev)
reps = int(os.environ.get("REPS", 10))
def timeit(fn):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
frames, L = M // 274, 274
res = torch.randn(frames, L, D, device=dev)
table = torch.randn(frames, 6 * D, device=dev)
rows = ops.Rows(res, 0, L, 18, 256)
ln = lambda n: (torch.ones(n, device=dev), torch.zeros(n, device=dev))
results = {}
only = sys.argv[1] if len(sys.argv) > 1 else None
for name, N, K in (("qkv", 3 * D, D), ("out", D, D), ("ff1", 4 * D, D), ("ff2", D, 4 * D)):
    if only and name != only: continue
    A, W, b = mk(M, K), mk(N, K), torch.randn(N, device=dev)
    fl = 2.0 * M * N * K
    cases = {}
    outb = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    cases["store_bf16"] = lambda: ops.linear(A, W, ops.epi_store(outb, b))
    if name == "qkv":
        qw, qb = ln(64)
        cases["qknorm"] = lambda: ops.linear(A, W, ops.epi_qknorm(outb, b, qw, qb, qw, qb, D, D))
    if name == "ff1":
        cases["gelu"] = lambda: ops.linear(A, W, ops.epi_store(outb, b, gelu=True))
    if N == D:
        cases["residual_gated"] = lambda: ops.linear(A, W, ops.epi_residual(rows, b, ops.Ada(table, 1, 1), 2 * D, 5 * D))
    cases["cublas"] = lambda: torch.matmul(A, W.t(), out=outb)
    for cname, fn in cases.items():
        ms = timeit(fn)
        results["%s/%s" % (name, cname)] = round(fl / ms / 1e9, 1)
        print("%-4s M=%d N=%d K=%d %-15s %8.3f ms %8.1f TFLOP/s" % (name, M, N, K, cname, ms, fl / ms / 1e9), flush=True)
print(json.dumps(results))
