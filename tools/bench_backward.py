"""Backward building blocks at the DiT shapes: dgrad / wgrad on the tcgen05 GEMM (incl. the operand transposes),
bias-gradient column sums, LayerNorm and GELU backward.  CUDA events, 10 reps after warm-up."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hivae_b200 import backward as B
dev = torch.device("cuda:0")
torch.manual_seed(0)


def timeit(fn, reps=10):
    fn(); fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


M = 35072
for name, N, K in (("qkv", 3072, 1024), ("out", 1024, 1024), ("ff1", 4096, 1024), ("ff2", 1024, 4096)):
    x = (torch.randn(M, K, device=dev) * 0.5).to(torch.bfloat16)
    w = (torch.randn(N, K, device=dev) * 0.05).to(torch.bfloat16)
    dy = (torch.randn(M, N, device=dev) * 0.1).to(torch.bfloat16)
    ms = timeit(lambda: B.linear_backward(dy, x, w))
    print("%-4s M=%d N=%d K=%d  linear_backward (dX + dW + db, transposes included) %8.3f ms  %7.1f TFLOP/s" % (name, M, N, K, ms, 4.0 * M * N * K / ms / 1e9), flush=True)
    ms = timeit(lambda: B.transpose2d(dy))
    print("     transpose dY (%d x %d bf16)  %8.3f ms  %7.1f GB/s" % (M, N, ms, 4.0 * M * N / ms / 1e6), flush=True)
    ms = timeit(lambda: B.colsum(dy))
    print("     colsum dY                    %8.3f ms  %7.1f GB/s" % (ms, 2.0 * M * N / ms / 1e6), flush=True)
x = torch.randn(M, 1024, device=dev)
dyf = torch.randn(M, 1024, device=dev).to(torch.bfloat16)
w = torch.ones(1024, device=dev)
ms = timeit(lambda: B.layernorm_backward(x, dyf, w, 1e-5))
print("layernorm_backward 35072x1024 (x fp32, dy bf16 -> dx fp32)  %8.3f ms  %7.1f GB/s" % (ms, M * 1024 * 10.0 / ms / 1e6))
pre = torch.randn(M, 4096, device=dev).to(torch.bfloat16)
d = torch.randn(M, 4096, device=dev).to(torch.bfloat16)
ms = timeit(lambda: B.gelu_tanh_backward(pre, d))
print("gelu_tanh_backward 35072x4096 bf16                           %8.3f ms  %7.1f GB/s" % (ms, M * 4096 * 6.0 / ms / 1e6))
