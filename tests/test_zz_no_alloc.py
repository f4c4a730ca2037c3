"""hv_mse_loss_static: the allocation-free variant of hv_mse_loss (scratch in a static __device__ array, 1024 fp64 partial
sums = 8 KB).  Collected last on purpose (file name).  Called through the C ABI (ctypes) and
compared with the existing entry point on the same inputs, on two calls in a row; passes alone in a fresh process and after
the other GPU tests (the scratch keeps the previous call's contents, nothing has to be cleared)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available(), "the gpu tier needs a CUDA device"
    from hivae_b200 import _lib
    _lib.require_device()
    return torch.device("cuda:0")


def _p(t):
    return C.c_void_p(t.data_ptr())


def _both(lib, pre, gt):
    """(existing entry point with a caller-provided workspace, static-scratch entry point) on the same inputs"""
    st = C.c_void_p(torch.cuda.current_stream(pre.device).cuda_stream)
    ws = torch.empty(lib.hv_reduce_workspace_doubles(), dtype=torch.float64, device=pre.device)
    ref = torch.full((1,), -1.0, device=pre.device)
    out = torch.full((1,), -2.0, device=pre.device)
    assert lib.hv_mse_loss(_p(pre), _p(gt), pre.numel(), _p(ws), _p(ref), st) == 0, lib.hv_last_error().decode()
    assert lib.hv_mse_loss_static(_p(pre), _p(gt), pre.numel(), _p(out), st) == 0, lib.hv_last_error().decode()
    torch.cuda.synchronize()
    return float(ref), float(out)


def test_mse_loss_static_matches_workspace_version_two_calls_in_a_row(dev):
    from hivae_b200 import _lib
    lib = _lib.load()
    g = torch.Generator().manual_seed(5)
    # call 1: the train step's loss shape (4 clips x 16 frames x 4 x 32 x 32 latents) -> 256 partial sums;
    # call 2: a ragged, much smaller input -> the scratch still holds call 1's partials beyond the ones call 2 writes
    for n in (4 * 16 * 4 * 32 * 32, 12345):
        pre, gt = torch.randn(n, generator=g).to(dev), torch.randn(n, generator=g).to(dev)
        ref, out = _both(lib, pre, gt)
        assert out == ref, (n, out, ref)  # same launch geometry and summation order: bit-identical
        want = float(((pre.double() - gt.double()) ** 2).mean())
        assert abs(out - want) <= 1e-6 * abs(want)  # the fp32 tolerance of the optimizer-tail tests
    # the largest accepted size, then the first call again: still identical
    nmax = lib.hv_mse_loss_static_max_elements()
    assert nmax == 1024 * 256 * 4
    pre, gt = torch.randn(nmax, generator=g).to(dev), torch.randn(nmax, generator=g).to(dev)
    ref, out = _both(lib, pre, gt)
    assert out == ref


def test_mse_loss_static_allocates_nothing_and_rejects_oversize(dev):
    from hivae_b200 import _lib
    lib = _lib.load()
    n = 1 << 18
    pre, gt = torch.randn(n, device=dev), torch.randn(n, device=dev)
    out = torch.zeros(1, device=dev)
    st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    assert lib.hv_mse_loss_static(_p(pre), _p(gt), n, _p(out), st) == 0  # set-up call (symbol address lookup)
    torch.cuda.synchronize()
    free0, alloc0 = torch.cuda.mem_get_info(dev)[0], torch.cuda.memory_allocated(dev)
    for _ in range(3):
        assert lib.hv_mse_loss_static(_p(pre), _p(gt), n, _p(out), st) == 0
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(dev)[0] == free0 and torch.cuda.memory_allocated(dev) == alloc0
    assert abs(float(out) - float(((pre - gt) ** 2).mean())) <= 1e-5 * float(out)
    # beyond the static scratch: a clear error, no launch, output untouched
    big = lib.hv_mse_loss_static_max_elements() * 4
    xb = torch.zeros(big, device=dev)
    out.fill_(7.0)
    rc = lib.hv_mse_loss_static(_p(xb), _p(xb), big, _p(out), st)
    torch.cuda.synchronize()
    assert rc != 0 and float(out) == 7.0
    msg = lib.hv_last_error().decode()
    assert "hv_mse_loss_static" in msg and "1024" in msg, msg
