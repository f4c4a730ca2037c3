"""GPU tier (-m gpu): the sm_100a kernels, called through the C ABI (ctypes -> libhivae_b200.so),
against the CPU oracle (oracle/amd_oracle.py) and the committed golden fixtures generated from the
unmodified reference (tests/golden/make_golden.py).

Tolerances (BASELINE.json north_star): DWT coefficients 1e-6 (fp32), latents / reconstructions 1e-4
relative in fp32 and 2e-2 in bf16, reconstruction PSNR within 0.05 dB.
"""
import math

import pytest
import torch

from oracle import amd_oracle as O
from hivae_b200 import configs, synth

pytestmark = pytest.mark.gpu
torch.set_grad_enabled(False)

TOL_DWT = 1e-6
TOL_F32 = 1e-4
TOL_BF16 = 2e-2


def rel(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return float((a - b).abs().max() / (b.abs().max() + 1e-12))


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available(), "the gpu tier needs a CUDA device"
    from hivae_b200 import _lib
    _lib.require_device()
    return torch.device("cuda:0")


def _gen(seed):
    return torch.Generator().manual_seed(seed)


# =============================================================================== wavelet
def test_dwt_idwt_golden(dev, golden):
    import hivae_b200 as hv
    g = golden("signal_ops")
    outs = hv.dwt_init(g["dwt_x"].to(dev))
    for name, o in zip(("LL", "HL", "LH", "HH"), outs):
        assert rel(o, g["dwt_" + name]) <= TOL_DWT, name
        assert torch.equal(o.cpu(), g["dwt_" + name]), name  # same fp32 operation order as wavelet.py:15-24
    rec = hv.iwt_init(torch.cat(outs, 0))
    assert rec.dtype == torch.float32
    assert rel(rec, g["iwt_out"]) <= TOL_DWT
    # module API
    assert torch.equal(hv.DWT()(g["dwt_x"].to(dev))[0], outs[0])
    assert torch.equal(hv.IWT()(torch.cat(outs, 0)), rec)


@pytest.mark.parametrize("shape", [(1, 1, 2, 2), (2, 3, 6, 10), (3, 2, 16, 24), (1, 3, 64, 34), (5, 1, 2, 258), (2, 2, 130, 8)])
def test_dwt_ragged_shapes_vs_oracle(dev, shape):
    import hivae_b200 as hv
    x = torch.randn(*shape, generator=_gen(sum(shape)))
    outs = hv.dwt_init(x.to(dev))
    for o, r in zip(outs, O.haar_dwt2(x)):
        assert rel(o, r) <= TOL_DWT
    rec = hv.iwt_init(torch.cat(outs, 0))
    assert rel(rec, O.haar_idwt2(torch.cat(O.haar_dwt2(x), 0))) <= TOL_DWT
    assert rel(rec, x) <= 1e-6


def test_dwt_edge_cases(dev):
    import hivae_b200 as hv
    e = hv.dwt_init(torch.empty(0, 3, 8, 8, device=dev))
    assert all(t.shape == (0, 3, 4, 4) for t in e)
    assert hv.iwt_init(torch.empty(0, 3, 4, 4, device=dev)).shape == (0, 3, 8, 8)
    with pytest.raises(RuntimeError):
        hv.dwt_init(torch.randn(1, 1, 5, 4, device=dev))  # odd H: the reference's strided slices would mis-broadcast
    with pytest.raises(TypeError):
        hv.dwt_init(torch.randn(1, 1, 4, 4, device=dev).half())
    # bf16 in -> bf16 sub-bands (dtype-preserving, wavelet.py:13-27), IWT always fp32 (wavelet.py:40-41)
    xb = torch.randn(2, 3, 32, 48, generator=_gen(3)).bfloat16()
    outs = hv.dwt_init(xb.to(dev))
    assert all(o.dtype == torch.bfloat16 for o in outs)
    for o, r in zip(outs, O.haar_dwt2(xb)):  # the reference evaluates every add/sub in bf16: bit-exact
        assert torch.equal(o.cpu(), r)
    rec = hv.iwt_init(torch.cat(outs, 0))
    assert rec.dtype == torch.float32
    assert torch.equal(rec.cpu(), O.haar_idwt2(torch.cat([o.cpu() for o in outs], 0)))


@pytest.mark.parametrize("T,H,levels", [(17, 256, 2), (33, 512, 3)])
def test_dwt_pyramid_full_size_roundtrip(dev, T, H, levels):
    """BASELINE config 4 sizes: size-independent properties (perfect reconstruction, energy
    conservation of the orthonormal Haar transform) + a spot comparison against the oracle."""
    import hivae_b200 as hv
    x = torch.randn(T, 3, H, H, generator=_gen(0))
    xd = x.to(dev)
    bands, ll = hv.dwt_pyramid(xd, levels)
    assert ll.shape == (T, 3, H >> levels, H >> levels)
    rec = hv.iwt_pyramid(bands, ll)
    assert rel(rec, x) <= 1e-6
    e_in = float((xd.double() ** 2).sum())
    e_out = float((ll.double() ** 2).sum()) + sum(float((b.double() ** 2).sum()) for lv in bands for b in lv)
    assert abs(e_in - e_out) <= 1e-6 * e_in
    ob, oll = O.haar_pyramid(x[:2], levels)
    assert rel(ll[:2], oll) <= TOL_DWT
    for lv, olv in zip(bands, ob):
        for b, r in zip(lv, olv):
            assert rel(b[:2], r) <= TOL_DWT


@pytest.mark.parametrize("shape,levels", [((3, 2, 64, 40), 2), ((33, 3, 512, 512), 3), ((2, 1, 24, 8), 3)])
def test_bf16_pyramid_inverse_mixed_dtypes(dev, shape, levels):
    """bf16 pyramid: level 1 of the inverse sees four bf16 bands, later levels an fp32 LL (iwt_init always
    returns fp32) with bf16 detail bands; the reference's torch.cat promotes -> bit-exact vs the oracle."""
    import hivae_b200 as hv
    x = torch.randn(*shape, generator=_gen(3)).bfloat16()
    bands, ll = hv.dwt_pyramid(x.to(dev), levels)
    ob, oll = O.haar_pyramid(x, levels)
    assert torch.equal(ll.cpu(), oll)
    rec = hv.iwt_pyramid(bands, ll)
    assert rec.dtype == torch.float32
    assert torch.equal(rec.cpu(), O.haar_pyramid_inverse(ob, oll))


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("shape,levels", [((2, 3, 16, 24), 1), ((17, 3, 256, 256), 2), ((3, 1, 64, 136), 3), ((2, 2, 20, 12), 2)])
def test_fused_pyramid_equals_per_level(dev, dtype, shape, levels):
    """one-launch pyramid kernels (hv_dwt2_haar_pyramid / hv_idwt2_haar_pyramid) are bit-identical to the per-level
    recursion and to the oracle; the last shape (W % 8 != 0) exercises the per-level fallback of the wrapper"""
    import hivae_b200 as hv
    from hivae_b200 import ops
    x = torch.randn(*shape, generator=_gen(5)).to(dtype)
    xd = x.to(dev)
    bands, ll = hv.dwt_pyramid(xd, levels)
    pb, pll = [], xd
    for _ in range(levels):
        pll, a, b, c = ops.dwt2_haar(pll)
        pb.append((a, b, c))
    assert torch.equal(ll, pll) and all(torch.equal(u, v) for lu, lv in zip(bands, pb) for u, v in zip(lu, lv))
    ob, oll = O.haar_pyramid(x, levels)
    assert torch.equal(ll.cpu(), oll) and all(torch.equal(u.cpu(), v) for lu, lv in zip(bands, ob) for u, v in zip(lu, lv))
    rec = hv.iwt_pyramid(bands, ll)
    prec = ll
    for hl, lh, hh in reversed(bands):
        prec = ops.idwt2_haar(prec, hl, lh, hh)
    assert rec.dtype == torch.float32 and torch.equal(rec, prec)
    assert torch.equal(rec.cpu(), O.haar_pyramid_inverse(ob, oll))


# =============================================================================== frequency split
def test_freq_split_golden(dev, golden):
    import hivae_b200 as hv
    g = golden("signal_ops")
    x = g["freq_x"].to(dev)
    for d in (5, 6):
        lpf = hv.gaussian_low_pass_filter([32, 32, 32], d / 10, d / 10)
        lo, hi = hv.freq_3d_filter(x, lpf)
        assert rel(lo, g["freq_lo_%d" % d]) <= 1e-5 and rel(hi, g["freq_hi_%d" % d]) <= 1e-5
        assert rel(lo + hi, g["freq_x"]) <= 1e-6
    # ragged even sizes + anisotropic cut-offs through an UNTAGGED mask tensor (rank-1 factorisation path)
    lpf2 = O.gaussian_lpf([8, 12, 20], 0.3, 0.7)
    lo2, hi2 = hv.freq_3d_filter(g["freq2_x"].to(dev), lpf2)
    assert rel(lo2, g["freq2_lo"]) <= 1e-5 and rel(hi2, g["freq2_hi"]) <= 1e-5


def test_freq_split_properties_full_size(dev):
    import hivae_b200 as hv
    from hivae_b200.frequency_utils import freq_split_gauss
    x = torch.randn(32, 4, 32, 32, 32, generator=_gen(5)).to(dev)
    lo, hi = freq_split_gauss(x, 0.6, 0.6)
    assert rel(lo + hi, x) <= 1e-6
    # linearity and idempotence-like contraction: filtering twice == squaring the mask
    lo2, _ = freq_split_gauss(2.0 * x, 0.6, 0.6)
    assert rel(lo2, 2.0 * lo) <= 1e-6
    olo, ohi = O.freq_split3d(x[:1].cpu(), O.gaussian_lpf([32, 32, 32], 0.6, 0.6))
    assert rel(lo[:1], olo) <= 1e-5 and rel(hi[:1], ohi) <= 1e-5
    z = torch.zeros(0, 4, 32, 32, 32, device=dev)
    assert freq_split_gauss(z, 0.6, 0.6)[0].shape == z.shape
    # zero cut-off -> zero mask -> low = 0, high = x (frequency_utils.py:58-59)
    lo0, hi0 = hv.freq_3d_filter(x[:1], hv.gaussian_low_pass_filter([32, 32, 32], 0.0, 0.5))
    assert float(lo0.abs().max()) == 0.0 and torch.equal(hi0, x[:1])


@pytest.mark.parametrize("T,N", [(32, 32), (16, 32), (8, 32), (32, 16), (16, 16), (8, 16)])
def test_freq_split_tensor_core_shapes(dev, T, N, monkeypatch):
    """every instance of the 3xTF32 tensor-core kernel (T in {8, 16, 32}, H = W in {16, 32}) against the oracle's FFT split
    (frequency_utils.py:84-108) with anisotropic cut-offs, and against the SIMT kernels on the same input"""
    import hivae_b200 as hv
    x = torch.randn(3, 2, T, N, N, generator=_gen(7))
    lpf = O.gaussian_lpf([T, N, N], 0.4, 0.7)
    olo, ohi = O.freq_split3d(x, lpf)
    monkeypatch.setenv("HV_FREQ_TC", "1")       # a launch this small would otherwise stay on the SIMT kernel
    lo, hi = hv.freq_3d_filter(x.to(dev), lpf)
    assert rel(lo, olo) <= 1e-5 and rel(hi, ohi) <= 1e-5 and rel(lo + hi, x) <= 1e-6
    monkeypatch.setenv("HV_FREQ_TC", "0")
    slo, shi = hv.freq_3d_filter(x.to(dev), lpf)
    assert rel(lo, slo) <= 3e-6 and rel(hi, shi) <= 3e-6


# =============================================================================== GEMM + epilogues
def _ref_linear(A, W, b):
    y = A.double() @ W.double().t()
    return (y + b.double()) if b is not None else y


@pytest.mark.parametrize("M,N,K", [(1, 64, 64), (37, 128, 128), (128, 256, 64), (300, 384, 512), (4384, 3072, 1024),
                                   (129, 1024, 4096), (257, 192, 72), (1000, 4096, 1024)])
def test_linear_bf16_tcgen05_store(dev, M, N, K):
    from hivae_b200 import ops
    g = _gen(M + N + K)
    A = (torch.randn(M, K, generator=g)).bfloat16().to(dev)
    W = (torch.randn(N, K, generator=g) / math.sqrt(K)).bfloat16().to(dev)
    b = torch.randn(N, generator=g).to(dev)
    ref = _ref_linear(A.cpu().float(), W.cpu().float(), b.cpu())
    out = torch.full((M, N), float("nan"), device=dev)
    ops.linear(A, W, ops.epi_store(out, b))
    assert rel(out, ref) <= 1e-5  # bf16 operands are exact in fp32; fp32 accumulation order only
    outb = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    ops.linear(A, W, ops.epi_store(outb, b, gelu=True))
    assert rel(outb, torch.nn.functional.gelu(ref.float(), approximate="tanh")) <= 8e-3


def test_linear_bf16_strided_views(dev):
    """A and out as column slices of wider matrices (the q|k|v views of engine.run_block)."""
    from hivae_b200 import ops
    g = _gen(9)
    big = torch.randn(200, 3 * 128, generator=g).bfloat16().to(dev)
    W = (torch.randn(128, 128, generator=g) / 11).bfloat16().to(dev)
    outbig = torch.zeros(200, 256, device=dev)
    ops.linear(big[:, 128:256], W, ops.epi_store(outbig[:, 128:], None))
    ref = _ref_linear(big[:, 128:256].cpu().float(), W.cpu().float(), None)
    assert rel(outbig[:, 128:], ref) <= 1e-5 and float(outbig[:, :128].abs().max()) == 0.0


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_linear_qknorm_epilogue(dev, dtype):
    from hivae_b200 import ops
    g = _gen(21)
    M, D = 333, 128
    A = torch.randn(M, D, generator=g)
    W = torch.randn(3 * D, D, generator=g) / math.sqrt(D)
    b = torch.randn(3 * D, generator=g) * 0.1
    qw, qb, kw, kb = (torch.randn(64, generator=g) * 0.2 + 1, torch.randn(64, generator=g) * 0.1,
                      torch.randn(64, generator=g) * 0.2 + 1, torch.randn(64, generator=g) * 0.1)
    if dtype == torch.bfloat16:
        A, W = A.bfloat16(), W.bfloat16()
    y = _ref_linear(A.float(), W.float(), b).float()
    q = torch.nn.functional.layer_norm(y[:, :D].view(M, 2, 64), (64,), qw, qb, 1e-6).view(M, D)
    k = torch.nn.functional.layer_norm(y[:, D:2 * D].view(M, 2, 64), (64,), kw, kb, 1e-6).view(M, D)
    ref = torch.cat([q, k, y[:, 2 * D:]], 1)
    out = torch.empty(M, 3 * D, device=dev, dtype=dtype)
    d = lambda t: t.to(dev)
    ops.linear(d(A), d(W), ops.epi_qknorm(out, d(b), d(qw), d(qb), d(kw), d(kb), D, D))
    assert rel(out, ref) <= (1e-5 if dtype == torch.float32 else 8e-3)


@pytest.mark.parametrize("geom", [(6, 50, 128, 256, 3, 40, 10, 25),        # one CTA per tile, blocks straddle 2 frames
                                  (5, 300, 384, 128, 18, 274, 18, 256),    # CTA pairs (M > 256, N > 128), the DiT row layout
                                  (37, 30, 200, 64, 5, 20, 8, 10),         # L < 32: a 32-row block touches 3 frames; N % 32 != 0
                                  (3, 128, 1024, 1024, 0, 128, 0, 128)])   # frame == tile
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_linear_gated_residual_epilogue(dev, dtype, geom):
    """res[f, row_off+i, :] += gate[cond(f), seg(i)] * (A W^T + b) with image / motion gate columns
    (bf16: TMA-reduce epilogue of the tcgen05 GEMM; fp32: the SIMT GEMM's read-modify-write)."""
    from hivae_b200 import ops
    g = _gen(33)
    frames, FR, D, K, row_off, L, img_start, img_len = geom
    A = torch.randn(frames * L, K, generator=g)
    W = torch.randn(D, K, generator=g) / math.sqrt(K)
    b = torch.randn(D, generator=g) * 0.1
    res = torch.randn(frames, FR, D, generator=g)
    table = torch.randn((frames + 1) // 2, 6 * D, generator=g)  # conditioning row = (f // 2) * 1
    if dtype == torch.bfloat16:
        A, W = A.bfloat16(), W.bfloat16()
    y = _ref_linear(A.float(), W.float(), b).float().view(frames, L, D)
    ref = res.clone()
    ai = row_off + torch.arange(L)
    img = ((ai >= img_start) & (ai < img_start + img_len)).view(1, L, 1)
    tf = table[torch.arange(frames) // 2]                                    # (frames, 6D)
    gate = torch.where(img, tf[:, None, 5 * D:6 * D], tf[:, None, 2 * D:3 * D])  # (frames, L, D)
    ref[:, row_off:row_off + L] += gate * y
    resd = res.to(dev)
    rows = ops.Rows(resd, row_off, L, img_start, img_len)
    ops.linear(A.to(dev), W.to(dev), ops.epi_residual(rows, b.to(dev), ops.Ada(table.to(dev), 2, 1), 2 * D, 5 * D))
    assert rel(resd, ref) <= 1e-5
    # plain residual add (no gate)
    resd2 = res.to(dev)
    ops.linear(A.to(dev), W.to(dev), ops.epi_residual(ops.Rows(resd2, row_off, L), b.to(dev)))
    ref2 = res.clone()
    ref2[:, row_off:row_off + L] += y
    assert rel(resd2, ref2) <= 1e-5


@pytest.mark.parametrize("M,N,K", [(5, 16, 1024), (300, 512, 16), (64, 70, 33), (2, 6144, 512)])
def test_linear_f32_simt(dev, M, N, K):
    from hivae_b200 import ops
    g = _gen(M * N + K)
    A, W, b = torch.randn(M, K, generator=g), torch.randn(N, K, generator=g) / math.sqrt(K), torch.randn(N, generator=g)
    out = torch.empty(M, N, device=dev)
    ops.linear(A.to(dev), W.to(dev), ops.epi_store(out, b.to(dev)))
    assert rel(out, _ref_linear(A, W, b)) <= 1e-5


# =============================================================================== attention
def _ref_attention(q, k, v, scale):
    s = torch.einsum("bhqd,bhkd->bhqk", q.double(), k.double()) * scale
    return torch.einsum("bhqk,bhkd->bhqd", torch.softmax(s, -1), v.double())


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("B,H,Lq,Lkv", [(3, 2, 274, 274), (2, 8, 264, 264), (5, 1, 16, 16), (2, 4, 10, 37), (1, 16, 64, 128),
                                        (2, 2, 100, 49)])
def test_attention_dense(dev, dtype, B, H, Lq, Lkv):
    from hivae_b200 import ops
    g = _gen(B * H + Lq + Lkv)
    D = H * 64
    q, k, v = (torch.randn(B * Lq, D, generator=g), torch.randn(B * Lkv, D, generator=g), torch.randn(B * Lkv, D, generator=g))
    q, k, v = q.to(dtype), k.to(dtype), v.to(dtype)
    ref = _ref_attention(q.float().view(B, Lq, H, 64).transpose(1, 2), k.float().view(B, Lkv, H, 64).transpose(1, 2),
                         v.float().view(B, Lkv, H, 64).transpose(1, 2), 0.125).transpose(1, 2).reshape(B * Lq, D)
    out = torch.empty(B * Lq, D, device=dev, dtype=dtype)
    ops.attention(q.to(dev), k.to(dev), v.to(dev), out, B, H, 64, Lq, Lkv, ops.seq_map(1, Lq, 0, 1), ops.seq_map(1, Lkv, 0, 1), 0.125)
    assert rel(out, ref) <= (2e-5 if dtype == torch.float32 else 1e-2)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_attention_strided_temporal(dev, dtype):
    """'(n t) s d -> (n s) t d' regrouping expressed as a sequence map (transformer.py:2022-2028)."""
    from hivae_b200 import engine, ops
    n, T, S, H = 2, 16, 24, 2
    D = H * 64
    g = _gen(77)
    qkv = torch.randn(n * T * S, 3 * D, generator=g).to(dtype)
    x = qkv.float().view(n, T, S, 3, H, 64).permute(3, 0, 2, 4, 1, 5)  # (3, n, S, H, T, 64)
    ref = _ref_attention(x[0].reshape(n * S, H, T, 64), x[1].reshape(n * S, H, T, 64), x[2].reshape(n * S, H, T, 64), 0.125)
    ref = ref.view(n, S, H, T, 64).permute(0, 3, 1, 2, 4).reshape(n * T * S, D)
    qd = qkv.to(dev)
    out = torch.empty(n * T * S, D, device=dev, dtype=dtype)
    geom = engine.strided_geom(n, S, T)
    ops.attention(qd[:, :D], qd[:, D:2 * D], qd[:, 2 * D:], out, geom.seqs, H, 64, T, T, geom.qmap, geom.kvmap, 0.125)
    assert rel(out, ref) <= (2e-5 if dtype == torch.float32 else 1e-2)


def test_rope_attention_module(dev, golden):
    """model/attention.py Attention: RoPE / no RoPE / masked with start_pos."""
    import hivae_b200 as hv
    from hivae_b200.attention import Attention
    g = golden("signal_ops")
    att = Attention(128, 40, 4)
    att.load_state_dict(synth.synth_state_dict(((k, v.shape) for k, v in att.state_dict().items()), 3))
    att = att.to(dev)
    x = g["ropeattn_x"].to(dev)
    assert rel(att(x), g["ropeattn_rope"]) <= TOL_F32
    assert rel(att(x, inference=False), g["ropeattn_norope"]) <= TOL_F32
    assert rel(att(x, mask=torch.tril(torch.ones(24, 24)), start_pos=3), g["ropeattn_rope_mask"]) <= TOL_F32
    assert hv.__version__


# =============================================================================== transformer blocks
def _blocks(dev, dtype):
    from hivae_b200 import modules as M
    D, H = 128, 2
    blocks = {"basic": M.BasicTransformerBlock(D, H, 64, attention_bias=True),
              "cross": M.BasicCrossTransformerBlock(D, D, H, 64, attention_bias=True),
              "joint": M.AMDTransformerBlock(D, H, 64, 512, attention_bias=True),
              "dit": M.BasicDiTBlock(D, H, 64, 512, attention_bias=True)}
    for blk in blocks.values():
        blk.load_state_dict(synth.synth_state_dict(((k, v.shape) for k, v in blk.state_dict().items()), 11))
        blk.to(dev)
        blk.hv_compute_dtype = dtype
    return blocks


@pytest.mark.parametrize("dtype,tol", [(torch.float32, TOL_F32), (torch.bfloat16, TOL_BF16)])
def test_blocks_golden(dev, golden, dtype, tol):
    """the four block classes of model/modules.py against the unmodified reference's outputs"""
    g = golden("blocks")
    b = _blocks(dev, dtype)
    xh, xe, te = g["x_hid"].to(dev), g["x_enc"].to(dev), g["temb"].to(dev)
    assert rel(b["basic"](xe), g["basic_out"]) <= tol
    assert rel(b["cross"](xh, xe), g["cross_out"]) <= tol
    oh, oe = b["joint"](xh, te, xe)
    assert rel(oh, g["joint_hid"]) <= tol and rel(oe, g["joint_enc"]) <= tol
    assert rel(b["dit"](xe, te), g["dit_out"]) <= tol


def test_patch_embed_module(dev):
    from hivae_b200 import modules as M
    pe = M.PatchEmbed(2, 8, 256)
    sd = synth.synth_state_dict(((k, v.shape) for k, v in pe.state_dict().items()), 4)
    pe.load_state_dict(sd)
    x = torch.randn(3, 8, 32, 32, generator=_gen(8))
    ref = O.patch_embed({"p.proj.weight": sd["proj.weight"], "p.proj.bias": sd["proj.bias"]}, "p", x, 2)
    assert rel(pe.to(dev)(x.to(dev)), ref) <= 1e-5


# =============================================================================== models
def _small(dev, dtype=torch.float32):
    import hivae_b200 as hv
    cfg = configs.SMALL
    m = hv.AMDModel_New(**cfg)
    sd = synth.synth_state_dict(((k, v.shape) for k, v in m.state_dict().items()), 0)
    m.load_state_dict(sd)
    m.to(dev).set_compute_dtype(dtype)
    orc = O.AMDOracle(sd, cfg["enc_nhead"], cfg["diffusion_attn_num_heads"], 16, 2, cfg["motion_token_num"])
    return m, orc


def _fix_noise(m, z0):
    orig = m.scheduler.get_train_tuple
    m.scheduler.get_train_tuple = lambda z0_=None, z1=None, time_step=None: orig(z0=z0.to(z1.device), z1=z1, time_step=time_step)


@pytest.mark.parametrize("dtype,tol", [(torch.float32, TOL_F32), (torch.bfloat16, TOL_BF16)])
def test_small_model_encode_decode_golden(dev, golden, dtype, tol):
    """BASELINE.json config 1: small AMD encoder + rectified-flow decoder, one clip, encode -> decode."""
    g = golden("small_model")
    m, _ = _small(dev, dtype)
    video, ref_img = synth.split_clip(synth.synth_clip_latents(1, seed=1))
    video, ref_img = video.to(dev), ref_img.to(dev)
    cam, osrc, otgt = m.encode(video, ref_img)
    assert cam.shape == (1, 16, 16, 8) and osrc.shape == (16, 8, 16) and otgt.shape == (16, 8, 16)
    assert rel(cam, g["camera"]) <= tol and rel(osrc, g["obj_src"]) <= tol and rel(otgt, g["obj_tgt"]) <= tol
    _fix_noise(m, g["z0"])
    meta = m.decode(video, ref_img, cam, osrc, otgt, start_step=1000, sample_step=4, return_meta_info=True)
    for i, p in enumerate(meta["pre_cache"]):
        assert rel(p, g["pre_%d" % i]) <= tol, i
    assert rel(meta["sample"], g["zt"]) <= tol
    zj = meta["zj"].cpu()
    assert abs(O.psnr(meta["sample"].cpu(), zj) - O.psnr(g["zt"], zj)) <= 0.05
    # sample == encode + decode and returns (zi, zt, zj) each (n, t, c, h, w)
    zi, zt, zj2 = m.sample(video, ref_img, sample_step=4)
    assert zi.shape == zt.shape == zj2.shape == (1, 16, 4, 32, 32)
    assert rel(zt, g["zt"]) <= tol
    assert torch.equal(zi, ref_img) and torch.equal(zj2, video)


def test_small_model_forward_golden(dev, golden):
    g = golden("small_model")
    m, _ = _small(dev)
    video, ref_img = synth.split_clip(synth.synth_clip_latents(1, seed=1))
    _fix_noise(m, g["z0"])
    m.prepare_timestep = lambda batch_size, device, time_step=None: torch.full((batch_size,), 437.0, device=device)
    pre, vel, ld = m(video.to(dev), ref_img.to(dev))
    assert rel(pre, g["fwd_pre"]) <= TOL_F32 and rel(vel, g["fwd_vel"]) <= 1e-6
    assert set(ld) == {"loss", "diff_loss", "rec_loss"}
    assert abs(float(ld["loss"]) - float(g["fwd_loss"][0])) <= 1e-4 * float(g["fwd_loss"][0])
    assert abs(float(ld["rec_loss"]) - float(g["fwd_loss"][2])) <= 1e-4 * float(g["fwd_loss"][2])


def test_small_model_batch_independence(dev):
    """clips are independent along the batch dimension (the sharding axis): a batch of 3
    equals three batches of 1."""
    m, _ = _small(dev)
    z = synth.synth_clip_latents(3, seed=5)
    video, ref_img = (t.to(dev) for t in synth.split_clip(z))
    z0 = torch.randn(48, 4, 32, 32, generator=_gen(2))
    _fix_noise(m, z0)
    _, zt, _ = m.sample(video, ref_img, sample_step=2)
    for i in range(3):
        _fix_noise_i = z0[16 * i:16 * (i + 1)]
        m.scheduler = type(m.scheduler)(m.scheduler.num_step)
        _fix_noise(m, _fix_noise_i)
        _, zti, _ = m.sample(video[i:i + 1], ref_img[i:i + 1], sample_step=2)
        assert rel(zti, zt[i:i + 1]) <= 1e-5


@pytest.mark.parametrize("dtype,tol", [(torch.float32, TOL_F32), (torch.bfloat16, TOL_BF16)])
def test_object_only_variant_golden(dev, golden, dtype, tol):
    """diffusion_model_type='default', use_camera=False -> AMDDiffusionTransformerModelTempMotion
    (transformer.py:1483-1743; object joint blocks only, per-frame timesteps kept): sample + training forward
    against the unmodified reference's outputs (tests/golden/make_golden_variants.py)."""
    import hivae_b200 as hv
    g = golden("small_object_only")
    m = hv.AMDModel_New(**configs.SMALL_OBJECT_ONLY)
    m.load_state_dict(synth.synth_state_dict(((k, v.shape) for k, v in m.state_dict().items()), 0))
    m.to(dev).set_compute_dtype(dtype)
    video, ref_img = (t.to(dev) for t in synth.split_clip(synth.synth_clip_latents(1, seed=1)))
    _fix_noise(m, g["z0"])
    meta = m.sample(video, ref_img, sample_step=3, return_meta_info=True)
    for i, p in enumerate(meta["pre_cache"]):
        assert rel(p, g["pre_%d" % i]) <= tol, i
    assert rel(meta["sample"], g["zt"]) <= tol
    m.prepare_timestep = lambda batch_size, device, time_step=None: g["ts"].to(device)
    pre, vel, ld = m(video, ref_img)
    assert rel(pre, g["fwd_pre"]) <= tol
    assert abs(float(ld["loss"]) - float(g["fwd_loss"][0])) <= (1e-4 if dtype == torch.float32 else 2e-2) * float(g["fwd_loss"][0])


@pytest.mark.parametrize("dtype,tol", [(torch.float32, TOL_F32), (torch.bfloat16, TOL_BF16)])
def test_temporal_motion_encoder_golden(dev, golden, dtype, tol):
    """MotionEncoderLearnTokenTemporalTransformer (transformer.py:195-367) vs the unmodified
    reference, n = 1 and n = 2 (the reference regroups the second half of the flat (n t) batch, 340)."""
    import hivae_b200 as hv
    from test_cpu_oracle_and_host import TENC
    g = golden("temporal_encoder")
    m = hv.MotionEncoderLearnTokenTemporalTransformer(**TENC)
    m.load_state_dict(synth.synth_state_dict(((k, v.shape) for k, v in m.state_dict().items()), 3))
    m.to(dev)
    for sub in m.modules():
        sub.hv_compute_dtype = dtype
    for n in (1, 2):
        y = m(g["x%d" % n].to(dev))
        assert y.shape == (n, 8, 4, 16)
        assert rel(y, g["y%d" % n]) <= tol, n


def test_small_model_sample_cross_and_refimg_motion(dev, golden):
    """a18: AMDModel_New.sample_cross (model_AMD.py:2198-2377; camera motion of video_1, no object motion in the
    denoiser) and .sample_with_refimg_motion (2701-2785; object source motion from the high band of the grey
    reference) against the unmodified reference, fp32."""
    g = golden("small_cross_refimg")
    m, _ = _small(dev)
    v1, ref_img = (t.to(dev) for t in synth.split_clip(synth.synth_clip_latents(1, seed=1)))
    v2 = synth.split_clip(synth.synth_clip_latents(1, seed=4))[0].to(dev)
    _fix_noise(m, g["z0"])
    zi, zt, zj = m.sample_cross(v1, v2, ref_img, sample_step=3)
    assert rel(zt, g["cross_zt"]) <= TOL_F32 and torch.equal(zj, v2) and torch.equal(zi, ref_img)
    ri, rt, rj = m.sample_with_refimg_motion(ref_img, g["cam"].to(dev), g["otgt"].to(dev), sample_step=3,
                                             ref_img_grey=g["grey"].to(dev))
    assert rel(rt, g["refimg_zt"]) <= TOL_F32
    assert rj.shape == (16, 4, 32, 32) and ri.shape == rt.shape == (1, 16, 4, 32, 32)  # zj stays flat (2774-2775)
    with pytest.raises(NameError):
        m.sample_with_refimg_motion(ref_img, g["cam"].to(dev), g["otgt"].to(dev), return_meta_info=True, ref_img_grey=ref_img)


@pytest.mark.parametrize("dtype,tol", [(torch.float32, TOL_F32), (torch.bfloat16, TOL_BF16)])
def test_imgspatial_dit_golden(dev, golden, dtype, tol):
    """a17 sibling AMDDiffusionTransformerModelImgSpatial (transformer.py:2044-2356) vs the unmodified reference:
    'plus', 'decouple' camera-only, and 'decouple' with object motions (camera rows swapped for object tokens after
    layer 5) with and without a camera source."""
    import hivae_b200 as hv
    from test_cpu_oracle_and_host import DITS
    g = golden("imgspatial_dit")
    t = lambda k: g[k].to(dev)
    for mt, cases in (("plus", [("obj", None, "os", "ot")]),
                      ("decouple", [("cam", None, None, None), ("cam_obj", None, "os", "ot"), ("camsrc_obj", "cs", "os", "ot")])):
        m = hv.AMDDiffusionTransformerModelImgSpatial(motion_type=mt, **DITS)
        m.load_state_dict(synth.synth_state_dict(((k, v.shape) for k, v in m.state_dict().items()), 5))
        m.to(dev)
        for sub in m.modules():
            sub.hv_compute_dtype = dtype
        for name, cs, os_, ot in cases:
            y = m(t("ct"), t("image"), t("ts"), camera_motion_source_hidden_states=t(cs) if cs else None,
                  object_motion_source_hidden_states=t(os_) if os_ else None,
                  object_motion_target_hidden_states=t(ot) if ot else None)
            assert rel(y, g["y_%s_%s" % (mt, name)]) <= tol, (mt, name)
    # AMDDiffusionTransformerModel (transformer.py:1010-1259): joint blocks only, layers 6-7 run in both passes of 'decouple'
    kw = dict({k: v for k, v in DITS.items() if k != "motion_target_num_frame"}, num_layers=9)
    for mt, cases in (("plus", [("camsrc_obj", "cs", "os", "ot")]), ("decouple", [("cam_obj", None, "os", "ot"), ("camsrc_obj", "cs", "os", "ot")])):
        m = hv.AMDDiffusionTransformerModel(motion_type=mt, **kw)
        m.load_state_dict(synth.synth_state_dict(((k, v.shape) for k, v in m.state_dict().items()), 6))
        m.to(dev)
        for sub in m.modules():
            sub.hv_compute_dtype = dtype
        for name, cs, os_, ot in cases:
            y = m(t("ct"), t("image"), t("ts"), camera_motion_source_hidden_states=t(cs) if cs else None,
                  object_motion_source_hidden_states=t(os_), object_motion_target_hidden_states=t(ot))
            assert rel(y, g["legacy_%s_%s" % (mt, name)]) <= tol, ("legacy", mt, name)


def test_unpatchify_both_feature_orders(dev):
    """(p,q,c) order of AMDModel*.unpatchify / MAE (einsum 'nhwpqc->nchpwq', model_AMD.py:2816-2830) and
    (c,p,q) order of the DiT heads (transformer.py:2038-2041); patchify is the exact inverse. Bit-exact (pure moves)."""
    from hivae_b200 import ops
    N, h, p, c = 3, 16, 2, 4
    x = torch.randn(N, h * h, p * p * c, generator=_gen(31))
    ref0 = torch.einsum("nhwpqc->nchpwq", x.reshape(N, h, h, p, p, c)).reshape(N, c, h * p, h * p)
    ref1 = x.reshape(N, 1, h, h, c, p, p).permute(0, 1, 4, 2, 5, 3, 6).flatten(5, 6).flatten(3, 4).squeeze(1)
    y0, y1 = ops.unpatchify(x.to(dev), p, order=0), ops.unpatchify(x.to(dev), p, order=1)
    assert torch.equal(y0.cpu(), ref0) and torch.equal(y1.cpu(), ref1)
    assert torch.equal(ops.patchify(y0, p, order=0).cpu(), x) and torch.equal(ops.patchify(y1, p, order=1).cpu(), x)
    m, _ = _small(dev)
    assert torch.equal(m.unpatchify(x.to(dev), p).cpu(), ref0)
    v = torch.randn(2, 5, 4, 8, 8, generator=_gen(32))
    assert m.prepare_encoder_input(v).shape == (8, 8, 8, 8)


@pytest.mark.parametrize("name", ["temporal", "legacy"])
def test_amdmodel_golden(dev, golden, name):
    """a18 sibling AMDModel (model_AMD.py:33-594) vs the unmodified reference, fp32: training forward (per-clip
    timesteps for 'spatial', per-frame for 'default'; camera mask on the low band) and a 2-step sample.
    'temporal' = temporal motion encoders + ImgSpatial 'decouple' (7 layers: swap after layer 5);
    'legacy' = per-frame encoders + camera_down + AMDDiffusionTransformerModel 'plus'."""
    import hivae_b200 as hv
    g = golden("amd_model")
    cfg = configs.AMDMODEL_TEMPORAL if name == "temporal" else configs.AMDMODEL_LEGACY
    m = hv.AMDModel(**cfg)
    m.load_state_dict(synth.synth_state_dict(((k, v.shape) for k, v in m.state_dict().items()), 7))
    m.to(dev)
    t = lambda k: g[k].to(dev)
    _fix_noise(m, g["z0"])
    ts = g[name + "_ts"]
    per_clip = cfg["diffusion_model_type"] != "default"
    m.prepare_timestep = (lambda batch_size, device, time_step=None: ts[::4].to(device)) if per_clip else \
        (lambda batch_size, device, time_step=None: ts.to(device))
    pre, vel, ld = m(t("video"), t("ref_img"), video_grey=t("vg"), ref_img_grey=t("rg"), camera_mask=t("mask") if cfg.get("use_mask") else None)
    assert rel(pre, g[name + "_pre"]) <= TOL_F32
    assert abs(float(ld["loss"]) - float(g[name + "_loss"][0])) <= 1e-4 * float(g[name + "_loss"][0])
    assert abs(float(ld["rec_loss"]) - float(g[name + "_loss"][1])) <= 1e-4 * float(g[name + "_loss"][1])
    zi, zt, zj = m.sample(t("video"), t("ref_img"), video_grey=t("vg"), ref_img_grey=t("rg"), sample_step=2)
    assert rel(zt, g[name + "_zt"]) <= TOL_F32 and torch.equal(zj, t("video"))


@pytest.mark.parametrize("dtype,tol", [(torch.float32, TOL_F32), (torch.bfloat16, TOL_BF16)])
def test_amdmodel_camera_and_rec_golden(dev, golden, dtype, tol):
    """f3: AMDModel_Camera (model_AMD.py:894-1661; both denoisers, 'decouple' camera stream) and AMDModel_Rec (2840-3125; the
    non-diffusion AMDReconstructTransformerModel, transformer.py:703-854, need_norm_out encoders) against the unmodified
    reference: training forward + losses, 3-step sample."""
    import hivae_b200 as hv
    g = golden("amd_model_camera")
    t = lambda k: g[k].to(dev)
    ts = torch.tensor([411.0]).repeat_interleave(4)
    for dmt in ("spatial", "default"):
        m = hv.AMDModel_Camera(**dict(configs.AMDMODEL_CAMERA, diffusion_model_type=dmt))
        m.load_state_dict(synth.synth_state_dict(((k, v.shape) for k, v in m.state_dict().items()), 8))
        m.to(dev).set_compute_dtype(dtype)
        _fix_noise(m, g["z0"])
        m.prepare_timestep = lambda batch_size, device, time_step=None: (torch.tensor([411.0]) if batch_size == 1 else ts).to(device)
        kw = dict(video_grey=t("video_grey"), ref_img_grey=t("ref_img_grey"), camera_mask=t("camera_mask"))
        pre, vel, ld = m(t("video"), t("ref_img"), **kw)
        assert rel(pre, g[dmt + "_pre"]) <= tol, dmt
        assert abs(float(ld["loss"]) - float(g[dmt + "_loss"][0])) <= tol * float(g[dmt + "_loss"][0])
        zi, zt, zj = m.sample(t("video"), t("ref_img"), sample_step=3, **kw)
        assert rel(zt, g[dmt + "_zt"]) <= tol and rel(zj, g[dmt + "_zj"]) <= 1e-5 and torch.equal(zi, t("ref_img_grey"))
    gr = golden("amd_model_rec")
    m = hv.AMDModel_Rec(**configs.AMDMODEL_REC)
    m.load_state_dict(synth.synth_state_dict(((k, v.shape) for k, v in m.state_dict().items()), 9))
    m.to(dev).set_compute_dtype(dtype)
    pre, zj, ld = m(gr["video"].to(dev), gr["ref_img"].to(dev))
    assert rel(pre, gr["pre"]) <= tol and set(ld) == {"loss", "rec_loss"}
    assert abs(float(ld["loss"]) - float(gr["loss"])) <= tol * float(gr["loss"])


def test_adaln_modules_standalone(dev):
    """AMDLayerNormZero / _OneVariable / 2Condition (modules.py:881-956) and AdaLayerNorm(chunk_dim=1)
    (958-1013) as stand-alone modules against the reference formulas, fp32 1e-5."""
    import torch.nn.functional as F
    import hivae_b200 as hv
    g = _gen(51)
    B, L, L2, D, E = 5, 33, 7, 256, 64
    x, c1, c2 = (torch.randn(B, n, D, generator=g) for n in (L, L2, 3))
    temb = torch.randn(B, E, generator=g)

    def ref(m, streams):
        ch = F.linear(F.silu(temb), m.linear.weight, m.linear.bias).chunk(3 * len(streams), dim=1)
        outs = [F.layer_norm(s, (D,), m.norm.weight, m.norm.bias, m.norm.eps) * (1 + ch[3 * i + 1])[:, None] + ch[3 * i][:, None]
                for i, s in enumerate(streams)]
        return outs, [ch[3 * i + 2][:, None] for i in range(len(streams))]

    for cls, streams in ((hv.AMDLayerNormZero_OneVariable, [x]), (hv.AMDLayerNormZero, [x, c1]), (hv.AMDLayerNormZero2Condition, [x, c1, c2])):
        torch.manual_seed(3)
        m = cls(E, D)
        with torch.no_grad():
            m.norm.weight.normal_(1.0, 0.2), m.norm.bias.normal_(0.0, 0.2)
        eo, eg = ref(m, streams)
        got = m.to(dev)(*[s.to(dev) for s in streams], temb.to(dev))
        n = len(streams)
        for i in range(n):
            assert rel(got[i], eo[i].detach()) <= 1e-5 and rel(got[n + i], eg[i].detach()) <= 1e-5, (cls.__name__, i)
    torch.manual_seed(4)
    m = hv.AdaLayerNorm(embedding_dim=E, output_dim=2 * D, norm_elementwise_affine=True, chunk_dim=1)
    shift, scale = F.linear(F.silu(temb), m.linear.weight, m.linear.bias).chunk(2, dim=1)
    exp = F.layer_norm(x, (D,), m.norm.weight, m.norm.bias, m.norm.eps) * (1 + scale)[:, None] + shift[:, None]
    assert rel(m.to(dev)(x.to(dev), temb=temb.to(dev)), exp.detach()) <= 1e-5


def _default(dev, dtype):
    import hivae_b200 as hv
    m = hv.AMDModel_New(**configs.DEFAULT)
    sd = synth.synth_state_dict(((k, v.shape) for k, v in m.state_dict().items()), 0)
    m.load_state_dict(sd)
    del sd
    return m.to(dev).set_compute_dtype(dtype)


@pytest.mark.parametrize("dtype,tol", [(torch.float32, TOL_F32), (torch.bfloat16, TOL_BF16)])
def test_default_model_golden(dev, golden, dtype, tol):
    """BASELINE.json config 2 architecture (695 M parameters): encode + one DiT evaluation at
    step 600 against the unmodified reference (fp32 CPU)."""
    g = golden("default_model")
    gs = golden("small_model")
    m = _default(dev, dtype)
    video, ref_img = synth.split_clip(synth.synth_clip_latents(1, seed=1))
    video, ref_img = video.to(dev), ref_img.to(dev)
    cam, osrc, otgt = m.encode(video, ref_img)
    assert rel(cam, g["camera"]) <= tol and rel(osrc, g["obj_src"]) <= tol and rel(otgt, g["obj_tgt"]) <= tol
    zi, zj = ref_img.flatten(0, 1), video.flatten(0, 1)
    step = torch.ones(16) * 600
    zt, _ = O.rf_train_tuple(zj.cpu(), step, gs["z0"])
    # feed the GOLDEN motions so the DiT comparison does not inherit encoder error
    pre = m.diffusion_transformer(torch.cat([zi, zt.to(dev)], 1), step.to(dev),
                                  camera_motion_target_hidden_states=g["camera"].to(dev),
                                  object_motion_source_hidden_states=g["obj_src"].to(dev),
                                  object_motion_target_hidden_states=g["obj_tgt"].to(dev))
    assert rel(pre, g["dit_pre_600"]) <= tol


def rms_rel(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return float((a - b).norm() / (b.norm() + 1e-30))


@pytest.mark.parametrize("dtype,tol", [(torch.float32, TOL_F32), (torch.bfloat16, TOL_BF16)])
def test_default_model_sample20_golden(dev, golden, dtype, tol):
    """THE HEADLINE COMPUTATION end to end (BASELINE.json config 2, the one bench.py times): the default 695 M model,
    encode + 20 Euler steps (20 x 36 blocks), against the UNMODIFIED reference run on CPU
    (tests/golden/make_golden_train.py sample20; model_AMD.py:2020-2195).  Final latents and four intermediate
    predictions: max-abs error over max-abs of the reference <= 1e-4 (fp32 mode) / 2e-2 (bf16 mode), PSNR(zt, zj) within
    0.05 dB; the RMS-relative error is printed beside each figure."""
    g = golden("default_sample20")
    m = _default(dev, dtype)
    video, ref_img = synth.split_clip(synth.synth_clip_latents(1, seed=1))
    _fix_noise(m, g["z0"])
    meta = m.sample(video.to(dev), ref_img.to(dev), sample_step=20, return_meta_info=True)
    assert len(meta["pre_cache"]) == 20
    for i in (0, 6, 13, 19):
        e, r = rel(meta["pre_cache"][i], g["pre_%d" % i]), rms_rel(meta["pre_cache"][i], g["pre_%d" % i])
        print("\n  %s step %2d pre: max-abs rel %.2e  rms rel %.2e" % (dtype, i, e, r))
        assert e <= tol, (i, e)
    e, r = rel(meta["sample"], g["zt"]), rms_rel(meta["sample"], g["zt"])
    zj = meta["zj"].cpu()
    dpsnr = abs(O.psnr(meta["sample"].cpu(), zj) - float(g["psnr"]))
    print("  %s final zt: max-abs rel %.2e  rms rel %.2e  |dPSNR| %.4f dB" % (dtype, e, r, dpsnr))
    assert e <= tol and dpsnr <= 0.05


def test_kernels_ran_natively(dev):
    """the product path launched our kernels (no silent eager fallback)"""
    from hivae_b200 import ops
    n0 = ops.launch_count()
    import hivae_b200 as hv
    hv.dwt_init(torch.randn(1, 1, 4, 4, device=dev))
    assert ops.launch_count() == n0 + 1


# =============================================================================== train-step tail (train_amd.py:489-496)
def _toy_params(seed, device):
    g = _gen(seed)
    shapes = [(64, 33), (129,), (7, 5, 3), (1,), (256, 256), (1000,)]  # ragged sizes: padded arena offsets, tails
    return [torch.nn.Parameter(torch.randn(*s, generator=g).to(device)) for s in shapes]


@pytest.mark.parametrize("max_norm", [1.0, 1e9])
def test_adamw_clip_matches_torch(dev, max_norm):
    """fused clip + AdamW over a flat arena vs torch.nn.utils.clip_grad_norm_ + torch.optim.AdamW (CPU fp32),
    4 steps with fresh gradients each step; max_norm=1e9 exercises the coef-clamped-to-1 branch"""
    from hivae_b200 import optim
    ref = _toy_params(3, "cpu")
    ours = [torch.nn.Parameter(p.detach().clone().to(dev)) for p in ref]
    topt = torch.optim.AdamW(ref, lr=1e-2)  # reference defaults: betas (0.9, 0.999), eps 1e-8, weight_decay 0.01
    arena = optim.FlatArena(ours)
    opt = optim.AdamW(arena, lr=1e-2)
    assert all(p.data_ptr() >= arena.flat.data_ptr() for p in ours)
    g = _gen(4)
    for step in range(4):
        grads = [torch.randn(*p.shape, generator=g) * (3.0 if step % 2 else 0.01) for p in ref]
        for p, q, gr in zip(ref, ours, grads):
            p.grad = gr.clone()
            q.grad.copy_(gr.to(dev))
        tn = torch.nn.utils.clip_grad_norm_(ref, max_norm)
        topt.step()
        norm = optim.clip_grad_norm_(arena, max_norm)
        opt.step()
        assert abs(float(norm) - float(tn)) <= 1e-6 * float(tn)
        for p, q in zip(ref, ours):
            assert rel(q.data, p.data) <= 2e-6
    for p, st in zip(ref, arena.offsets):
        assert rel(opt.exp_avg[st:st + p.numel()].view_as(p), topt.state[p]["exp_avg"]) <= 2e-6
        assert rel(opt.exp_avg_sq[st:st + p.numel()].view_as(p), topt.state[p]["exp_avg_sq"]) <= 2e-6


def test_clip_materialize_and_mse(dev):
    from hivae_b200 import optim
    ref = _toy_params(8, "cpu")
    ours = [torch.nn.Parameter(p.detach().clone().to(dev)) for p in ref]
    arena = optim.FlatArena(ours)
    g = _gen(9)
    for p, q in zip(ref, ours):
        gr = torch.randn(*p.shape, generator=g)
        p.grad = gr.clone()
        q.grad.copy_(gr.to(dev))
    tn = torch.nn.utils.clip_grad_norm_(ref, 0.5)
    norm = optim.clip_grad_norm_(arena, 0.5, materialize=True)
    assert abs(float(norm) - float(tn)) <= 1e-6 * float(tn) and arena.clip is None
    for p, q in zip(ref, ours):
        assert rel(q.grad, p.grad) <= 1e-6
    a, b = torch.randn(8 * 16, 4, 32, 32, generator=g), torch.randn(8 * 16, 4, 32, 32, generator=g)
    assert abs(float(optim.mse_loss(a.to(dev), b.to(dev))) - float(torch.nn.functional.mse_loss(a, b))) <= 1e-6
    a, b = torch.randn(1003, generator=g), torch.randn(1003, generator=g)  # ragged tail
    assert abs(float(optim.mse_loss(a.to(dev), b.to(dev))) - float(torch.nn.functional.mse_loss(a, b))) <= 1e-6


# =============================================================================== backward building blocks
@pytest.mark.gpu
@pytest.mark.parametrize("shape,dtype", [((1000, 520), torch.bfloat16), ((77, 129), torch.float32), ((4096, 1024), torch.bfloat16), ((2501, 520), torch.bfloat16)])
def test_transpose_and_colsum(dev, shape, dtype):
    from hivae_b200 import backward as B
    x = torch.randn(*shape, generator=_gen(61)).to(dtype)
    assert torch.equal(B.transpose2d(x.to(dev)).cpu(), x.t().contiguous())       # pure data movement: bit-exact
    wide = torch.randn(shape[0], shape[1] + 8, generator=_gen(62)).to(dtype).to(dev)
    assert torch.equal(B.transpose2d(wide[:, 3:3 + shape[1]]).cpu(), wide[:, 3:3 + shape[1]].t().contiguous().cpu())  # strided view
    assert rel(B.colsum(x.to(dev)), x.double().sum(0).float()) <= 1e-5
    xt, cs = B.transpose2d_colsum(x.to(dev))       # fused kernel for aligned bf16, the two separate kernels otherwise
    assert torch.equal(xt.cpu(), x.t().contiguous()) and rel(cs, x.double().sum(0).float()) <= 1e-5


@pytest.mark.gpu
@pytest.mark.parametrize("M,N,K", [(1024, 512, 256), (4096, 1024, 1024), (2304, 3072, 1024)])
def test_linear_backward_vs_autograd(dev, M, N, K):
    """dX, dW, db of y = x W^T + b on the tcgen05 GEMM against torch autograd (fp64 on the same bf16-rounded operands);
    bf16 tolerance 2e-2 for the bf16 dX, 1e-3 for the fp32 dW / db"""
    from hivae_b200 import backward as B
    g = _gen(63)
    x = (torch.randn(M, K, generator=g) * 0.5).to(torch.bfloat16)
    w = (torch.randn(N, K, generator=g) * 0.05).to(torch.bfloat16)
    dy = (torch.randn(M, N, generator=g) * 0.1).to(torch.bfloat16)
    with torch.enable_grad():
        xr, wr, br = x.double().requires_grad_(), w.double().requires_grad_(), torch.zeros(N, dtype=torch.float64, requires_grad=True)
        (torch.nn.functional.linear(xr, wr, br) * dy.double()).sum().backward()
    dx, dw, db = B.linear_backward(dy.to(dev), x.to(dev), w.to(dev))
    assert dx.dtype == torch.bfloat16 and dw.dtype == torch.float32
    assert rel(dx, xr.grad.float()) <= TOL_BF16
    assert rel(dw, wr.grad.float()) <= 1e-3 and rel(db, br.grad.float()) <= 1e-3
    dx32, _, _ = B.linear_backward(dy.to(dev), x.to(dev), w.to(dev), dx_dtype=torch.float32)
    assert rel(dx32, xr.grad.float()) <= 1e-3


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_gelu_and_layernorm_backward_vs_autograd(dev, dtype):
    from hivae_b200 import backward as B
    g = _gen(64)
    pre = (torch.randn(333, 1024, generator=g) * 2).to(dtype)
    dy = torch.randn(333, 1024, generator=g).to(dtype)
    with torch.enable_grad():
        p = pre.double().requires_grad_()
        (torch.nn.functional.gelu(p, approximate="tanh") * dy.double()).sum().backward()
    assert rel(B.gelu_tanh_backward(pre.to(dev), dy.to(dev)), p.grad.float()) <= (1e-5 if dtype == torch.float32 else 1e-2)
    for D, rows in ((1024, 1000), (512, 77), (128, 9)):
        x = torch.randn(rows, D, generator=g) * 1.5 + 0.3
        w, b = torch.randn(D, generator=g) * 0.3 + 1.0, torch.randn(D, generator=g) * 0.1
        d = torch.randn(rows, D, generator=g).to(dtype)
        with torch.enable_grad():
            xr, wr, br = x.double().requires_grad_(), w.double().requires_grad_(), b.double().requires_grad_()
            (torch.nn.functional.layer_norm(xr, (D,), wr, br, 1e-5) * d.double()).sum().backward()
        dx, dw, db = B.layernorm_backward(x.to(dev), d.to(dev), w.to(dev), 1e-5)
        assert rel(dx, xr.grad.float()) <= 1e-5 and rel(dw, wr.grad.float()) <= 1e-5 and rel(db, br.grad.float()) <= 1e-5, (D, rows)
        base = torch.randn(rows, D, generator=g)
        acc, _, _ = B.layernorm_backward(x.to(dev), d.to(dev), w.to(dev), 1e-5, dx=base.clone().to(dev), accumulate=True)
        assert rel(acc, base + xr.grad.float()) <= 1e-5


@pytest.mark.gpu
@pytest.mark.parametrize("dtype,tol", [(torch.float32, 2e-4), (torch.bfloat16, TOL_BF16)])
@pytest.mark.parametrize("B,H,Lq,Lkv", [(3, 2, 274, 274), (2, 4, 100, 49), (5, 2, 16, 16), (2, 8, 264, 264), (2, 1, 1, 3), (1, 2, 7, 600), (1, 2, 650, 33)])
def test_attention_backward_vs_autograd(dev, dtype, tol, B, H, Lq, Lkv):
    """dq, dk, dv of the dense attention against torch autograd of the reference formula (fp64 on the same rounded inputs)"""
    from hivae_b200 import backward as Bk, engine, ops
    if dtype == torch.float32 and max(Lq, Lkv) > 400:
        pytest.skip("the fp32 (SIMT) kernels keep a whole fp32 head in shared memory: L <= ~430")
    g = _gen(71)
    D = H * 64
    q = (torch.randn(B * Lq, D, generator=g) * 0.8).to(dtype)
    k = (torch.randn(B * Lkv, D, generator=g) * 0.8).to(dtype)
    v = torch.randn(B * Lkv, D, generator=g).to(dtype)
    do = torch.randn(B * Lq, D, generator=g).to(dtype)
    with torch.enable_grad():
        qr, kr, vr = (t.double().requires_grad_() for t in (q, k, v))
        hq = qr.view(B, Lq, H, 64).transpose(1, 2)
        hk, hv_ = kr.view(B, Lkv, H, 64).transpose(1, 2), vr.view(B, Lkv, H, 64).transpose(1, 2)
        o = (torch.softmax(hq @ hk.transpose(-1, -2) * 0.125, -1) @ hv_).transpose(1, 2).reshape(B * Lq, D)
        (o * do.double()).sum().backward()
    qm, km = ops.seq_map(1, Lq, 0, 1), ops.seq_map(1, Lkv, 0, 1)
    out = o.detach().to(dtype).to(dev)   # the forward output the backward receives
    dq, dk, dv = Bk.attention_backward(q.to(dev), k.to(dev), v.to(dev), out, do.to(dev), B, H, Lq, Lkv, qm, km, 0.125)
    assert rel(dq, qr.grad.float()) <= tol and rel(dk, kr.grad.float()) <= tol and rel(dv, vr.grad.float()) <= tol


@pytest.mark.gpu
@pytest.mark.parametrize("H", [4, 3])
@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-5), (torch.bfloat16, TOL_BF16)])
def test_headnorm64_backward_vs_autograd(dev, dtype, tol, H):
    """QK-norm backward: per-head LayerNorm(64, eps 1e-6) with one affine shared by all heads, on a column slice of a wider
    matrix (q inside qkv), against torch autograd in fp64"""
    from hivae_b200 import backward as Bk
    g = _gen(81)
    T_ = 517
    qkv = (torch.randn(T_, 3 * H * 64, generator=g) * 1.3 + 0.2).to(dtype)
    dy = torch.randn(T_, H * 64, generator=g).to(dtype)
    w, b = torch.randn(64, generator=g) * 0.3 + 1.0, torch.randn(64, generator=g) * 0.1
    with torch.enable_grad():
        xr = qkv[:, H * 64:2 * H * 64].double().requires_grad_()
        wr, br = w.double().requires_grad_(), b.double().requires_grad_()
        y = torch.nn.functional.layer_norm(xr.view(T_, H, 64), (64,), wr, br, 1e-6).reshape(T_, H * 64)
        (y * dy.double()).sum().backward()
    dx, dw, db = Bk.headnorm64_backward(qkv.to(dev)[:, H * 64:2 * H * 64], dy.to(dev), w.to(dev), H)
    assert rel(dx, xr.grad.float()) <= tol
    assert rel(dw, wr.grad.float()) <= (1e-5 if dtype == torch.float32 else 1e-2) and rel(db, br.grad.float()) <= 1e-5


@pytest.mark.gpu
def test_accumulating_reductions_equal_temporary_plus_add(dev):
    """hv_colsum_acc / hv_headnorm64_bwd_acc add the reduced sums straight into their destinations (the gradient-arena slots);
    the result is the very fp32 value `dest + temporary` gives: bit-exact"""
    from hivae_b200 import backward as Bk
    g = _gen(82)
    for M, N, parts in ((2501, 3072, (1024, 1024, 1024)), (517, 520, (520,)), (4096, 1024, (512, 512)), (33, 129, (100, 29))):
        x = torch.randn(M, N, generator=g).to(torch.bfloat16).to(dev)
        cs = Bk.colsum(x)
        base = [torch.randn(n, generator=g).to(dev) for n in parts]
        want = torch.cat(base) + cs
        dests = [t.clone() for t in base]
        Bk.colsum_into(x, dests)
        assert torch.equal(torch.cat(dests), want), (M, N, parts)
        Bk.colsum_into(x, dests, accumulate=False)
        assert torch.equal(torch.cat(dests), cs)
    H, T_ = 4, 777
    for dtype in (torch.bfloat16, torch.float32):
        qkv = (torch.randn(T_, 3 * H * 64, generator=g) * 1.3).to(dtype).to(dev)
        dy = torch.randn(T_, H * 64, generator=g).to(dtype).to(dev)
        w = (torch.randn(64, generator=g) * 0.3 + 1.0).to(dev)
        dx, dw, db = Bk.headnorm64_backward(qkv[:, :H * 64], dy, w, H)
        bw, bb = torch.randn(64, generator=g).to(dev), torch.randn(64, generator=g).to(dev)
        aw, ab = bw.clone(), bb.clone()
        dx2, _, _ = Bk.headnorm64_backward(qkv[:, :H * 64], dy, w, H, dw_acc=aw, db_acc=ab)
        assert torch.equal(dx, dx2) and torch.equal(aw, bw + dw) and torch.equal(ab, bb + db)


def _ref_block(x, P, heads, enc=None):
    """modules.py:721-743 (basic) / 817-840 (cross: K/V from the un-normalised encoder_hidden_state) with the diffusers
    Attention / FeedForward of modules.py:699-719 written out (fp64 autograd reference)"""
    Fn = torch.nn.functional
    B, L, D = x.shape
    split = lambda t: t.view(B, -1, heads, 64).transpose(1, 2)
    xn = Fn.layer_norm(x, (D,), P["norm1.weight"], P["norm1.bias"], 1e-5)
    src = xn if enc is None else enc
    q = Fn.layer_norm(split(Fn.linear(xn, P["attn1.to_q.weight"], P["attn1.to_q.bias"])), (64,), P["attn1.norm_q.weight"], P["attn1.norm_q.bias"], 1e-6)
    k = Fn.layer_norm(split(Fn.linear(src, P["attn1.to_k.weight"], P["attn1.to_k.bias"])), (64,), P["attn1.norm_k.weight"], P["attn1.norm_k.bias"], 1e-6)
    v = split(Fn.linear(src, P["attn1.to_v.weight"], P["attn1.to_v.bias"]))
    o = (torch.softmax(q @ k.transpose(-1, -2) * 0.125, -1) @ v).transpose(1, 2).reshape(B, L, D)
    x = x + Fn.linear(o, P["attn1.to_out.0.weight"], P["attn1.to_out.0.bias"])
    xn = Fn.layer_norm(x, (D,), P["norm2.weight"], P["norm2.bias"], 1e-5)
    return x + Fn.linear(Fn.gelu(Fn.linear(xn, P["ff.net.0.proj.weight"], P["ff.net.0.proj.bias"]), approximate="tanh"),
                         P["ff.net.2.weight"], P["ff.net.2.bias"])


@pytest.mark.gpu
@pytest.mark.parametrize("B,L,Lk,D,cross,heads", [(4, 264, 0, 512, 0, 8), (3, 40, 0, 128, 0, 2), (6, 16, 16, 512, 512, 8), (3, 40, 24, 128, 64, 2)])
def test_block_backward_vs_autograd(dev, B, L, Lk, D, cross, heads):
    """whole-block backward of BasicTransformerBlock (object motion encoder) and BasicCrossTransformerBlock (camera motion
    encoder; Lk > 0) through the autograd.Function wrapper: output, dx, d encoder_hidden_state and every parameter
    gradient against torch autograd of the reference formula in fp64 on the same bf16-rounded GEMM weights;
    bf16 tolerance 2e-2 (max-abs relative)"""
    from hivae_b200 import backward as Bk
    from hivae_b200.modules import BasicTransformerBlock, BasicCrossTransformerBlock
    g = _gen(91)
    blk = BasicCrossTransformerBlock(D, cross, heads, 64, attention_bias=True) if Lk else BasicTransformerBlock(D, heads, 64, attention_bias=True)
    with torch.no_grad():
        for n, p in blk.named_parameters():
            if p.dim() == 2:
                p.copy_((torch.randn(p.shape, generator=g) / p.shape[1] ** 0.5).to(torch.bfloat16).float())
            elif "norm" in n and n.endswith("weight"):
                p.copy_(1.0 + 0.2 * torch.randn(p.shape, generator=g))
            else:
                p.copy_(0.1 * torch.randn(p.shape, generator=g))
    x = torch.randn(B, L, D, generator=g)
    dy = torch.randn(B, L, D, generator=g)
    enc = torch.randn(B, Lk, cross, generator=g) if Lk else None
    with torch.enable_grad():
        P = {n: p.detach().double().requires_grad_() for n, p in blk.named_parameters()}
        xr = x.double().requires_grad_()
        er = enc.double().requires_grad_() if Lk else None
        yr = _ref_block(xr, P, heads, er)
        (yr * dy.double()).sum().backward()
    blk.to(dev)
    blk.hv_compute_dtype = torch.bfloat16
    with torch.enable_grad():
        xg = x.to(dev).requires_grad_()
        eg = enc.to(dev).requires_grad_() if Lk else None
        y = Bk.cross_block_apply(blk, xg, eg) if Lk else Bk.basic_block_apply(blk, xg)
        (y * dy.to(dev)).sum().backward()
    assert rel(y, yr.detach().float()) <= TOL_BF16
    y_inf = blk(x.to(dev), enc.to(dev)) if Lk else blk(x.to(dev))
    assert rel(y_inf, y) <= 1e-6          # the training forward is the inference forward
    errs = {"dx": rel(xg.grad, xr.grad.float())}
    if Lk:
        errs["denc"] = rel(eg.grad, er.grad.float())
    for n, p in blk.named_parameters():
        assert p.grad is not None and p.grad.shape == p.shape, n
        errs[n] = rel(p.grad, P[n].grad.float())
    # softmax is invariant to a per-query shift of the scores, so sum_j dk_j = 0 in every (sequence, head): d norm_k.bias is
    # identically zero and d to_k.bias is a cancellation residue; both are judged on the scale of their q-side counterparts
    for kn, qn in (("attn1.norm_k.bias", "attn1.norm_q.bias"), ("attn1.to_k.bias", "attn1.to_q.bias")):
        pk = dict(blk.named_parameters())[kn]
        errs[kn] = float((pk.grad.cpu() - P[kn].grad.float()).abs().max() / P[qn].grad.float().abs().max())
    print({k: "%.1e" % v for k, v in errs.items()})
    assert max(errs.values()) <= TOL_BF16, errs


@pytest.mark.gpu
@pytest.mark.parametrize("dtype,tol", [(torch.float32, 2e-4), (torch.bfloat16, TOL_BF16)])
def test_attention_backward_strided_geometry(dev, dtype, tol):
    """the temporal regrouping '(n t) s d -> (n s) t d' of transformer.py:2022 as a sequence map (no data movement):
    sequence (o, s) owns the rows (o*T + t)*inner + s; gradients against fp64 autograd of the regrouped formula"""
    from hivae_b200 import backward as Bk, engine
    g = _gen(73)
    outer, inner, T_, H = 2, 5, 16, 2
    D, rows = H * 64, 2 * 16 * 5
    q, k, v, do = ((torch.randn(rows, D, generator=g) * 0.8).to(dtype) for _ in range(4))
    with torch.enable_grad():
        qr, kr, vr = (t.double().requires_grad_() for t in (q, k, v))
        grp = lambda t: t.view(outer, T_, inner, H, 64).permute(0, 2, 3, 1, 4)          # (o, s, h, t, 64)
        o = torch.softmax(grp(qr) @ grp(kr).transpose(-1, -2) * 0.125, -1) @ grp(vr)
        o = o.permute(0, 3, 1, 2, 4).reshape(rows, D)
        (o * do.double()).sum().backward()
    geom = engine.strided_geom(outer, inner, T_)
    dq, dk, dv = Bk.attention_backward(q.to(dev), k.to(dev), v.to(dev), o.detach().to(dtype).to(dev), do.to(dev),
                                       geom.seqs, H, T_, T_, geom.qmap, geom.kvmap, 0.125)
    assert rel(dq, qr.grad.float()) <= tol and rel(dk, kr.grad.float()) <= tol and rel(dv, vr.grad.float()) <= tol


@pytest.mark.parametrize("B,H,Lq,Lkv", [(3, 2, 274, 274), (2, 4, 256, 256), (5, 3, 288, 288), (2, 2, 129, 200), (3, 1, 160, 48), (1, 2, 20, 274)])
@pytest.mark.parametrize("qscale", [1.0, 12.0])
def test_attention_tc2_bound_and_exact_paths(dev, B, H, Lq, Lkv, qscale):
    """attn_tc2.cu subtracts a Cauchy-Schwarz bound |q_i| max_j |k_j| scale instead of the row maximum (softmax is invariant to
    the constant) and falls back to the exact maximum per tile when the bound could underflow: qscale 12 makes
    |q| |k| scale log2(e) ~ 138 > 56 and forces that path; both must match fp64 within the bf16 tolerance."""
    from hivae_b200 import ops
    g = _gen(B * H + Lq + Lkv)
    D = H * 64
    q = (torch.randn(B * Lq, D, generator=g) * qscale).bfloat16()
    k, v = torch.randn(B * Lkv, D, generator=g).bfloat16(), torch.randn(B * Lkv, D, generator=g).bfloat16()
    ref = _ref_attention(q.float().view(B, Lq, H, 64).transpose(1, 2), k.float().view(B, Lkv, H, 64).transpose(1, 2),
                         v.float().view(B, Lkv, H, 64).transpose(1, 2), 0.125).transpose(1, 2).reshape(B * Lq, D)
    out = torch.empty(B * Lq, D, device=dev, dtype=torch.bfloat16)
    ops.attention(q.to(dev), k.to(dev), v.to(dev), out, B, H, 64, Lq, Lkv, ops.seq_map(1, Lq, 0, 1), ops.seq_map(1, Lkv, 0, 1), 0.125)
    assert rel(out, ref) <= 1e-2


@pytest.mark.gpu
@pytest.mark.parametrize("B,H,Lq,Lkv", [(3, 2, 274, 274), (2, 4, 100, 49), (2, 8, 264, 264), (2, 2, 256, 256), (1, 2, 20, 274)])
@pytest.mark.parametrize("qscale", [0.8, 12.0])
def test_attention_forward_lse_handover_to_backward(dev, B, H, Lq, Lkv, qscale):
    """hv_attention_lse + hv_attention_bwd_lse: the forward leaves the per-row log-sum-exp in the backward's workspace and the
    tcgen05 dq kernel skips its own maximum / sum passes.  The handed-over values equal the reference log-sum-exp, and dq / dk /
    dv still match torch autograd (fp64 on the same rounded inputs); qscale 12 takes the forward's exact-maximum path."""
    from hivae_b200 import backward as Bk, ops
    g = _gen(73)
    D = H * 64
    q = (torch.randn(B * Lq, D, generator=g) * qscale).bfloat16()
    k = (torch.randn(B * Lkv, D, generator=g) * 0.8).bfloat16()
    v = torch.randn(B * Lkv, D, generator=g).bfloat16()
    do = torch.randn(B * Lq, D, generator=g).bfloat16()
    with torch.enable_grad():
        qr, kr, vr = (t.double().requires_grad_() for t in (q, k, v))
        hq = qr.view(B, Lq, H, 64).transpose(1, 2)
        hk, hv_ = kr.view(B, Lkv, H, 64).transpose(1, 2), vr.view(B, Lkv, H, 64).transpose(1, 2)
        s = hq @ hk.transpose(-1, -2) * 0.125
        o = (torch.softmax(s, -1) @ hv_).transpose(1, 2).reshape(B * Lq, D)
        (o * do.double()).sum().backward()
    qm, km = ops.seq_map(1, Lq, 0, 1), ops.seq_map(1, Lkv, 0, 1)
    qd, kd, vd = q.to(dev), k.to(dev), v.to(dev)
    out = torch.empty(B * Lq, D, device=dev, dtype=torch.bfloat16)
    stats = torch.full((B * H * Lq * 2,), float("nan"), device=dev)
    wrote = ops.attention(qd, kd, vd, out, B, H, 64, Lq, Lkv, qm, km, 0.125, lse=stats)
    assert rel(out, o.detach().float()) <= 1e-2
    if Lkv >= 48:   # the tcgen05 forward handles these shapes and must hand the values over
        assert wrote is True
        lse2 = (torch.logsumexp(s.detach(), -1) * 1.4426950408889634).float()          # (B, H, Lq), log2 domain
        got = stats.view(B, H, Lq, 2)[..., 0].cpu()
        assert float((got - lse2).abs().max()) <= 2e-3 * max(1.0, float(lse2.abs().max()))
    dq, dk, dv = Bk.attention_backward(qd, kd, vd, out, do.to(dev), B, H, Lq, Lkv, qm, km, 0.125, stats=stats, have_lse=wrote is True)
    assert rel(dq, qr.grad.float()) <= TOL_BF16 and rel(dk, kr.grad.float()) <= TOL_BF16 and rel(dv, vr.grad.float()) <= TOL_BF16
