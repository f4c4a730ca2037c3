"""Training path of the reference's train step (train_amd.py:459-496) on the sm_100a kernels:

    pre, gt, loss_dict = amd_model(z, ref_img, ...)      -> TrainRun.forward   (AMDModel_New.forward, model_AMD.py:1844-2017)
    accelerator.backward(loss)                           -> TrainRun.backward  (hand-chained backward, no autograd graph)
    DDP bucketed gradient all-reduce                     -> optim.GradAllReducer (buckets issued as their gradients retire)
    clip_grad_norm_ / AdamW.step / zero_grad             -> optim.clip_grad_norm_ / optim.AdamW

The forward is the inference forward with the activations kept; each forward stage pushes its backward closure on a
tape, and the tape is replayed in reverse.  Everything lives on the same (frames, frame_rows, D) fp32 residual layout as
the inference path: the gradient of the residual stream is a second frames buffer that the block backwards update in
place, so the reference's torch.cat / chunk / rearrange of the token streams has no backward either.

Parameter gradients are ACCUMULATED into a "sink" {parameter name: fp32 tensor} — views of the flat gradient arena in the
fused step (TrainStep), fresh zero tensors behind the autograd.Function that makes `AMDModel_New.forward` differentiable.

Compute dtype: bf16 = the reference's `--mp bf16` autocast semantics (bf16 GEMM / attention operands, fp32 accumulation,
fp32 residual stream and LayerNorm statistics, fp32 gradients of the parameters); fp32 = the parity mode on the SIMT GEMM.
No CPU fallback: every op raises without the CUDA library."""
import torch

from . import backward as B
from . import engine, ops
from .ops import Ada, Rows


class Tape:
    """backward closures in forward order; `done` lists the parameters whose gradient is final after each closure"""

    def __init__(self):
        self.items = []

    def push(self, fn, done=()):
        self.items.append((fn, tuple(done)))

    def run(self, on_done=None):
        for fn, done in reversed(self.items):
            fn()
            if on_done is not None and done:
                on_done(done)
        self.items = []


def _names(pfx, kind, cross=False):
    n1, n2 = ("norm1.norm.", "norm2.norm.") if kind in ("joint", "dit") else ("norm1.", "norm2.")
    return n1, n2


def _block_param_names(pfx, kind):
    out = []
    ada = kind in ("joint", "dit")
    for n in ("norm1", "norm2"):
        out += [pfx + n + (".norm" if ada else "") + ".weight", pfx + n + (".norm" if ada else "") + ".bias"]
    for n in ("to_q", "to_k", "to_v", "norm_q", "norm_k", "to_out.0"):
        out += [pfx + "attn1." + n + ".weight", pfx + "attn1." + n + ".bias"]
    out += [pfx + "ff.net.0.proj.weight", pfx + "ff.net.0.proj.bias", pfx + "ff.net.2.weight", pfx + "ff.net.2.bias"]
    return out


def _ada_cols(bp):
    D = bp.D
    if bp.kind == "joint":
        o1, o2 = bp.ada_off1, bp.ada_off2
        c1 = dict(shift_mot=o1, scale_mot=o1 + D, shift_img=o1 + 3 * D, scale_img=o1 + 4 * D)
        c2 = dict(shift_mot=o2, scale_mot=o2 + D, shift_img=o2 + 3 * D, scale_img=o2 + 4 * D)
        return c1, c2, (o1 + 2 * D, o1 + 5 * D), (o2 + 2 * D, o2 + 5 * D)
    if bp.kind == "dit":
        o1, o2 = bp.ada_off1, bp.ada_off2
        c1 = dict(shift_mot=o1, scale_mot=o1 + D, shift_img=o1, scale_img=o1 + D)
        c2 = dict(shift_mot=o2, scale_mot=o2 + D, shift_img=o2, scale_img=o2 + D)
        return c1, c2, (o1 + 2 * D, o1 + 2 * D), (o2 + 2 * D, o2 + 2 * D)
    return {}, {}, (0, 0), (0, 0)


def block_train(tape, ctx, bp, rows, drows, geom, sink, pfx, tag, ada=None, dada=None, kv=None, dkv=None):
    """Training forward of one transformer block over `rows` (in place, same kernels as engine.run_block) that keeps the
    activations and pushes the block's backward.  modules.py:721-743 (basic), 817-840 (cross: K/V from the dense `kv`,
    its gradient accumulated into the fp32 `dkv`), 1065-1109 (joint), 1160-1188 (dit).
    drows addresses the gradient of the residual stream with the geometry of `rows`."""
    M, D, cd = rows.M, bp.D, ctx.dtype
    hd = D // bp.heads
    scale = 1.0 / hd ** 0.5
    gated = bp.kind in ("joint", "dit")
    if not gated:
        ada = dada = None
    c1, c2, g1, g2 = _ada_cols(bp)
    sv = lambda name, shape, dtype=None: ctx.buf("sv_%s_%s" % (tag, name), shape, dtype or cd)
    F = bp.w1.shape[0]

    x0 = sv("x0", (M, D), torch.float32)   # the rows entering norm1, written by the LayerNorm kernel itself
    xn1 = ops.ln_modulate(rows, bp.n1w, bp.n1b, bp.eps, sv("xn1", (M, D)), ada, save=x0, **c1)
    if bp.kind == "cross":
        Mk = kv.shape[0]
        q_pre, q = sv("q_pre", (M, D)), sv("q", (M, D))
        ops.linear(xn1, bp.wq, ops.epi_qknorm(q, bp.bq, bp.qn_w, bp.qn_b, None, None, D, 0, pre=q_pre))
        kv_pre, kvn = sv("kv_pre", (Mk, 2 * D)), sv("kv", (Mk, 2 * D))
        ops.linear(kv, bp.wkv, ops.epi_qknorm(kvn, bp.bkv, None, None, bp.kn_w, bp.kn_b, 0, D, pre=kv_pre))
        qv, kk, vv, k_pre = q, kvn[:, :D], kvn[:, D:], kv_pre[:, :D]
    else:
        Mk = M
        qkv_pre, qkv = sv("qkv_pre", (M, 3 * D)), sv("qkv", (M, 3 * D))
        # ONE GEMM, two outputs: the normalised q / k (and v) for the attention, the pre-norm projections for the backward
        ops.linear(xn1, bp.wqkv, ops.epi_qknorm(qkv, bp.bqkv, bp.qn_w, bp.qn_b, bp.kn_w, bp.kn_b, D, D, pre=qkv_pre))
        qv, kk, vv = qkv[:, :D], qkv[:, D:2 * D], qkv[:, 2 * D:]
        q_pre, k_pre = qkv_pre[:, :D], qkv_pre[:, D:2 * D]
    att = sv("att", (M, D))
    # the forward hands its per-row log-sum-exp to the backward where its kernel produces it (tcgen05 path): the dq kernel then
    # skips its row-maximum pass and both cross-warp exchanges
    att_stats = sv("att_stats", (geom.seqs * bp.heads * geom.Lq * 2,), torch.float32) if cd == torch.bfloat16 else None
    have_lse = ops.attention(qv, kk, vv, att, geom.seqs, bp.heads, hd, geom.Lq, geom.Lkv, geom.qmap, geom.kvmap, scale, lse=att_stats) is True
    y1 = y2 = None
    if gated:
        y1 = sv("y1", (M, D))
        ops.linear(att, bp.wo, ops.epi_store(y1, bp.bo))
        B.gate_residual(rows, y1, ada, *g1)
    else:
        ops.linear(att, bp.wo, ops.epi_residual(rows, bp.bo))
    x1 = sv("x1", (M, D), torch.float32)
    xn2 = ops.ln_modulate(rows, bp.n2w, bp.n2b, bp.eps, sv("xn2", (M, D)), ada, save=x1, **c2)
    hpre, h = sv("hpre", (M, F)), sv("h", (M, F))
    ops.linear(xn2, bp.w1, ops.epi_store(h, bp.b1, gelu=True, pre=hpre))   # GELU output + pre-GELU activations
    if gated:
        y2 = sv("y2", (M, D))
        ops.linear(h, bp.w2, ops.epi_store(y2, bp.b2))
        B.gate_residual(rows, y2, ada, *g2)
    else:
        ops.linear(h, bp.w2, ops.epi_residual(rows, bp.b2))

    nn1, nn2 = _names(pfx, bp.kind)
    g = lambda name: sink.get(pfx + name)

    def bwd():
        dY = B.gate_bwd(drows, y2, ada, g2[0], g2[1], ctx.buf("bw_dY", (M, D)), dada)
        dh = B.linear_grads(ctx, dY, h, bp.w2, g("ff.net.2.weight"), g("ff.net.2.bias"), tag="bw_w2")
        dhpre = B.gelu_tanh_backward(hpre, dh)
        dxn2 = B.linear_grads(ctx, dhpre, xn2, bp.w1, g("ff.net.0.proj.weight"), g("ff.net.0.proj.bias"), tag="bw_w1")
        B.ln_modulate_bwd(x1, dxn2, bp.n2w, bp.n2b, bp.eps, drows, ada, dada, g(nn2 + "weight"), g(nn2 + "bias"), **c2)
        dY = B.gate_bwd(drows, y1, ada, g1[0], g1[1], ctx.buf("bw_dY", (M, D)), dada)
        datt = B.linear_grads(ctx, dY, att, bp.wo, g("attn1.to_out.0.weight"), g("attn1.to_out.0.bias"), tag="bw_wo")
        cross = bp.kind == "cross"
        dproj = ctx.buf("bw_dproj", (Mk, 2 * D if cross else 3 * D))
        dqp = ctx.buf("bw_dqp", (M, D)) if cross else dproj[:, :D]
        dkp, dvv = (dproj[:, :D], dproj[:, D:]) if cross else (dproj[:, D:2 * D], dproj[:, 2 * D:])
        dq, dk, _ = B.attention_backward(qv, kk, vv, att, datt, geom.seqs, bp.heads, geom.Lq, geom.Lkv, geom.qmap, geom.kvmap,
                                         scale, dq=ctx.buf("bw_dq", (M, D)), dk=ctx.buf("bw_dk", (Mk, D)), dv=dvv,
                                         stats=att_stats, have_lse=have_lse)
        for pre_, d_, w_, out_, nm in ((q_pre, dq, bp.qn_w, dqp, "attn1.norm_q."), (k_pre, dk, bp.kn_w, dkp, "attn1.norm_k.")):
            gw, gb = g(nm + "weight"), g(nm + "bias")
            if gw is not None and gb is not None and gw.is_contiguous() and gb.is_contiguous():
                B.headnorm64_backward(pre_, d_, w_, bp.heads, out=out_, dw_acc=gw.view(-1), db_acc=gb.view(-1))  # += into the gradient slots
            else:
                _, dw_, db_ = B.headnorm64_backward(pre_, d_, w_, bp.heads, out=out_)
                if gw is not None:
                    ops.axpy(1.0, dw_, gw)
                if gb is not None:
                    ops.axpy(1.0, db_, gb)
        if cross:
            dxn1 = B.linear_grads(ctx, dqp, xn1, bp.wq, g("attn1.to_q.weight"), g("attn1.to_q.bias"), tag="bw_wq")
            B.linear_grads(ctx, dproj, kv, bp.wkv, _lst(g, "attn1.to_k.weight", "attn1.to_v.weight"),
                           _lst(g, "attn1.to_k.bias", "attn1.to_v.bias"), need_dx=dkv is not None, dx_accum=dkv, tag="bw_wkv")
        else:
            dxn1 = B.linear_grads(ctx, dproj, xn1, bp.wqkv, _lst(g, "attn1.to_q.weight", "attn1.to_k.weight", "attn1.to_v.weight"),
                                  _lst(g, "attn1.to_q.bias", "attn1.to_k.bias", "attn1.to_v.bias"), tag="bw_wqkv")
        B.ln_modulate_bwd(x0, dxn1, bp.n1w, bp.n1b, bp.eps, drows, ada, dada, g(nn1 + "weight"), g(nn1 + "bias"), **c1)

    tape.push(bwd, [n for n in _block_param_names(pfx, bp.kind) if n in sink])


def _lst(g, *names):
    ts = [g(n) for n in names]
    return None if any(t is None for t in ts) else ts


def _acc(dst, src):
    """dst (fp32 parameter gradient, any shape) += src"""
    if dst is not None:
        ops.axpy(1.0, src.contiguous().view(-1), dst.view(-1))


# ================================================================================================ encoders
def object_encoder_train(tape, enc, x, sink, pfx, G, key, mask_idx=None):
    """MotionEncoderLearnTokenTransformer.forward (transformer.py:116-162) with activations kept.
    x (N, T, C, H, W) -> motion tokens (N, T, Lm, c) fp32; the backward reads d(motion tokens) from G[key]."""
    if enc.need_norm_out:
        raise NotImplementedError("training with motion_need_norm_out=True is not implemented (false in every shipped script)")
    N, T, C, H, W = x.shape
    P = enc._prepared(x.device)
    ctx = P.ctx
    D, p = P.ew.shape[0], enc.patch_embed.patch_size
    Lm, S = P.tok.shape[0], enc.img_token_len
    assert S == (H // p) * (W // p), "img_token_len should be equal!"
    keep = S if mask_idx is None else mask_idx.shape[1]
    frames, FR = N * T, Lm + keep
    x = x.reshape(frames, C, H, W).float().contiguous()
    res = ctx.buf("tr_enc_res", (frames, FR, D), torch.float32)
    dres = ctx.buf("tr_enc_dres", (frames, FR, D), torch.float32)
    tok = engine.small_linear(ctx, P.tok, P.ew, P.eb, "tr_enc_tok")
    ops.rows_scatter(tok, 0, 1, None, Rows(res, 0, Lm))
    if mask_idx is None:
        ops.patch_embed(x, None, p, P.pw, P.pb, P.pos, None, res, FR, Lm, D)
    else:  # random_masking (transformer.py:164-192): keep the tokens mask_idx[f] of every frame, in that order
        full = ctx.buf("tr_enc_full", (frames, S, D), torch.float32)
        ops.patch_embed(x, None, p, P.pw, P.pb, P.pos, None, full, S, 0, D)
        B.rows_index(full, res[:, Lm:], mask_idx, frames, S, keep, D, S, FR)
    rows, drows, geom = Rows(res, 0, FR), Rows(dres, 0, FR), engine.dense_geom(frames, FR)

    def embed_bwd():
        dtok = B.rows_reduce(dres, D, Lm, 0, 1, frames, FR)
        dmt = B.linear_grads(ctx, dtok, P.tok, P.ew, sink.get(pfx + "motion_embed.weight"), sink.get(pfx + "motion_embed.bias"), tag="bw_me")
        _acc(sink.get(pfx + "motion_token"), dmt)
        if sink.get(pfx + "patch_embed.proj.weight") is None:
            return
        if mask_idx is None:
            dpe = ops.rows_gather(Rows(dres, Lm, S), ctx.buf("bw_dpe", (frames * S, D), torch.float32))
        else:
            dpe = ctx.buf("bw_dpe", (frames * S, D), torch.float32).zero_()
            B.rows_index(dres[:, Lm:], dpe, mask_idx, frames, S, keep, D, FR, S, scatter=True)
        X = ops.patchify(x, p, order=1).view(frames * S, -1)
        wv = sink[pfx + "patch_embed.proj.weight"]
        B.linear_grads(ctx, dpe, X, P.pw, wv.view(D, -1), sink.get(pfx + "patch_embed.proj.bias"), need_dx=False, tag="bw_pe")

    tape.push(embed_bwd, [pfx + n for n in ("motion_token", "motion_embed.weight", "motion_embed.bias", "patch_embed.proj.weight",
                                            "patch_embed.proj.bias")])
    for i, bp in enumerate(P.blocks):
        block_train(tape, ctx, bp, rows, drows, geom, sink, "%stransformer_blocks.%d." % (pfx, i), "oe%d" % i)
    mrows = Rows(res, 0, Lm)
    xf = ops.rows_gather(mrows, ctx.buf("sv_oe_xf", (frames * Lm, D), torch.float32))
    y = ops.ln_modulate(mrows, P.nfw, P.nfb, enc.norm_final.eps, ctx.buf("sv_oe_fin", (frames * Lm, D), torch.float32))
    m = engine.small_linear(ctx, y, P.ow, P.ob, "tr_enc_out").view(N, T, Lm, -1).clone()

    def head_bwd():
        dres.zero_()
        dm = G[key].reshape(frames * Lm, -1).float().contiguous()
        dy = B.linear_grads(ctx, dm, y, P.ow, sink.get(pfx + "proj_out.weight"), sink.get(pfx + "proj_out.bias"), tag="bw_po")
        B.ln_modulate_bwd(xf, dy, P.nfw, P.nfb, enc.norm_final.eps, Rows(dres, 0, Lm), dw=sink.get(pfx + "norm_final.weight"),
                          db=sink.get(pfx + "norm_final.bias"))

    tape.push(head_bwd, [pfx + n for n in ("proj_out.weight", "proj_out.bias", "norm_final.weight", "norm_final.bias")])
    return m


def camera_encoder_train(tape, enc, x, sink, pfx, G, key, dkey, mask_idx=None):
    """MotionEncoderLearnTokenOnlyTemporalTransformer.forward (transformer.py:495-556) with activations kept.
    x (N, T, C, H, W) -> (N, T, S, c); backward reads d(out) from G[key] and leaves d(x) in G[dkey] (the input comes from
    camera_down, whose parameters need it)."""
    if enc.need_norm_out:
        raise NotImplementedError("training with motion_need_norm_out=True is not implemented (false in every shipped script)")
    N, T, C, H, W = x.shape
    P = enc._prepared(x.device)
    ctx = P.ctx
    D, p = P.ew.shape[0], enc.patch_embed.patch_size
    S0, t0 = enc.img_token_len, P.tok.shape[0]
    assert S0 == (H // p) * (W // p), "img_token_len should be equal!"
    assert T <= enc.video_frames and T % t0 == 0
    rep = T // t0
    x = x.reshape(N * T, C, H, W).float().contiguous()
    S = S0 if mask_idx is None else mask_idx.shape[1]
    cd = ctx.dtype
    # K/V tokens: rows (n, t, s) = patch embed + 2-D sincos + temporal sincos (506-514); masking keeps the spatial locations
    # mask_idx[n] of every frame of clip n (558-586)
    img = ctx.buf("tr_cam_img", (N * T * S, D))
    if mask_idx is None:
        ops.patch_embed(x, None, p, P.pw, P.pb, P.pos, P.tpos[:T].contiguous(), img.view(N * T, S, D), S, 0, D)
    else:
        full = ctx.buf("tr_cam_full", (N * T, S0, D), torch.float32)
        ops.patch_embed(x, None, p, P.pw, P.pb, P.pos, P.tpos[:T].contiguous(), full, S0, 0, D)
        idx_f = mask_idx.repeat_interleave(T, 0).contiguous()
        kept = B.rows_index(full, ctx.buf("tr_cam_kept", (N * T, S, D), torch.float32), idx_f, N * T, S0, S, D, S0, S)
        ops.convert(kept, img.view(N * T, S, D))
    dimg = ctx.buf("tr_cam_dimg", (N * T * S, D), torch.float32)
    res = ctx.buf("tr_cam_res", (N * T, S, D), torch.float32)
    dres = ctx.buf("tr_cam_dres", (N * T, S, D), torch.float32)
    tok = engine.small_linear(ctx, P.tok, P.ew, P.eb, "tr_cam_tok")  # (t0, D)
    tt = ctx.buf("tr_cam_tt", (1, T, D), torch.float32)
    for f in range(t0):
        ops.rows_scatter(tok[f:f + 1], 0, 0, P.tpos[f * rep:(f + 1) * rep].contiguous(), Rows(tt, f * rep, rep))
    tab = ctx.buf("tr_cam_tab", (1, T * S, D), torch.float32)
    ops.rows_scatter(tt.view(T, D), 1, 0, None, Rows(tab.view(T, S, D), 0, S))
    ops.rows_scatter(tab.view(T * S, D), 0, 1, None, Rows(res.view(N, T * S, D), 0, T * S))
    rows, drows, geom = Rows(res, 0, S), Rows(dres, 0, S), engine.strided_geom(N, S, T)

    def embed_bwd():
        # queries: res[(n, t, s)] = tok[t // rep] + temporal[t]  ->  d tok[j] = sum over n, the rep frames of j, s
        dtok = B.rows_reduce(dres, D, t0, 0, rep * S, N, T * S, rep * S, 1)
        dmt = B.linear_grads(ctx, dtok, P.tok, P.ew, sink.get(pfx + "motion_embed.weight"), sink.get(pfx + "motion_embed.bias"), tag="bw_cme")
        _acc(sink.get(pfx + "motion_token"), dmt)
        if mask_idx is None:
            dpe = dimg
        else:
            dpe = ctx.buf("bw_cam_dpe", (N * T * S0, D), torch.float32).zero_()
            B.rows_index(dimg, dpe, idx_f, N * T, S0, S, D, S, S0, scatter=True)
        X = ops.patchify(x, p, order=1).view(N * T * S0, -1)
        wv = sink.get(pfx + "patch_embed.proj.weight")
        dX = B.linear_grads(ctx, dpe, X, P.pw, None if wv is None else wv.view(D, -1), sink.get(pfx + "patch_embed.proj.bias"), tag="bw_cpe")
        G[dkey] = ops.unpatchify(dX.view(N * T, S0, -1).float(), p, C, order=1)

    tape.push(embed_bwd, [pfx + n for n in ("motion_token", "motion_embed.weight", "motion_embed.bias", "patch_embed.proj.weight",
                                            "patch_embed.proj.bias")])
    for i, bp in enumerate(P.blocks):
        block_train(tape, ctx, bp, rows, drows, geom, sink, "%stransformer_blocks.%d." % (pfx, i), "ce%d" % i, kv=img, dkv=dimg)
    xf = ops.rows_gather(rows, ctx.buf("sv_ce_xf", (N * T * S, D), torch.float32))
    y = ops.ln_modulate(rows, P.nfw, P.nfb, enc.norm_final.eps, ctx.buf("sv_ce_fin", (N * T * S, D), torch.float32))
    m = engine.small_linear(ctx, y, P.ow, P.ob, "tr_cam_out").view(N, T, S, -1).clone()

    def head_bwd():
        dres.zero_()
        dimg.zero_()
        dm = G[key].reshape(N * T * S, -1).float().contiguous()
        dy = B.linear_grads(ctx, dm, y, P.ow, sink.get(pfx + "proj_out.weight"), sink.get(pfx + "proj_out.bias"), tag="bw_cpo")
        B.ln_modulate_bwd(xf, dy, P.nfw, P.nfb, enc.norm_final.eps, drows, dw=sink.get(pfx + "norm_final.weight"),
                          db=sink.get(pfx + "norm_final.bias"))

    tape.push(head_bwd, [pfx + n for n in ("proj_out.weight", "proj_out.bias", "norm_final.weight", "norm_final.bias")])
    return m


def camera_down_train(tape, model, x, sink, G, key):
    """camera_down (model_AMD.py:1788-1793): conv 3x3 s2 -> conv 3x3 s1 -> maxpool 2; backward reads d(out) from G[key]"""
    c0, c1 = model.camera_down[0], model.camera_down[1]
    f = lambda t: t.detach().float().contiguous()
    y0 = ops.conv3x3(x, f(c0.weight), f(c0.bias), 2)
    y1 = ops.conv3x3(y0, f(c1.weight), f(c1.bias), 1)
    out = ops.maxpool2(y1)

    def bwd():
        dy1 = B.maxpool2_backward(y1, G[key].reshape(out.shape).float().contiguous())
        dy0 = B.conv3x3_backward(y0, f(c1.weight), dy1, 1, sink.get("camera_down.1.weight"), sink.get("camera_down.1.bias"))
        B.conv3x3_backward(x, f(c0.weight), dy0, 2, sink.get("camera_down.0.weight"), sink.get("camera_down.0.bias"), need_dx=False)

    tape.push(bwd, ["camera_down.%d.%s" % (i, n) for i in (0, 1) for n in ("weight", "bias")])
    return out


# ================================================================================================ the denoiser
def dit_train(tape, dit, zi, zt, timestep, cam, osrc, otgt, sink, pfx, G):
    """AMDDiffusionTransformerModelImgSpatialTempMotion.forward (transformer.py:1911-2043; also the object-only sibling)
    with activations kept.  zi, zt (N, C, H, W); cam (n, T, Lc, dc) or None; osrc / otgt (N, l, do) or None.
    Returns pre (N, C_out, H, W); the backward reads G['dpre'] and leaves G['dcam'], G['dosrc'], G['dotgt']."""
    N, _, Hi, Wi = zi.shape
    P = dit._prepared(zi.device)
    ctx = P.ctx
    cd = ctx.dtype
    T, p = dit.target_frame, dit.image_patch_size
    n = N // T
    D = P.hw.shape[1]
    S = (Hi // p) * (Wi // p)
    l = otgt.shape[1] if osrc is not None else 0
    Lo = 2 * l + 2 if osrc is not None else 0
    Lc = cam.shape[2] if cam is not None else 0
    FR = Lo + S + Lc
    res = ctx.buf("tr_dit_res", (N, FR, D), torch.float32)
    dres = ctx.buf("tr_dit_dres", (N, FR, D), torch.float32)
    img = torch.cat([zi, zt], 1).float().contiguous()
    ops.patch_embed(zi, zt, p, P.pw, P.pb, P.pos, P.tpos, res, FR, Lo, D)
    if Lo:
        src_in, tgt_in = osrc.reshape(N * l, -1).float().contiguous(), otgt.reshape(N * l, -1).float().contiguous()
        src = engine.small_linear(ctx, src_in, P.ow, P.ob, "tr_dit_osrc")
        ops.rows_scatter(src, l, 1, P.mpos[1:1 + l], Rows(res, 1, l))
        tgt = engine.small_linear(ctx, tgt_in, P.ow, P.ob, "tr_dit_otgt")
        ops.rows_scatter(tgt, l, 1, P.mpos[l + 2:2 * l + 2], Rows(res, l + 2, l))
        ops.rows_scatter(P.stok, 0, 0, P.mpos[0:1], Rows(res, 0, 1))
        ops.rows_scatter(P.ttok, 0, 0, P.mpos[l + 1:l + 2], Rows(res, l + 1, 1))
    if Lc:
        cam_in = cam.reshape(N * Lc, -1).float().contiguous()
        ops.rows_scatter(engine.small_linear(ctx, cam_in, P.cw, P.cb, "tr_dit_cam"), Lc, 1, None, Rows(res, Lo + S, Lc))
    # timestep embedding -> temb -> every adaLN table of the step in ONE GEMM (transformer.py:1939-1945)
    ts = timestep.to(device=zi.device, dtype=torch.float32).contiguous()
    te = ops.timestep_embed(ts, D, ctx.buf("tr_dit_te", (N, D)))
    h = engine.small_linear(ctx, te, P.tw1, P.tb1, "tr_dit_te_h")
    hs = ops.silu(h, ctx.buf("tr_dit_te_hs", h.shape))
    emb = engine.small_linear(ctx, hs, P.tw2, P.tb2, "tr_dit_emb")
    semb = ops.silu(emb, ctx.buf("tr_dit_semb", emb.shape, P.ada.w.dtype))
    table = ctx.buf("tr_dit_ada", (N, P.ada.N), torch.float32)
    ops.linear(semb, P.ada.w, ops.epi_store(table, P.ada.b))
    dtable = ctx.buf("tr_dit_dada", (N, P.ada.N), torch.float32)
    ada_f, ada_s = Ada(table, 1, 1), Ada(table, T, T)

    def embed_bwd():
        # ---- adaLN table -> packed Linear -> SiLU -> TimestepEmbedding
        E = emb.shape[1]
        names = dit_ada_names(dit, pfx)
        dtc = dtable if P.ada.w.dtype == torch.float32 else ops.convert(dtable, ctx.buf("bw_dtc", dtable.shape, P.ada.w.dtype))
        ws, bs = [sink.get(nm + ".weight") for nm in names], [sink.get(nm + ".bias") for nm in names]
        dsemb = ctx.buf("bw_dsemb", (N, E), torch.float32).zero_()
        B.linear_grads(ctx, dtc, semb, P.ada.w, None if any(w is None for w in ws) else ws,
                       None if any(b is None for b in bs) else bs, dx_accum=dsemb, tag="bw_ada", splitk=1)
        demb = B.silu_backward(emb, dsemb, ctx.buf("bw_demb", (N, E), P.tw2.dtype))
        dhs = B.linear_grads(ctx, demb, hs, P.tw2, sink.get(pfx + "time_embedding.linear_2.weight"),
                             sink.get(pfx + "time_embedding.linear_2.bias"), tag="bw_t2")
        dh = B.silu_backward(h, dhs, ctx.buf("bw_dh", h.shape, P.tw1.dtype))
        B.linear_grads(ctx, dh, te, P.tw1, sink.get(pfx + "time_embedding.linear_1.weight"),
                       sink.get(pfx + "time_embedding.linear_1.bias"), need_dx=False, tag="bw_t1")
        # ---- token streams
        if Lo:
            d = ops.rows_gather(Rows(dres, 1, l), ctx.buf("bw_dosrc", (N * l, D), torch.float32))
            wq, bq = sink.get(pfx + "object_motion_patch_embed.weight"), sink.get(pfx + "object_motion_patch_embed.bias")
            G["dosrc"] = B.linear_grads(ctx, d, src_in, P.ow, wq, bq, tag="bw_os").float().view(osrc.shape).clone()
            d = ops.rows_gather(Rows(dres, l + 2, l), ctx.buf("bw_dosrc", (N * l, D), torch.float32))
            G["dotgt"] = B.linear_grads(ctx, d, tgt_in, P.ow, wq, bq, tag="bw_os").float().view(otgt.shape).clone()
            if sink.get(pfx + "source_token") is not None:
                _acc(sink[pfx + "source_token"], B.rows_reduce(dres, D, 1, 0, 0, N, FR))
                _acc(sink[pfx + "target_token"], B.rows_reduce(dres, D, 1, l + 1, 0, N, FR))
        if Lc:
            d = ops.rows_gather(Rows(dres, Lo + S, Lc), ctx.buf("bw_dcam", (N * Lc, D), torch.float32))
            G["dcam"] = B.linear_grads(ctx, d, cam_in, P.cw, sink.get(pfx + "camera_motion_patch_embed.weight"),
                                       sink.get(pfx + "camera_motion_patch_embed.bias"), tag="bw_cm").float().view(cam.shape).clone()
        wv = sink.get(pfx + "image_patch_embed.proj.weight")
        if wv is not None:
            dpe = ops.rows_gather(Rows(dres, Lo, S), ctx.buf("bw_ddpe", (N * S, D), torch.float32))
            X = ops.patchify(img, p, order=1).view(N * S, -1)
            B.linear_grads(ctx, dpe, X, P.pw, wv.view(D, -1), sink.get(pfx + "image_patch_embed.proj.bias"), need_dx=False, tag="bw_ipe")

    emb_names = [nm + sfx for nm in dit_ada_names(dit, pfx) for sfx in (".weight", ".bias")]
    emb_names += [pfx + nm for nm in ("time_embedding.linear_1.weight", "time_embedding.linear_1.bias", "time_embedding.linear_2.weight",
                                      "time_embedding.linear_2.bias", "object_motion_patch_embed.weight", "object_motion_patch_embed.bias",
                                      "source_token", "target_token", "camera_motion_patch_embed.weight", "camera_motion_patch_embed.bias",
                                      "image_patch_embed.proj.weight", "image_patch_embed.proj.bias")]
    tape.push(embed_bwd, [nm for nm in emb_names if nm in sink])

    obj_rows, dobj = (Rows(r, 0, Lo + S, img_start=Lo, img_len=S) for r in (res, dres))
    cam_rows, dcam = (Rows(r, Lo, S + Lc, img_start=Lo, img_len=S) for r in (res, dres))
    img_rows, dimg = (Rows(r, Lo, S, img_start=Lo, img_len=S) for r in (res, dres))
    g_obj, g_cam, g_tmp = engine.dense_geom(N, Lo + S), engine.dense_geom(N, S + Lc), engine.strided_geom(n, S, T)
    for i, (bo, bc, bs) in enumerate(zip(P.obj, P.cam, P.sp)):
        if Lo:
            block_train(tape, ctx, bo, obj_rows, dobj, g_obj, sink, "%sobject_transformer_blocks.%d." % (pfx, i), "do%d" % i, ada_f, dtable)
        if Lc:
            block_train(tape, ctx, bc, cam_rows, dcam, g_cam, sink, "%scamera_transformer_blocks.%d." % (pfx, i), "dc%d" % i, ada_f, dtable)
        if bs is not None:
            block_train(tape, ctx, bs, img_rows, dimg, g_tmp, sink, "%sspatial_blocks.%d." % (pfx, i), "ds%d" % i, ada_s, dtable)
    # head: norm_final -> AdaLayerNorm (shift, scale) -> Linear(D -> C p p) -> unpatchify with (c, p, q) features (2032-2041)
    off = P.ada.extra_off[0]
    Co = dit.out_channels
    xf = ops.rows_gather(img_rows, ctx.buf("sv_dit_xf", (N * S, D), torch.float32))
    n1 = ops.ln_modulate(img_rows, P.nfw, P.nfb, dit.norm_final.eps, ctx.buf("sv_dit_n1", (N * S, D), torch.float32))
    n1rows = Rows(n1.view(N, S, D), 0, S)
    hc = dict(shift_mot=off, scale_mot=off + D, shift_img=off, scale_img=off + D)
    n2 = ops.ln_modulate(n1rows, P.now, P.nob, dit.norm_out.norm.eps, ctx.buf("sv_dit_n2", (N * S, D), torch.float32), ada_f, **hc)
    tokens = engine.small_linear(ctx, n2, P.hw, P.hb, "tr_dit_tok")
    pre = ops.unpatchify(tokens.view(N, S, -1), p, Co, order=1)

    def head_bwd():
        dres.zero_()
        dtable.zero_()
        dtok = ops.patchify(G["dpre"].float().contiguous(), p, order=1).view(N * S, -1)
        dn2 = B.linear_grads(ctx, dtok, n2, P.hw, sink.get(pfx + "proj_out.weight"), sink.get(pfx + "proj_out.bias"), tag="bw_ho")
        dn1 = ctx.buf("bw_dn1", (N, S, D), torch.float32).zero_()
        B.ln_modulate_bwd(n1, dn2, P.now, P.nob, dit.norm_out.norm.eps, Rows(dn1, 0, S), ada_f, dtable,
                          sink.get(pfx + "norm_out.norm.weight"), sink.get(pfx + "norm_out.norm.bias"), **hc)
        B.ln_modulate_bwd(xf, dn1.view(N * S, D), P.nfw, P.nfb, dit.norm_final.eps, dimg, dw=sink.get(pfx + "norm_final.weight"),
                          db=sink.get(pfx + "norm_final.bias"))

    tape.push(head_bwd, [pfx + nm for nm in ("proj_out.weight", "proj_out.bias", "norm_out.norm.weight", "norm_out.norm.bias",
                                             "norm_final.weight", "norm_final.bias") if pfx + nm in sink])
    return pre


def dit_ada_names(dit, pfx):
    """the adaLN Linear layers in the column order of the packed table (engine.AdaPack built in the denoiser's _prepare)"""
    names = []
    nl = len(dit.spatial_blocks) if dit._has_temporal_blocks else len(dit.object_transformer_blocks)
    for i in range(nl):
        if dit.use_object:
            names += ["%sobject_transformer_blocks.%d.norm%d.linear" % (pfx, i, k) for k in (1, 2)]
        if dit.use_camera:
            names += ["%scamera_transformer_blocks.%d.norm%d.linear" % (pfx, i, k) for k in (1, 2)]
        if dit._has_temporal_blocks:
            names += ["%sspatial_blocks.%d.norm%d.linear" % (pfx, i, k) for k in (1, 2)]
    return names + [pfx + "norm_out.linear"]


# ================================================================================================ the model
class TrainRun:
    """One differentiable evaluation of AMDModel_New.forward (model_AMD.py:1844-2017)."""

    def __init__(self, model):
        self.model = model
        self.tape = Tape()
        self.G = {}

    def forward(self, video, ref_img, time_step, zt, sink, cam_idx=None, obj_idx=None):
        """video, ref_img (n, t, c, h, w); time_step (n*t,) float; zt (n*t, c, h, w) the noised latents of
        scheduler.get_train_tuple -> pre (n*t, c, h, w) fp32.
        cam_idx / obj_idx: kept token indices of random_masking for the two encoders (None = no masking)."""
        m, tape, G = self.model, self.tape, self.G
        n, t, c, h, w = video.shape
        from .frequency_utils import freq_split_gauss
        rv = torch.cat([ref_img, video], dim=1).float().contiguous()
        cam = osrc = otgt = None
        if m.use_filter:
            vol = rv.permute(0, 2, 1, 3, 4).contiguous()
            lo, _hi = freq_split_gauss(vol, 0.6, 0.6)  # the 0.5 high band of forward() is computed and discarded upstream (1904-1908)
            lf_video = lo.permute(0, 2, 1, 3, 4)[:, t:].contiguous()
            if m.use_camera:
                if m.use_camera_down:
                    lf = camera_down_train(tape, m, lf_video.view(n * t, c, h, w), sink, G, "dcamin")
                    lf_video = lf.view(n, t, *lf.shape[1:])
                cam = camera_encoder_train(tape, m.camera_motion_encoder, lf_video, sink, "camera_motion_encoder.", G, "dcam", "dcamin", cam_idx)
            if m.use_object:
                obj = object_encoder_train(tape, m.object_motion_encoder, rv, sink, "object_motion_encoder.", G, "dobj", obj_idx)
                osrc, otgt = obj[:, :t].flatten(0, 1).contiguous(), obj[:, t:].flatten(0, 1).contiguous()
                self._obj_shape = obj.shape
        else:
            x = rv
            if m.use_camera_down:
                y = camera_down_train(tape, m, x.view(n * 2 * t, c, h, w), sink, G, "dcamin_all")
                x = y.view(n, 2 * t, *y.shape[1:])
                raise NotImplementedError("training without use_filter but with use_camera_down is not implemented")
            cam = camera_encoder_train(tape, m.camera_motion_encoder, x[:, t:].contiguous(), sink, "camera_motion_encoder.", G, "dcam", "dcamin", cam_idx)
        zi, zt = ref_img.flatten(0, 1).float().contiguous(), zt.float().contiguous()

        def join_bwd():  # the object encoder saw [ref | video]: its gradient is the two halves side by side
            if m.use_object and m.use_filter:
                d = torch.empty(self._obj_shape, dtype=torch.float32, device=zi.device)
                d[:, :t] = G["dosrc"].view(n, t, *d.shape[2:])
                d[:, t:] = G["dotgt"].view(n, t, *d.shape[2:])
                G["dobj"] = d

        tape.push(join_bwd)
        pre = dit_train(tape, m.diffusion_transformer, zi, zt, time_step.float(), cam if m.use_camera else None,
                        osrc if m.use_object else None, otgt if m.use_object else None, sink, "diffusion_transformer.", G)
        return pre

    def backward(self, dpre, on_done=None):
        self.G["dpre"] = dpre
        self.tape.run(on_done)
        self.G.clear()


def new_sink(model):
    """fresh zero gradients for every trainable parameter"""
    return {k: torch.zeros_like(p, dtype=torch.float32) for k, p in model.named_parameters() if p.requires_grad}


class _ForwardFn(torch.autograd.Function):
    """makes AMDModel_New.forward differentiable for torch: pre = f(parameters); backward replays the tape"""

    @staticmethod
    def forward(fctx, model, video, ref_img, time_step, zt, cam_idx, obj_idx, *params):
        run = TrainRun(model)
        sink = new_sink(model)
        pre = run.forward(video, ref_img, time_step, zt, sink, cam_idx, obj_idx)
        fctx.run, fctx.sink = run, sink
        fctx.names = [k for k, p in model.named_parameters() if p.requires_grad]
        return pre

    @staticmethod
    def backward(fctx, dpre):
        fctx.run.backward(dpre.float().contiguous())
        grads = tuple(fctx.sink[k] for k in fctx.names)
        fctx.run = fctx.sink = None
        return (None,) * 7 + grads


class _MseFn(torch.autograd.Function):
    """model/loss.py:13-14 with a device-side backward"""

    @staticmethod
    def forward(fctx, pre, gt):
        from . import optim
        pre, gt = pre.detach().float().contiguous(), gt.detach().float().contiguous()
        fctx.save_for_backward(pre, gt)
        return optim.mse_loss(pre, gt)

    @staticmethod
    def backward(fctx, g):
        pre, gt = fctx.saved_tensors
        return B.mse_backward(pre, gt, g.float().reshape(1).contiguous()), None


def forward_with_grad(model, video, ref_img, time_step, zt, cam_idx=None, obj_idx=None):
    params = [p for p in model.parameters() if p.requires_grad]
    return _ForwardFn.apply(model, video, ref_img, time_step, zt, cam_idx, obj_idx, *params)


def mse(pre, gt):
    return _MseFn.apply(pre, gt)
