"""CPU ORACLE — TEST INFRASTRUCTURE ONLY (never imported by the product path).

Functional, state-dict driven fp32 PyTorch restatement of the Hi-VAE hot path
(reference = HuaizeLiu/Hi-VAE, paths relative to /root/reference):

  * Haar DWT / IDWT                     model/wavelet.py:13-48
  * Gaussian low-pass + FFT band split  model/frequency_utils.py:45-108
  * rectified-flow schedule             model/rectified_flow.py:18-45,70-75
  * RoPE attention                      model/attention.py:6-117
  * transformer blocks                  model/modules.py:651-1188
  * motion encoders                     model/transformer.py:36-162, 406-556
  * default DiT                         model/transformer.py:1746-2043
  * AMDModel_New encode/decode/sample   model/model_AMD.py:1844-2195, 2379-2536
  * diffusers 0.30.0 Attention / FeedForward / Timesteps / sincos tables
    (third-party, absent from /root/reference; semantics restated, see oracle/shim)

Parity status: PINNED against the unmodified reference run through oracle/shim in
the build container (tests/golden/make_golden.py asserts agreement and writes the
fixtures in tests/golden/); the reference itself ships no golden vectors, so the
diffusers semantics inside the shim remain "parity unpinned" upstream.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this file.
"""
import math

import numpy as np
import torch
import torch.nn.functional as F


# ----------------------------------------------------------------------------- wavelet
def haar_dwt2(x):
    """wavelet.py:13-27 — returns (LL, HL, LH, HH), each (B,C,H/2,W/2)."""
    a = x[:, :, 0::2, 0::2] / 2
    b = x[:, :, 1::2, 0::2] / 2
    c = x[:, :, 0::2, 1::2] / 2
    d = x[:, :, 1::2, 1::2] / 2
    return a + b + c + d, -a - b + c + d, -a + b - c + d, a - b - c + d


def haar_idwt2(x):
    """wavelet.py:31-48 — x = cat(LL,HL,LH,HH) on dim 0; output always fp32."""
    B = x.shape[0] // 4
    x1, x2, x3, x4 = (x[i * B:(i + 1) * B] / 2 for i in range(4))
    h = torch.zeros(B, x.shape[1], 2 * x.shape[2], 2 * x.shape[3], dtype=torch.float32, device=x.device)
    h[:, :, 0::2, 0::2] = x1 - x2 - x3 + x4
    h[:, :, 1::2, 0::2] = x1 - x2 + x3 - x4
    h[:, :, 0::2, 1::2] = x1 + x2 - x3 - x4
    h[:, :, 1::2, 1::2] = x1 + x2 + x3 + x4
    return h


def haar_pyramid(x, levels):
    """Multi-level analysis by recursing on LL (the caller-side loop of
    frenquence_wavelet.py:86-89).  Returns [(HL,LH,HH) per level], final LL."""
    bands = []
    ll = x
    for _ in range(levels):
        ll, hl, lh, hh = haar_dwt2(ll)
        bands.append((hl, lh, hh))
    return bands, ll


def haar_pyramid_inverse(bands, ll):
    for hl, lh, hh in reversed(bands):
        ll = haar_idwt2(torch.cat([ll, hl, lh, hh], 0))
    return ll


# ----------------------------------------------------------------------------- frequency
def gaussian_lpf(shape, d_s=0.25, d_t=0.25):
    """frequency_utils.py:45-81 (3-D mask only)."""
    T, H, W = shape[-3], shape[-2], shape[-1]
    if d_s == 0 or d_t == 0:
        return torch.zeros(T, H, W)
    t = torch.arange(T).float() * 2 / T - 1
    h = torch.arange(H).float() * 2 / H - 1
    w = torch.arange(W).float() * 2 / W - 1
    gt, gh, gw = torch.meshgrid(t, h, w, indexing="ij")
    d2 = (gt * (1 / d_t)).pow(2) + (gh * (1 / d_s)).pow(2) + (gw * (1 / d_s)).pow(2)
    return torch.exp(-0.5 * d2)


def freq_split3d(x, lpf):
    """frequency_utils.py:84-108."""
    dims = (-3, -2, -1)
    xf = torch.fft.fftshift(torch.fft.fftn(x, dim=dims), dim=dims)
    lo = torch.fft.ifftn(torch.fft.ifftshift(xf * lpf, dim=dims), dim=dims).real
    hi = torch.fft.ifftn(torch.fft.ifftshift(xf * (1 - lpf), dim=dims), dim=dims).real
    return lo, hi


# ----------------------------------------------------------------------------- rectified flow
def rf_time(step, num_step=1000):
    """rectified_flow.py:70-75."""
    t = (num_step - step) / num_step
    return t.view(-1, 1, 1, 1) if t.dim() == 1 else t


def rf_train_tuple(z1, step, z0=None, num_step=1000):
    """rectified_flow.py:18-40 — draws z0 with torch.randn_like when absent."""
    t = rf_time(step, num_step)
    if z0 is None:
        z0 = torch.randn_like(z1)
    return t * z1 + (1.0 - t) * z0, z1 - z0


def rf_target(zt, vel, step, num_step=1000):
    """rectified_flow.py:42-45."""
    return zt + (1.0 - rf_time(step, num_step)) * vel


def sample_steps(start_step, sample_step):
    """model_AMD.py:2134-2135."""
    seq = np.linspace(0, start_step, num=sample_step + 1, endpoint=True, dtype=int)
    return [int(s) for s in reversed(seq[1:])]


# ----------------------------------------------------------------------------- embeddings (diffusers 0.30.0)
def sincos_1d(dim, pos):
    omega = np.arange(dim // 2, dtype=np.float64) / (dim / 2.0)
    omega = 1.0 / 10000 ** omega
    out = np.einsum("m,d->md", np.asarray(pos, dtype=np.float64).reshape(-1), omega)
    return np.concatenate([np.sin(out), np.cos(out)], axis=1)


def sincos_2d(dim, gh, gw, base_size=16):
    grid_h = np.arange(gh, dtype=np.float32) / (gh / base_size)
    grid_w = np.arange(gw, dtype=np.float32) / (gw / base_size)
    grid = np.stack(np.meshgrid(grid_w, grid_h), axis=0).reshape([2, 1, gw, gh])
    return np.concatenate([sincos_1d(dim // 2, grid[0]), sincos_1d(dim // 2, grid[1])], axis=1)


def timestep_embedding(t, dim):
    """Timesteps(dim, flip_sin_to_cos=True, freq_shift=0) — [cos | sin]."""
    half = dim // 2
    freqs = torch.exp(-math.log(10000) * torch.arange(half, dtype=torch.float32, device=t.device) / half)
    arg = t[:, None].float() * freqs[None]
    return torch.cat([torch.cos(arg), torch.sin(arg)], dim=-1)


# ----------------------------------------------------------------------------- blocks
def _lin(sd, p, x):
    return F.linear(x, sd[p + ".weight"], sd.get(p + ".bias"))


def _ln(sd, p, x, eps=1e-5):
    return F.layer_norm(x, (x.shape[-1],), sd.get(p + ".weight"), sd.get(p + ".bias"), eps)


def attention(sd, p, x, enc, heads):
    """diffusers Attention with qk_norm='layer_norm' eps 1e-6 (modules.py:699-707)."""
    B = x.shape[0]
    enc = x if enc is None else enc
    q, k, v = _lin(sd, p + ".to_q", x), _lin(sd, p + ".to_k", enc), _lin(sd, p + ".to_v", enc)
    hd = q.shape[-1] // heads
    q, k, v = (z.view(B, -1, heads, hd).transpose(1, 2) for z in (q, k, v))
    q = _ln(sd, p + ".norm_q", q, 1e-6)
    k = _ln(sd, p + ".norm_k", k, 1e-6)
    s = (q @ k.transpose(-1, -2)) / math.sqrt(hd)
    o = torch.softmax(s, dim=-1) @ v
    o = o.transpose(1, 2).reshape(B, -1, heads * hd)
    return _lin(sd, p + ".to_out.0", o)


def feed_forward(sd, p, x):
    return _lin(sd, p + ".net.2", F.gelu(_lin(sd, p + ".net.0.proj", x), approximate="tanh"))


def basic_block(sd, p, x, heads, enc=None):
    """BasicTransformerBlock / BasicCrossTransformerBlock (modules.py:721-743, 817-840)."""
    x = x + attention(sd, p + ".attn1", _ln(sd, p + ".norm1", x), enc, heads)
    return x + feed_forward(sd, p + ".ff", _ln(sd, p + ".norm2", x))


def _ada(sd, p, temb, k):
    return _lin(sd, p + ".linear", F.silu(temb)).chunk(k, dim=1)


def joint_block(sd, p, hid, enc, temb, heads):
    """AMDTransformerBlock (modules.py:1065-1109): hid = motion, enc = image."""
    Li = enc.shape[1]
    sh, sc, g, esh, esc, eg = _ada(sd, p + ".norm1", temb, 6)
    nh = _ln(sd, p + ".norm1.norm", hid) * (1 + sc)[:, None] + sh[:, None]
    ne = _ln(sd, p + ".norm1.norm", enc) * (1 + esc)[:, None] + esh[:, None]
    a = attention(sd, p + ".attn1", torch.cat([ne, nh], 1), None, heads)
    hid = hid + g[:, None] * a[:, Li:]
    enc = enc + eg[:, None] * a[:, :Li]
    sh, sc, g, esh, esc, eg = _ada(sd, p + ".norm2", temb, 6)
    nh = _ln(sd, p + ".norm2.norm", hid) * (1 + sc)[:, None] + sh[:, None]
    ne = _ln(sd, p + ".norm2.norm", enc) * (1 + esc)[:, None] + esh[:, None]
    f = feed_forward(sd, p + ".ff", torch.cat([ne, nh], 1))
    return hid + g[:, None] * f[:, Li:], enc + eg[:, None] * f[:, :Li]


def dit_block(sd, p, x, temb, heads):
    """BasicDiTBlock (modules.py:1160-1188)."""
    sh, sc, g = _ada(sd, p + ".norm1", temb, 3)
    x = x + g[:, None] * attention(sd, p + ".attn1", _ln(sd, p + ".norm1.norm", x) * (1 + sc)[:, None] + sh[:, None], None, heads)
    sh, sc, g = _ada(sd, p + ".norm2", temb, 3)
    return x + g[:, None] * feed_forward(sd, p + ".ff", _ln(sd, p + ".norm2.norm", x) * (1 + sc)[:, None] + sh[:, None])


def patch_embed(sd, p, x, patch):
    """PatchEmbed (modules.py:858-879) for a 4-D input: (B,C,H,W) -> (B, hw, D)."""
    y = F.conv2d(x, sd[p + ".proj.weight"], sd[p + ".proj.bias"], stride=patch)
    return y.flatten(2).transpose(1, 2)


def _count_layers(sd, prefix):
    n = 0
    while (prefix + ".%d.norm1.weight" % n) in sd or (prefix + ".%d.norm1.linear.weight" % n) in sd:
        n += 1
    return n


# ----------------------------------------------------------------------------- encoders / DiT
def mask_keep(noise, mask_ratio):
    """random_masking (transformer.py:164-192, 369-397, 558-586): indices of the int(L (1 - ratio)) smallest noise values"""
    return torch.argsort(noise, dim=1)[:, :int(noise.shape[1] * (1 - mask_ratio))]


def object_encoder(sd, p, x, heads, patch=2, need_norm_out=False, keep=None):
    """MotionEncoderLearnTokenTransformer.forward (transformer.py:116-162); keep (N*T, len_keep) = kept image tokens per frame."""
    N, T, C, H, W = x.shape
    D = sd[p + ".motion_embed.weight"].shape[0]
    mt = _lin(sd, p + ".motion_embed", sd[p + ".motion_token"]).repeat(N * T, 1, 1)
    img = patch_embed(sd, p + ".patch_embed", x.reshape(N * T, C, H, W), patch)
    pos = torch.from_numpy(sincos_2d(D, H // patch, W // patch)).float().to(x.device)
    img = img + pos[None]
    if keep is not None:
        img = torch.gather(img, 1, keep.to(x.device)[..., None].expand(-1, -1, D))
    h = torch.cat([mt, img], 1)
    for i in range(_count_layers(sd, p + ".transformer_blocks")):
        h = basic_block(sd, p + ".transformer_blocks.%d" % i, h, heads)
    m = _lin(sd, p + ".proj_out", _ln(sd, p + ".norm_final", h[:, :mt.shape[1]]))
    if need_norm_out:
        m = F.layer_norm(m, (m.shape[-1],))
    return m.view(N, T, *m.shape[1:])


def temporal_motion_encoder(sd, p, x, heads, video_frames, patch=2, need_norm_out=True, keep=None):
    """MotionEncoderLearnTokenTemporalTransformer.forward (transformer.py:298-367); keep (N*T, len_keep) as above.  Keeps the
    reference's two different splits: frames of each clip at 305 (dim=1), the flat (n t) batch at 340 (dim=0)."""
    N, T, C, H, W = x.shape
    D = sd[p + ".motion_embed.weight"].shape[0]
    tok = _lin(sd, p + ".motion_embed", sd[p + ".motion_token"])  # (1, l, D)
    l, half = tok.shape[1], T // 2
    tem = torch.from_numpy(sincos_1d(D, np.arange(video_frames * l))).float().to(x.device)
    mt = tok[None].repeat(N, T, 1, 1).clone()
    mt[:, half:] += tem[:half * l].view(1, half, l, D)
    mt = mt.reshape(N * T, l, D)
    img = patch_embed(sd, p + ".patch_embed", x.reshape(N * T, C, H, W), patch)
    pos = torch.from_numpy(sincos_2d(D, H // patch, W // patch)).float().to(x.device)
    img = img + pos[None]
    if keep is not None:
        img = torch.gather(img, 1, keep.to(x.device)[..., None].expand(-1, -1, D))
    h = torch.cat([mt, img], 1)
    for i in range(_count_layers(sd, p + ".transformer_blocks")):
        h = basic_block(sd, p + ".transformer_blocks.%d" % i, h, heads)
        src, tgt = h[:, :l].chunk(2, dim=0)
        tgt = tgt.reshape(N, half, l, D).permute(0, 2, 1, 3).reshape(N * l, half, D)
        tgt = basic_block(sd, p + ".motion_blocks.%d" % i, tgt, heads)  # AMDTransformerMotionBlock, temb=None (modules.py:1246-1285)
        tgt = tgt.view(N, l, half, D).permute(0, 2, 1, 3).reshape(N * half, l, D)
        h = torch.cat([torch.cat([src, tgt], 0), h[:, l:]], 1)
    m = _lin(sd, p + ".proj_out", _ln(sd, p + ".norm_final", h[:, :l]))
    if need_norm_out:
        m = F.layer_norm(m, (m.shape[-1],))
    return m.view(N, T, *m.shape[1:])


def camera_encoder(sd, p, x, heads, video_frames, patch=2, need_norm_out=False, keep=None):
    """MotionEncoderLearnTokenOnlyTemporalTransformer.forward (transformer.py:495-556); keep (N, len_keep) = the spatial
    locations a clip keeps in every frame (558-586)."""
    N, T, C, H, W = x.shape
    D = sd[p + ".motion_embed.weight"].shape[0]
    S = (H // patch) * (W // patch)
    img = patch_embed(sd, p + ".patch_embed", x.reshape(N * T, C, H, W), patch)
    pos = torch.from_numpy(sincos_2d(D, H // patch, W // patch)).float().to(x.device)
    tem = torch.from_numpy(sincos_1d(D, np.arange(video_frames))).float().to(x.device)
    img = (img + pos[None]).view(N, T, S, D) + tem[None, :T, None]
    if keep is not None:
        S = keep.shape[1]
        img = torch.gather(img, 2, keep.to(x.device)[:, None, :, None].expand(-1, T, -1, D))
    img = img.permute(0, 2, 1, 3).reshape(N * S, T, D)  # (n s) t c
    tok = _lin(sd, p + ".motion_embed", sd[p + ".motion_token"])[0]  # (t0, D)
    if tok.shape[0] != T:
        tok = tok.repeat_interleave(T // tok.shape[0], dim=0)
    m = (tok + tem[:T])[None].repeat(N * S, 1, 1)
    for i in range(_count_layers(sd, p + ".transformer_blocks")):
        m = basic_block(sd, p + ".transformer_blocks.%d" % i, m, heads, enc=img)
    m = _lin(sd, p + ".proj_out", _ln(sd, p + ".norm_final", m))
    if need_norm_out:
        m = F.layer_norm(m, (m.shape[-1],))
    return m.view(N, S, T, -1).permute(0, 2, 1, 3).contiguous()  # n t s c


def camera_down(sd, p, x):
    """model_AMD.py:1788-1793."""
    x = F.conv2d(x, sd[p + ".0.weight"], sd[p + ".0.bias"], stride=2, padding=1)
    x = F.conv2d(x, sd[p + ".1.weight"], sd[p + ".1.bias"], stride=1, padding=1)
    return F.max_pool2d(x, 2, 2)


def dit_forward(sd, p, image, timestep, cam, osrc, otgt, heads, T, patch=2, motion_token_num=12):
    """AMDDiffusionTransformerModelImgSpatialTempMotion.forward (transformer.py:1911-2043)."""
    N, Ci, Hi, Wi = image.shape
    n = N // T
    D = sd[p + ".image_patch_embed.proj.weight"].shape[0]
    S = (Hi // patch) * (Wi // patch)
    img = patch_embed(sd, p + ".image_patch_embed", image, patch)
    temb = timestep_embedding(timestep, D)
    emb = _lin(sd, p + ".time_embedding.linear_2", F.silu(_lin(sd, p + ".time_embedding.linear_1", temb)))
    emb_s = emb.view(n, T, -1)[:, :1].repeat(1, S, 1).flatten(0, 1)
    dev = image.device
    if cam is not None:
        cmo = _lin(sd, p + ".camera_motion_patch_embed", cam.flatten(0, 1))
    if osrc is not None:
        Lo = otgt.shape[1]
        mpos = torch.from_numpy(sincos_1d(D, np.arange(2 + 2 * motion_token_num))).float().to(dev)
        omo = torch.cat([sd[p + ".source_token"].repeat(N, 1, 1), _lin(sd, p + ".object_motion_patch_embed", osrc),
                         sd[p + ".target_token"].repeat(N, 1, 1), _lin(sd, p + ".object_motion_patch_embed", otgt)], 1)
        omo = omo + mpos[None, :2 * Lo + 2]
    pos = torch.from_numpy(sincos_2d(D, Hi // patch, Wi // patch)).float().to(dev)
    tpos = torch.from_numpy(sincos_1d(D, np.arange(T))).float().to(dev)
    img = (img + pos[None]).view(n, T, S, D) + tpos[None, :, None]
    img = img.view(N, S, D)
    has_temporal = _count_layers(sd, p + ".spatial_blocks") > 0  # absent in ...TempMotion (transformer.py:1483-1743)
    n_layers = max(_count_layers(sd, p + ".spatial_blocks"), _count_layers(sd, p + ".object_transformer_blocks"))
    for i in range(n_layers):
        if osrc is not None:
            omo, img = joint_block(sd, p + ".object_transformer_blocks.%d" % i, omo, img, emb, heads)
        if cam is not None:
            cmo, img = joint_block(sd, p + ".camera_transformer_blocks.%d" % i, cmo, img, emb, heads)
        if has_temporal:
            x = img.view(n, T, S, D).permute(0, 2, 1, 3).reshape(n * S, T, D)
            x = dit_block(sd, p + ".spatial_blocks.%d" % i, x, emb_s, heads)
            img = x.view(n, S, T, D).permute(0, 2, 1, 3).reshape(N, S, D)
    img = _ln(sd, p + ".norm_final", img)
    shift, scale = _lin(sd, p + ".norm_out.linear", F.silu(emb)).chunk(2, dim=1)
    img = _ln(sd, p + ".norm_out.norm", img) * (1 + scale)[:, None] + shift[:, None]
    out = _lin(sd, p + ".proj_out", img)
    C = out.shape[-1] // (patch * patch)
    out = out.reshape(N, Hi // patch, Wi // patch, C, patch, patch).permute(0, 3, 1, 4, 2, 5)
    return out.reshape(N, C, Hi, Wi)


def dit_imgspatial_forward(sd, p, cam_tgt, image, timestep, heads, T, patch=2, motion_token_num=12, motion_type="plus",
                           cam_src=None, osrc=None, otgt=None, legacy=False):
    """AMDDiffusionTransformerModelImgSpatial.forward (transformer.py:2186-2356): one motion stream; 'decouple'
    with object motions swaps the camera rows for object tokens after layer 5 (2259-2322) and forms the object
    target rows from the object SOURCE embedding (2262-2263).
    legacy=True: AMDDiffusionTransformerModel.forward (transformer.py:1130-1259) — no temporal blocks / table,
    'plus' embeds camera + object, camera stream for layers [0, 8), object stream for layers [6, n)."""
    N, Ci, Hi, Wi = image.shape
    L = cam_tgt.shape[1]
    T = 1 if legacy else T
    n, dev = N // T, image.device
    D = sd[p + ".image_patch_embed.proj.weight"].shape[0]
    S = (Hi // patch) * (Wi // patch)
    img = patch_embed(sd, p + ".image_patch_embed", image, patch)
    emb = _lin(sd, p + ".time_embedding.linear_2", F.silu(_lin(sd, p + ".time_embedding.linear_1", timestep_embedding(timestep, D))))
    emb_s = emb.view(n, T, -1)[:, :1].repeat(1, S, 1).flatten(0, 1)
    mpos = torch.from_numpy(sincos_1d(D, np.arange(2 + 2 * motion_token_num))).float().to(dev)
    stok, ttok = sd[p + ".source_token"].repeat(N, 1, 1), sd[p + ".target_token"].repeat(N, 1, 1)
    me = lambda x: _lin(sd, p + ".motion_patch_embed", x)
    if motion_type == "plus":
        ms, mt = (cam_src + osrc, cam_tgt + otgt) if legacy else (osrc, otgt)
        mo = torch.cat([stok, me(ms), ttok, me(mt)], 1) + mpos[None, :2 * L + 2]
    else:
        if cam_src is not None:
            mo = torch.cat([stok, me(cam_src), ttok, me(cam_tgt)], 1) + mpos[None, :2 * L + 2]
        else:
            mo = torch.cat([ttok, me(cam_tgt)], 1) + mpos[None, :L + 1]
        if osrc is not None:
            o_s = me(osrc) + mpos[None, 1:L + 1]
            o_t = o_s + mpos[None, L + 2:2 * L + 2]
    pos = torch.from_numpy(sincos_2d(D, Hi // patch, Wi // patch)).float().to(dev)
    img = img + pos[None]
    if not legacy:
        tpos = torch.from_numpy(sincos_1d(D, np.arange(T))).float().to(dev)
        img = (img.view(n, T, S, D) + tpos[None, :, None]).view(N, S, D)
    nl = _count_layers(sd, p + ".transformer_blocks")
    cam_end, obj_start = (8, 6) if legacy else (6, 6)

    def layer(i, mo, img):
        mo, img = joint_block(sd, p + ".transformer_blocks.%d" % i, mo, img, emb, heads)
        if legacy:
            return mo, img
        x = img.view(n, T, S, D).permute(0, 2, 1, 3).reshape(n * S, T, D)
        x = dit_block(sd, p + ".spatial_blocks.%d" % i, x, emb_s, heads)
        return mo, x.view(n, S, T, D).permute(0, 2, 1, 3).reshape(N, S, D)

    if motion_type == "decouple" and osrc is not None:
        for i in range(min(cam_end, nl)):
            mo, img = layer(i, mo, img)
        if cam_src is not None:
            s_tok, t_tok = mo[:, :1], mo[:, 1 + L:2 + L]
        else:
            s_tok, t_tok = stok, mo[:, :1]
        mo = torch.cat([s_tok, o_s, t_tok, o_t], 1)
        for i in range(obj_start, nl):
            mo, img = layer(i, mo, img)
    else:
        for i in range(nl):
            mo, img = layer(i, mo, img)
    img = _ln(sd, p + ".norm_final", img)
    shift, scale = _lin(sd, p + ".norm_out.linear", F.silu(emb)).chunk(2, dim=1)
    img = _ln(sd, p + ".norm_out.norm", img) * (1 + scale)[:, None] + shift[:, None]
    out = _lin(sd, p + ".proj_out", img)
    C = out.shape[-1] // (patch * patch)
    out = out.reshape(N, Hi // patch, Wi // patch, C, patch, patch).permute(0, 3, 1, 4, 2, 5)
    return out.reshape(N, C, Hi, Wi)


# ----------------------------------------------------------------------------- model
class AMDOracle:
    """AMDModel_New (model_AMD.py:1663-2838) for use_filter / use_camera_down /
    use_camera / use_object = True, diffusion_model_type='spatial', use_grey=False."""

    def __init__(self, sd, enc_nhead, dit_nhead, video_frames=16, patch=2, motion_token_num=12,
                 num_step=1000, need_norm_out=False):
        self.sd = {k: v.float() for k, v in sd.items()}
        self.eh, self.dh, self.T, self.patch = enc_nhead, dit_nhead, video_frames, patch
        self.mtn, self.num_step, self.nno = motion_token_num, num_step, need_norm_out

    def encode(self, video, ref_img, d_lo=0.6, cam_keep=None, obj_keep=None):
        n, t, c, h, w = video.shape
        rv = torch.cat([ref_img, video], 1)
        lpf = gaussian_lpf([2 * t, h, w], d_lo, d_lo).to(video.device)
        lo, _ = freq_split3d(rv.permute(0, 2, 1, 3, 4), lpf)
        cam = None
        if "camera_motion_encoder.motion_token" in self.sd:  # use_camera
            lf = lo.permute(0, 2, 1, 3, 4)[:, t:]
            lf = camera_down(self.sd, "camera_down", lf.flatten(0, 1))
            lf = lf.view(n, t, *lf.shape[1:])
            cam = camera_encoder(self.sd, "camera_motion_encoder", lf, self.eh, self.T, self.patch, self.nno, cam_keep)
        obj = object_encoder(self.sd, "object_motion_encoder", rv, self.eh, self.patch, self.nno, obj_keep)
        return cam, obj[:, :t].flatten(0, 1), obj[:, t:].flatten(0, 1)

    def dit(self, zi, zt, step, cam, osrc, otgt):
        return dit_forward(self.sd, "diffusion_transformer", torch.cat([zi, zt], 1), step, cam, osrc, otgt,
                           self.dh, self.T, self.patch, self.mtn)

    def decode(self, video, ref_img, cam, osrc, otgt, start_step, sample_step=50, z0=None, trace=None):
        n, t = video.shape[:2]
        zi, zj = ref_img.flatten(0, 1), video.flatten(0, 1)
        step = torch.ones(zj.shape[0], device=video.device) * start_step
        zt, _ = rf_train_tuple(zj, step, z0, self.num_step)
        dt = 1.0 / sample_step
        for i in sample_steps(start_step, sample_step):
            pre = self.dit(zi, zt, torch.ones(zt.shape[0], device=zt.device) * i, cam, osrc, otgt)
            zt = zt + pre * dt
            if trace is not None:
                trace.append(pre)
        shp = (n, t) + tuple(zi.shape[1:])
        return zi.view(shp), zt.view(shp), zj.view(shp)

    def sample(self, video, ref_img, sample_step=50, start_step=None, z0=None, trace=None):
        start_step = self.num_step if start_step is None else start_step
        cam, osrc, otgt = self.encode(video, ref_img)
        return self.decode(video, ref_img, cam, osrc, otgt, start_step, sample_step, z0, trace)

    def sample_cross(self, video_1, video_2, ref_img, sample_step=50, start_step=None, z0=None, trace=None):
        """model_AMD.py:2198-2377 (use_filter, use_camera, use_grey=False): camera motion from the low band of
        video_1 (cut-off 0.5 over its own t-frame volume, 2219-2221); the denoiser gets NO object motion (2349-2351)."""
        n, t, c, h, w = video_1.shape
        start_step = self.num_step if start_step is None else start_step
        lo, _ = freq_split3d(video_1.permute(0, 2, 1, 3, 4), gaussian_lpf([t, h, w], 0.5, 0.5).to(video_1.device))
        lf = camera_down(self.sd, "camera_down", lo.permute(0, 2, 1, 3, 4).flatten(0, 1))
        cam = camera_encoder(self.sd, "camera_motion_encoder", lf.view(n, t, *lf.shape[1:]), self.eh, self.T, self.patch, self.nno)
        return self.decode(video_2, ref_img, cam, None, None, start_step, sample_step, z0, trace)

    def sample_with_refimg_motion(self, ref_img, cam, otgt, ref_img_grey, sample_step=10, z0=None):
        """model_AMD.py:2701-2785: object source motion from the HIGH band (cut-off 0.6, t-frame volume) of
        ref_img_grey; zj = zi; zj is returned flat (2774-2775)."""
        n, t, c, h, w = ref_img.shape
        _, hi = freq_split3d(ref_img_grey.permute(0, 2, 1, 3, 4), gaussian_lpf([t, h, w], 0.6, 0.6).to(ref_img.device))
        osrc = object_encoder(self.sd, "object_motion_encoder", hi.permute(0, 2, 1, 3, 4), self.eh, self.patch, self.nno).flatten(0, 1)
        zi, zt, zj = self.decode(ref_img, ref_img, cam, osrc, otgt, self.num_step, sample_step, z0)
        return zi, zt, zj.flatten(0, 1)

    def forward(self, video, ref_img, time_step, z0=None):
        """Training forward (model_AMD.py:1844-2017) with an explicit per-frame time_step
        (b,) and noise z0; returns (pre, vel, loss dict)."""
        n, t = video.shape[:2]
        cam, osrc, otgt = self.encode(video, ref_img)
        zi, zj = ref_img.flatten(0, 1), video.flatten(0, 1)
        zt, vel = rf_train_tuple(zj, time_step, z0, self.num_step)
        pre = self.dit(zi, zt, time_step, cam, osrc, otgt)
        diff = F.mse_loss(pre, vel)
        rec = F.mse_loss(rf_target(zt, pre, time_step, self.num_step), zj)
        return pre, vel, {"loss": diff, "diff_loss": diff, "rec_loss": rec}


class AMDModelOracle:
    """AMDModel (model_AMD.py:33-594) for use_filter + use_grey (or neither), use_regularizers=False,
    need_motion_transformer=False: motions() = 272-343 / 441-500, forward = 258-423, sample = 426-594."""

    def __init__(self, sd, cfg):
        self.sd, self.c = {k: v.float() for k, v in sd.items()}, cfg
        self.T, self.num_step = cfg["video_frames"], cfg.get("scheduler_num_step", 1000)

    def _enc(self, p, x):
        c = self.c
        if c.get("use_motiontemporal", True):
            return temporal_motion_encoder(self.sd, p, x, c["enc_nhead"], self.T, c.get("image_patch_size", 2), c.get("motion_need_norm_out", False))
        return object_encoder(self.sd, p, x, c["enc_nhead"], c.get("image_patch_size", 2), c.get("motion_need_norm_out", False))

    def motions(self, video, ref_img, video_grey=None, ref_img_grey=None, camera_mask=None):
        c = self.c
        n, t, ch, h, w = video.shape
        rv = torch.cat([ref_img, video], 1)
        if c.get("use_filter", False):
            grey = torch.cat([ref_img_grey, video_grey], 1).permute(0, 2, 1, 3, 4)
            lo, _ = freq_split3d(grey, gaussian_lpf([2 * t, h, w], c.get("filter_num", 0.4), c.get("filter_num", 0.4)))
            _, hi = freq_split3d(grey, gaussian_lpf([2 * t, h, w], c.get("high_filter_num", 0.6), c.get("high_filter_num", 0.6)))
            lf, hf = lo.permute(0, 2, 1, 3, 4), hi.permute(0, 2, 1, 3, 4)
            if camera_mask is not None:
                lf = lf * camera_mask
            if c.get("use_camera_down", False):
                y = camera_down(self.sd, "camera_down", lf.flatten(0, 1))
                lf = y.view(n, 2 * t, *y.shape[1:])
            obj, cam = self._enc("object_motion_encoder", hf), self._enc("camera_motion_encoder", lf)
        else:
            obj, cam = self._enc("object_motion_encoder", rv), self._enc("camera_motion_encoder", rv)
        osrc, otgt, csrc, ctgt = obj[:, :t].flatten(0, 1), obj[:, t:].flatten(0, 1), cam[:, :t].flatten(0, 1), cam[:, t:].flatten(0, 1)
        if "camera_motion_map.weight" in self.sd:
            csrc, ctgt = _lin(self.sd, "camera_motion_map", csrc), _lin(self.sd, "camera_motion_map", ctgt)
        if "object_motion_map.weight" in self.sd:
            osrc, otgt = _lin(self.sd, "object_motion_map", osrc), _lin(self.sd, "object_motion_map", otgt)
        return csrc, ctgt, osrc, otgt

    def dit(self, zi, zt, step, m):
        c = self.c
        csrc, ctgt, osrc, otgt = m
        return dit_imgspatial_forward(self.sd, "diffusion_transformer", ctgt, torch.cat([zi, zt], 1), step, c["diffusion_attn_num_heads"],
                                      self.T, c.get("image_patch_size", 2), c.get("motion_token_num", 12), c.get("motion_type", "plus"),
                                      csrc, osrc, otgt, legacy=c.get("diffusion_model_type", "default") == "default")

    def forward(self, video, ref_img, time_step, z0, video_grey=None, ref_img_grey=None, camera_mask=None):
        m = self.motions(video, ref_img, video_grey, ref_img_grey, camera_mask)
        zi, zj = ref_img.flatten(0, 1), video.flatten(0, 1)
        zt, vel = rf_train_tuple(zj, time_step, z0, self.num_step)
        pre = self.dit(zi, zt, time_step, m)
        return pre, vel, {"loss": F.mse_loss(pre, vel), "rec_loss": F.mse_loss(rf_target(zt, pre, time_step, self.num_step), zj)}

    def sample(self, video, ref_img, sample_step, z0, video_grey=None, ref_img_grey=None, start_step=None):
        start_step = self.num_step if start_step is None else start_step
        m = self.motions(video, ref_img, video_grey, ref_img_grey)
        zi, zj = ref_img.flatten(0, 1), video.flatten(0, 1)
        zt, _ = rf_train_tuple(zj, torch.ones(zj.shape[0]) * start_step, z0, self.num_step)
        for i in sample_steps(start_step, sample_step):
            zt = zt + self.dit(zi, zt, torch.ones(zt.shape[0]) * i, m) / sample_step
        return zt.view(video.shape)


class AMDModelCameraOracle:
    """AMDModel_Camera (model_AMD.py:894-1661) in its runnable configuration: use_filter, use_motiontemporal=False (the
    temporal branch splits a t-frame motion into [:t] / [t:] and asserts, 1147-1163), no camera_down (the encoder is sized
    for the full image, 960-972), motion_type 'decouple' (camera tokens only).  zi = ref_img_grey, target = low band."""

    def __init__(self, sd, cfg):
        self.sd, self.c = {k: v.float() for k, v in sd.items()}, cfg
        self.T, self.num_step = cfg["video_frames"], cfg.get("scheduler_num_step", 1000)

    def motions(self, video, ref_img, video_grey, ref_img_grey, camera_mask):
        c = self.c
        n, t, ch, h, w = video.shape
        src = torch.cat([ref_img_grey, video_grey], 1) if c.get("use_grey", False) else torch.cat([ref_img, video], 1)
        lo, _ = freq_split3d(src.permute(0, 2, 1, 3, 4), gaussian_lpf([2 * t, h, w], c.get("filter_num", 0.4), c.get("filter_num", 0.4)))
        lf = lo.permute(0, 2, 1, 3, 4)
        gt = lf[:, t:]
        if c.get("use_mask", True):
            lf = lf * camera_mask
        cam = object_encoder(self.sd, "camera_motion_encoder", lf, c["enc_nhead"], c.get("image_patch_size", 2), c.get("motion_need_norm_out", False))
        csrc, ctgt = cam[:, :t].flatten(0, 1), cam[:, t:].flatten(0, 1)
        if "camera_motion_map.weight" in self.sd:
            csrc, ctgt = _lin(self.sd, "camera_motion_map", csrc), _lin(self.sd, "camera_motion_map", ctgt)
        return csrc, ctgt, gt.flatten(0, 1)

    def dit(self, zi, zt, step, csrc, ctgt):
        c = self.c
        return dit_imgspatial_forward(self.sd, "diffusion_transformer", ctgt, torch.cat([zi, zt], 1), step, c["diffusion_attn_num_heads"],
                                      self.T, c.get("image_patch_size", 2), c.get("motion_token_num", 12), c.get("motion_type", "plus"),
                                      csrc, None, None, legacy=c.get("diffusion_model_type", "default") == "default")

    def forward(self, video, ref_img, time_step, z0, video_grey, ref_img_grey, camera_mask):
        csrc, ctgt, zj = self.motions(video, ref_img, video_grey, ref_img_grey, camera_mask)
        zi = ref_img_grey.flatten(0, 1)
        zt, vel = rf_train_tuple(zj, time_step, z0, self.num_step)
        pre = self.dit(zi, zt, time_step, csrc, ctgt)
        return pre, vel, {"loss": F.mse_loss(pre, vel), "rec_loss": F.mse_loss(rf_target(zt, pre, time_step, self.num_step), zj)}

    def sample(self, video, ref_img, sample_step, z0, video_grey, ref_img_grey, camera_mask, start_step=None):
        start_step = self.num_step if start_step is None else start_step
        csrc, ctgt, zj = self.motions(video, ref_img, video_grey, ref_img_grey, camera_mask)
        zi = ref_img_grey.flatten(0, 1)
        zt, _ = rf_train_tuple(zj, torch.ones(zj.shape[0]) * start_step, z0, self.num_step)
        for i in sample_steps(start_step, sample_step):
            zt = zt + self.dit(zi, zt, torch.ones(zt.shape[0]) * i, csrc, ctgt) / sample_step
        return zt.view(video.shape), zj.view(video.shape)


def rec_transformer_forward(sd, p, src, tgt, image, heads, patch=2):
    """AMDReconstructTransformerModel.forward (transformer.py:805-854): BasicTransformerBlocks over [image | source_token,
    source, target_token, target], norm_final + proj_out on the image rows, (c, p, q) unpatchify."""
    N, _, Hi, Wi = image.shape
    L = src.shape[1]
    D = sd[p + ".motion_patch_embed.weight"].shape[0]
    ms, mt = _lin(sd, p + ".motion_patch_embed", src), _lin(sd, p + ".motion_patch_embed", tgt)
    img = patch_embed(sd, p + ".image_patch_embed", image, patch)
    mo = torch.cat([sd[p + ".source_token"].repeat(N, 1, 1), ms, sd[p + ".target_token"].repeat(N, 1, 1), mt], 1)
    # the table is sized by the denoiser's motion_token_num (2 + 2 * num rows); the forward only reads the first 2 L + 2
    mo = mo + torch.from_numpy(sincos_1d(D, np.arange(2 * L + 2))).float().to(image.device)[None]
    img = img + torch.from_numpy(sincos_2d(D, Hi // patch, Wi // patch)).float().to(image.device)[None]
    h = torch.cat([img, mo], 1)
    for i in range(_count_layers(sd, p + ".transformer_blocks")):
        h = basic_block(sd, p + ".transformer_blocks.%d" % i, h, heads)
    S = img.shape[1]
    y = _lin(sd, p + ".proj_out", _ln(sd, p + ".norm_final", h[:, :S]))
    C = y.shape[-1] // (patch * patch)
    y = y.reshape(N, 1, Hi // patch, Wi // patch, C, patch, patch)
    return y.permute(0, 1, 4, 2, 5, 3, 6).flatten(5, 6).flatten(3, 4).squeeze(1)


class AMDModelRecOracle:
    """AMDModel_Rec.forward (model_AMD.py:2940-3007): two MotionEncoderLearnTokenTransformer over [ref | video], motions added,
    one pass of the reconstruction transformer on [zi | zt_token]; loss = l2(pre, zj)."""

    def __init__(self, sd, cfg):
        self.sd, self.c = {k: v.float() for k, v in sd.items()}, cfg

    def forward(self, video, ref_img):
        c = self.c
        n, t = video.shape[:2]
        rv = torch.cat([ref_img, video], 1)
        nno = c.get("motion_need_norm_out", True)
        obj = object_encoder(self.sd, "object_motion_encoder", rv, c["enc_nhead"], c.get("image_patch_size", 2), nno)
        cam = object_encoder(self.sd, "camera_motion_encoder", rv, c["enc_nhead"], c.get("image_patch_size", 2), nno)
        src = obj[:, :t].flatten(0, 1) + cam[:, :t].flatten(0, 1)
        tgt = obj[:, t:].flatten(0, 1) + cam[:, t:].flatten(0, 1)
        zi, zj = ref_img.flatten(0, 1), video.flatten(0, 1)
        zt = self.sd["zt_token"].repeat(zj.shape[0], 1, 1, 1)
        pre = rec_transformer_forward(self.sd, "transformer", src, tgt, torch.cat([zi, zt], 1), c["diffusion_attn_num_heads"], c.get("image_patch_size", 2))
        return pre, zj, {"loss": F.mse_loss(pre, zj), "rec_loss": F.mse_loss(pre, zj)}


def psnr(a, b):
    """model/metric.py:8-10: 20*log10(1/sqrt(mse))."""
    return float(20 * torch.log10(1.0 / torch.sqrt(F.mse_loss(a, b))))


# ----------------------------------------------------------------------------- RoPE attention (model/attention.py)
def rope_attention(x, wq, bq, wk, bk, wv, bv, n_heads, max_seq_len, start_pos=0, inference=True, mask=None):
    """attention.py:85-117 (note: `self.out` is never applied there)."""
    bs, L, dim = x.shape
    hd = dim // n_heads
    q, k, v = (F.linear(x, w, b).view(bs, L, n_heads, hd) for w, b in ((wq, bq), (wk, bk), (wv, bv)))
    if inference:
        i = torch.arange(1, hd // 2 + 1, dtype=torch.float32)
        theta = 10000.0 ** (-2 * (i - 1) / hd)
        ang = torch.outer(torch.arange(max_seq_len * 2).float(), theta)[start_pos:start_pos + L]
        cos, sin = torch.cos(ang)[None, :, None], torch.sin(ang)[None, :, None]

        def rot(z):
            ze, zo = z[..., 0::2], z[..., 1::2]
            return torch.stack([ze * cos - zo * sin, ze * sin + zo * cos], -1).flatten(3)
        q, k = rot(q), rot(k)
    q, k, v = (z.transpose(1, 2) for z in (q, k, v))
    s = q @ k.transpose(2, 3) / math.sqrt(hd)
    if mask is not None:
        s = s.masked_fill(mask == 0, -1e9)
    o = torch.softmax(s.float(), dim=-1) @ v
    return o.transpose(1, 2).contiguous().view(bs, L, -1)
