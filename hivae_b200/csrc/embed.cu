// Patchify / unpatchify ends of the token pipeline, camera_down convs, RoPE.
#include "hv_common.cuh"

// ---------------------------------------------------------------------------------------------
// PatchEmbed = Conv2d(k=s=p) as a K=(C1+C2)*p*p tiny-K linear, fused with bias + 2-D sincos +
// temporal sincos adds (model/modules.py:858-879, transformer.py:1985-1988).  HBM-bound on the
// output write: TOK tokens per CTA so the weight matrix is re-read from L2 only once per 32 tokens.
template <typename OutT, int TOK>
__global__ void __launch_bounds__(256) patch_embed_kernel(
    const float* __restrict__ in1, int C1, const float* __restrict__ in2, int C2, long long frames, int H, int W, int p,
    const float* __restrict__ w, const float* __restrict__ bias, const float* __restrict__ pos,
    const float* __restrict__ tpos, int Tmod, OutT* __restrict__ out, long long frame_rows, int row_off, int D) {
    extern __shared__ float xs[];  // TOK x K
    const int Wp = W / p, S = (H / p) * Wp, K = (C1 + C2) * p * p;
    const long long tok0 = blockIdx.x * (long long)TOK;
    const long long ntok = frames * S;
    for (int idx = threadIdx.x; idx < TOK * K; idx += blockDim.x) {
        const int t = idx / K, k = idx - t * K;
        const long long tok = tok0 + t;
        float v = 0.f;
        if (tok < ntok) {
            const long long f = tok / S;
            const int s = (int)(tok - f * S);
            const int c = k / (p * p), ij = k - c * p * p, i = ij / p, j = ij - i * p;
            const int y = (s / Wp) * p + i, x = (s % Wp) * p + j;
            v = (c < C1) ? in1[((f * C1 + c) * H + y) * (long long)W + x]
                         : in2[((f * C2 + (c - C1)) * H + y) * (long long)W + x];
        }
        xs[idx] = v;
    }
    __syncthreads();
    if (K == 32 || K == 16) {
        // the DiT (K = 8 ch x 2 x 2) and encoder (K = 4 ch x 2 x 2) patch sizes: the thread keeps its weight row in
        // registers and streams the tokens with 16-byte broadcast reads (4x fewer shared-memory instructions)
        for (int d = threadIdx.x; d < D; d += blockDim.x) {
            float4 wk[8];
            const float4* wr4 = reinterpret_cast<const float4*>(w + (long long)d * K);
#pragma unroll
            for (int k = 0; k < 8; ++k) wk[k] = (4 * k < K) ? __ldg(wr4 + k) : make_float4(0.f, 0.f, 0.f, 0.f);
            const float b = bias ? bias[d] : 0.f;
#pragma unroll 4
            for (int t = 0; t < TOK; ++t) {
                const long long tok = tok0 + t;
                if (tok >= ntok) break;
                const float4* x4 = reinterpret_cast<const float4*>(xs + t * K);
                float acc = 0.f;
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    if (4 * k < K) {
                        const float4 xv = x4[k];
                        acc = fmaf(wk[k].x, xv.x, acc); acc = fmaf(wk[k].y, xv.y, acc); acc = fmaf(wk[k].z, xv.z, acc); acc = fmaf(wk[k].w, xv.w, acc);
                    }
                const long long f = tok / S;
                const int s = (int)(tok - f * S);
                float v = acc + b;
                if (pos) v += pos[(long long)s * D + d];
                if (tpos) v += tpos[(long long)(f % Tmod) * D + d];
                hv_st(out + (f * frame_rows + row_off + s) * D + d, v);
            }
        }
        return;
    }
    for (int d = threadIdx.x; d < D; d += blockDim.x) {
        float acc[TOK];
#pragma unroll
        for (int t = 0; t < TOK; ++t) acc[t] = 0.f;
        const float* wr = w + (long long)d * K;
        for (int k = 0; k < K; ++k) {
            const float wk = wr[k];
#pragma unroll
            for (int t = 0; t < TOK; ++t) acc[t] = fmaf(wk, xs[t * K + k], acc[t]);
        }
        const float b = bias ? bias[d] : 0.f;
#pragma unroll
        for (int t = 0; t < TOK; ++t) {
            const long long tok = tok0 + t;
            if (tok < ntok) {
                const long long f = tok / S;
                const int s = (int)(tok - f * S);
                float v = acc[t] + b;
                if (pos) v += pos[(long long)s * D + d];
                if (tpos) v += tpos[(long long)(f % Tmod) * D + d];
                hv_st(out + (f * frame_rows + row_off + s) * D + d, v);
            }
        }
    }
}

extern "C" int hv_patch_embed(const float* in1, int C1, const float* in2, int C2, int64_t frames, int H, int W, int p,
                              const float* w, const float* bias, const float* pos, const float* tpos, int Tmod,
                              void* out, int out_dtype, int64_t frame_rows, int row_off, int D, void* stream) {
    HV_REQUIRE(p > 0 && H % p == 0 && W % p == 0, "hv_patch_embed: H,W must be multiples of the patch size");
    HV_REQUIRE(C1 > 0 && C2 >= 0 && (C2 == 0 || in2 != nullptr), "hv_patch_embed: bad channel split");
    HV_REQUIRE(!tpos || Tmod > 0, "hv_patch_embed: Tmod must be > 0 with a temporal table");
    const int S = (H / p) * (W / p), K = (C1 + C2) * p * p;
    HV_REQUIRE(row_off >= 0 && row_off + S <= frame_rows, "hv_patch_embed: bad row range");
    HV_REQUIRE(K <= 256, "hv_patch_embed: K=%d > 256 unsupported", K);
    if (frames == 0) return HV_OK;
    constexpr int TOK = 32;
    const long long ntok = (long long)frames * S;
    dim3 grid((unsigned)((ntok + TOK - 1) / TOK));
    size_t smem = (size_t)TOK * K * sizeof(float);
    cudaStream_t st = hv_stream(stream);
    if (out_dtype == HV_F32)
        patch_embed_kernel<float, TOK><<<grid, 256, smem, st>>>(in1, C1, in2, C2, frames, H, W, p, w, bias, pos, tpos, Tmod,
                                                                (float*)out, frame_rows, row_off, D);
    else if (out_dtype == HV_BF16)
        patch_embed_kernel<__nv_bfloat16, TOK><<<grid, 256, smem, st>>>(in1, C1, in2, C2, frames, H, W, p, w, bias, pos, tpos, Tmod,
                                                                        (__nv_bfloat16*)out, frame_rows, row_off, D);
    else HV_REQUIRE(false, "hv_patch_embed: bad dtype");
    HV_LAUNCH_CHECK();
    return HV_OK;
}

// ---------------------------------------------------------------------------------------------
// DiT head: LN(affine) -> LN(affine)*(1+scale)+shift -> Linear(D -> O=C*p*p) -> unpatchify (c,p,q)
// -> optional Euler update.  One warp per token; the O dot products are reduced by shuffles.
template <int MAXNV>
__global__ void __launch_bounds__(256) final_head_kernel(
    const float* __restrict__ res, long long rows, HvRowMap map, int S, int D,
    const float* __restrict__ n1w, const float* __restrict__ n1b, const float* __restrict__ n2w, const float* __restrict__ n2b,
    float eps, const float* __restrict__ ada, long long ada_ld, int shift_col, int scale_col,
    const float* __restrict__ w, const float* __restrict__ bias, int C, int p, int Wp,
    float* __restrict__ pre, float* __restrict__ zt, float dt) {
    const int lane = threadIdx.x & 31;
    const long long r = blockIdx.x * (long long)(blockDim.x >> 5) + (threadIdx.x >> 5);
    if (r >= rows) return;
    const int nv = D >> 7;
    int frame, abs_i;
    const long long rr = map.res_row(r, frame, abs_i);
    const float4* xp = reinterpret_cast<const float4*>(res + rr * D);
    float4 v[MAXNV];
    float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int j = 0; j < MAXNV; ++j) v[j] = (j < nv) ? xp[j * 32 + lane] : zero;
    const float invD = 1.f / (float)D;
#pragma unroll
    for (int pass = 0; pass < 2; ++pass) {
        float sum = 0.f;
#pragma unroll
        for (int j = 0; j < MAXNV; ++j) if (j < nv) sum += (v[j].x + v[j].y) + (v[j].z + v[j].w);
        const float mean = hv_warp_sum(sum) * invD;
        float sq = 0.f;
#pragma unroll
        for (int j = 0; j < MAXNV; ++j) if (j < nv) {
            float a = v[j].x - mean, b = v[j].y - mean, c = v[j].z - mean, d = v[j].w - mean;
            sq += (a * a + b * b) + (c * c + d * d);
        }
        const float rstd = rsqrtf(hv_warp_sum(sq) * invD + eps);
        const float4* wp = reinterpret_cast<const float4*>(pass == 0 ? n1w : n2w);
        const float4* bp = reinterpret_cast<const float4*>(pass == 0 ? n1b : n2b);
#pragma unroll
        for (int j = 0; j < MAXNV; ++j) if (j < nv) {
            const int c4 = j * 32 + lane;
            float4 y;
            y.x = (v[j].x - mean) * rstd; y.y = (v[j].y - mean) * rstd;
            y.z = (v[j].z - mean) * rstd; y.w = (v[j].w - mean) * rstd;
            if (wp) { float4 g = wp[c4]; y.x *= g.x; y.y *= g.y; y.z *= g.z; y.w *= g.w; }
            if (bp) { float4 b = bp[c4]; y.x += b.x; y.y += b.y; y.z += b.z; y.w += b.w; }
            v[j] = y;
        }
    }
    if (ada) {
        const float* arow = ada + (long long)frame * ada_ld;
        const float4* sh = reinterpret_cast<const float4*>(arow + shift_col);
        const float4* sc = reinterpret_cast<const float4*>(arow + scale_col);
#pragma unroll
        for (int j = 0; j < MAXNV; ++j) if (j < nv) {
            const int c4 = j * 32 + lane;
            float4 s = sc[c4], t = sh[c4];
            v[j].x = v[j].x * (1.f + s.x) + t.x; v[j].y = v[j].y * (1.f + s.y) + t.y;
            v[j].z = v[j].z * (1.f + s.z) + t.z; v[j].w = v[j].w * (1.f + s.w) + t.w;
        }
    }
    const int O = C * p * p;
    float mine = 0.f;
    for (int o = 0; o < O; ++o) {
        const float4* wr = reinterpret_cast<const float4*>(w + (long long)o * D);
        float acc = 0.f;
#pragma unroll
        for (int j = 0; j < MAXNV; ++j) if (j < nv) {
            float4 g = wr[j * 32 + lane];
            acc = fmaf(g.x, v[j].x, acc); acc = fmaf(g.y, v[j].y, acc);
            acc = fmaf(g.z, v[j].z, acc); acc = fmaf(g.w, v[j].w, acc);
        }
        acc = hv_warp_sum(acc);
        if (lane == o) mine = acc + (bias ? bias[o] : 0.f);
    }
    if (lane < O) {
        const int s = abs_i - map.row_off;
        const int c = lane / (p * p), ij = lane - c * p * p, i = ij / p, j = ij - i * p;
        const int y = (s / Wp) * p + i, x = (s % Wp) * p + j;
        const int Hh = (S / Wp) * p, Ww = Wp * p;
        const long long o = (((long long)frame * C + c) * Hh + y) * Ww + x;
        if (pre) pre[o] = mine;
        if (zt) zt[o] = fmaf(mine, dt, zt[o]);
    }
}

extern "C" int hv_final_head(const float* res, int64_t frames, int64_t frame_rows, int row_off, int S, int D,
                             const float* n1_w, const float* n1_b, const float* n2_w, const float* n2_b, float eps,
                             const float* ada, int64_t ada_ld, int shift_col, int scale_col,
                             const float* w, const float* bias, int C, int p, int Wp,
                             float* pre, float* zt, float dt, void* stream) {
    HV_REQUIRE(D > 0 && D % 128 == 0 && D <= 2048, "hv_final_head: D=%d must be a multiple of 128 and <= 2048", D);
    HV_REQUIRE(C * p * p <= 32 && C > 0 && p > 0, "hv_final_head: C*p*p must be <= 32");
    HV_REQUIRE(Wp > 0 && S % Wp == 0, "hv_final_head: S must be a multiple of Wp");
    HV_REQUIRE(row_off >= 0 && row_off + S <= frame_rows, "hv_final_head: bad row range");
    HV_REQUIRE(!ada || (((shift_col | scale_col) & 3) == 0 && ada_ld % 4 == 0), "hv_final_head: adaLN offsets must be multiples of 4");
    const long long rows = (long long)frames * S;
    if (rows == 0) return HV_OK;
    HvRowMap map{S, (long long)frame_rows, row_off};
    dim3 grid((unsigned)((rows + 7) / 8));
    cudaStream_t st = hv_stream(stream);
    const int nv = D / 128;
#define FH(MAXNV) final_head_kernel<MAXNV><<<grid, 256, 0, st>>>(res, rows, map, S, D, n1_w, n1_b, n2_w, n2_b, eps, ada, ada_ld, \
        shift_col, scale_col, w, bias, C, p, Wp, pre, zt, dt)
    if (nv <= 4) FH(4); else if (nv <= 8) FH(8); else FH(16);
#undef FH
    HV_LAUNCH_CHECK();
    return HV_OK;
}

// ---------------------------------------------------------------------------------------------
// camera_down (model_AMD.py:1788-1793): tiny direct 3x3 conv (pad 1) and 2x2 max-pool.
__global__ void conv3x3_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                               float* __restrict__ y, long long B, int Cin, int Cout, int H, int W, int stride, int Ho, int Wo) {
    const long long total = B * Cout * Ho * Wo;
    for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
        const int ox = (int)(idx % Wo);
        long long t = idx / Wo;
        const int oy = (int)(t % Ho); t /= Ho;
        const int co = (int)(t % Cout);
        const long long b = t / Cout;
        float acc = bias ? bias[co] : 0.f;
        for (int ci = 0; ci < Cin; ++ci) {
            const float* xp = x + (b * Cin + ci) * (long long)H * W;
            const float* wp = w + ((long long)co * Cin + ci) * 9;
#pragma unroll
            for (int ky = 0; ky < 3; ++ky) {
                const int iy = oy * stride - 1 + ky;
                if (iy < 0 || iy >= H) continue;
#pragma unroll
                for (int kx = 0; kx < 3; ++kx) {
                    const int ix = ox * stride - 1 + kx;
                    if (ix < 0 || ix >= W) continue;
                    acc = fmaf(wp[ky * 3 + kx], xp[(long long)iy * W + ix], acc);
                }
            }
        }
        y[idx] = acc;
    }
}
extern "C" int hv_conv3x3(const float* x, const float* w, const float* bias, float* y,
                          int64_t B, int Cin, int Cout, int H, int W, int stride, void* stream) {
    HV_REQUIRE(stride == 1 || stride == 2, "hv_conv3x3: stride must be 1 or 2");
    HV_REQUIRE(B >= 0 && Cin > 0 && Cout > 0 && H > 0 && W > 0, "hv_conv3x3: bad sizes");
    if (B == 0) return HV_OK;
    const int Ho = (H + 2 - 3) / stride + 1, Wo = (W + 2 - 3) / stride + 1;
    long long total = (long long)B * Cout * Ho * Wo;
    int grid = (int)((total + 255) / 256);
    if (grid > 148 * 16) grid = 148 * 16;
    conv3x3_kernel<<<grid, 256, 0, hv_stream(stream)>>>(x, w, bias, y, B, Cin, Cout, H, W, stride, Ho, Wo);
    HV_LAUNCH_CHECK();
    return HV_OK;
}

__global__ void maxpool2_kernel(const float* __restrict__ x, float* __restrict__ y, long long planes, int H, int W) {
    const int Ho = H / 2, Wo = W / 2;
    const long long total = planes * Ho * Wo;
    for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
        const int ox = (int)(idx % Wo);
        long long t = idx / Wo;
        const int oy = (int)(t % Ho);
        const long long pl = t / Ho;
        const float* p0 = x + (pl * H + 2 * oy) * (long long)W + 2 * ox;
        y[idx] = fmaxf(fmaxf(p0[0], p0[1]), fmaxf(p0[W], p0[W + 1]));
    }
}
extern "C" int hv_maxpool2(const float* x, float* y, int64_t planes, int H, int W, void* stream) {
    HV_REQUIRE(planes >= 0 && H >= 2 && W >= 2, "hv_maxpool2: bad sizes");
    if (planes == 0) return HV_OK;
    long long total = (long long)planes * (H / 2) * (W / 2);
    int grid = (int)((total + 255) / 256);
    if (grid > 148 * 16) grid = 148 * 16;
    maxpool2_kernel<<<grid, 256, 0, hv_stream(stream)>>>(x, y, planes, H, W);
    HV_LAUNCH_CHECK();
    return HV_OK;
}

// ---------------------------------------------------------------------------------------------
// RoPE (model/attention.py:6-62): theta_i = base^(-2 i / hd), i = 0..hd/2-1; pair (2i,2i+1) of
// every head at position pos is multiplied by e^{j (start_pos+pos) theta_i}.
__global__ void rope_kernel(float* __restrict__ q, float* __restrict__ k, long long batch, int L, int heads, int hd,
                            long long ld, int start_pos, float base) {
    const int half = hd >> 1;
    const long long total = batch * L * heads * half;
    for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
        const int i = (int)(idx % half);
        long long t = idx / half;
        const int h = (int)(t % heads); t /= heads;
        const int pos = (int)(t % L);
        const long long row = t;  // batch*L + pos
        const float theta = powf(base, -2.0f * (float)i / (float)hd);
        float s, c;
        sincosf((float)(start_pos + pos) * theta, &s, &c);
        const long long o = row * ld + (long long)h * hd + 2 * i;
        float a = q[o], b = q[o + 1];
        q[o] = a * c - b * s; q[o + 1] = a * s + b * c;
        a = k[o]; b = k[o + 1];
        k[o] = a * c - b * s; k[o + 1] = a * s + b * c;
    }
}
extern "C" int hv_rope(float* q, float* k, int64_t batch, int L, int heads, int head_dim, int64_t ld,
                       int start_pos, float base, void* stream) {
    HV_REQUIRE(head_dim > 0 && head_dim % 2 == 0 && batch >= 0 && L >= 0 && heads > 0, "hv_rope: bad sizes");
    long long total = (long long)batch * L * heads * (head_dim / 2);
    if (total == 0) return HV_OK;
    int grid = (int)((total + 255) / 256);
    if (grid > 148 * 16) grid = 148 * 16;
    rope_kernel<<<grid, 256, 0, hv_stream(stream)>>>(q, k, batch, L, heads, head_dim, ld, start_pos, base);
    HV_LAUNCH_CHECK();
    return HV_OK;
}


// tokens (N, h*w, p*p*c) -> image (N, c, h*p, w*p).  order 0: token features are (p, q, c) — `AMDModel*.unpatchify`
// (model_AMD.py:2816-2830) and `MaskedAutoencoderViT.unpatchify` (model_MAE.py:131-145), einsum 'nhwpqc->nchpwq';
// order 1: (c, p, q) — the DiT heads' reshape/permute (transformer.py:2038-2041).  One thread per output pixel,
// consecutive threads walk an output row: coalesced 4-byte stores, reads strided by the feature order (c or 1).
// `inverse` != 0 runs the same index map the other way (patchify, model_MAE.py:116-129).
namespace {
__global__ void unpatchify_kernel(const float* __restrict__ in, float* __restrict__ out, long long total, int c, int h, int w,
                                  int p, int order, int inverse) {
    const int H = h * p, W = w * p;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int x = (int)(i % W), y = (int)((i / W) % H), ch = (int)((i / ((long long)W * H)) % c);
        const long long n = i / ((long long)W * H * c);
        const int ty = y / p, py = y - ty * p, tx = x / p, px = x - tx * p;
        const int feat = order == 0 ? (py * p + px) * c + ch : (ch * p + py) * p + px;
        const long long tok = ((n * h + ty) * w + tx) * (long long)(p * p * c) + feat;
        if (inverse) out[tok] = in[i]; else out[i] = in[tok];
    }
}
}  // namespace

extern "C" int hv_unpatchify(const float* in, float* out, int64_t N, int c, int h, int w, int p, int order, int inverse, void* stream) {
    HV_REQUIRE(N >= 0 && c > 0 && h > 0 && w > 0 && p > 0 && (order == 0 || order == 1), "hv_unpatchify: bad arguments");
    const long long total = (long long)N * c * h * p * w * p;
    if (total == 0) return HV_OK;
    long long grid = (total + 255) / 256;
    if (grid > 148 * 32) grid = 148 * 32;
    unpatchify_kernel<<<(unsigned)grid, 256, 0, hv_stream(stream)>>>(in, out, total, c, h, w, p, order, inverse);
    HV_LAUNCH_CHECK();
    return HV_OK;
}
