// Backward of the adaLN-Zero blocks and of the small layers around them (autograd of
// AMDTransformerBlock / BasicDiTBlock, model/modules.py:1065-1109, 1160-1188, reached from
// `accelerator.backward(loss)`, train_amd.py:489).  All kernels are HBM-bound row kernels on the same
// (frames, frame_rows, D) residual layout as the forward; every reduction is two-stage with a fixed grid and a fixed
// summation order, so gradients are bit-reproducible.
//
// A frame's row range [row_off, row_off + L) is split into up to three segments [motion | image | motion]
// (the two modulation sets of AMDLayerNormZero); a CTA owns one chunk of RCH rows of ONE segment of ONE frame, so
// it has a single conditioning set and its column sums are exactly the per-(frame, segment) sums that the
// gradients of shift / scale / gate need.
#include "hv_common.cuh"
#include <type_traits>

namespace {

constexpr int RCH = 32;  // rows per CTA chunk

struct SegPlan {
    int a, b, c;     // segment lengths: motion rows before the image rows, image rows, motion rows after
    int ga, gb, gc;  // chunks per segment
};

SegPlan make_plan(int row_off, int L, int img_start, int img_len) {
    const int lo = row_off, hi = row_off + L;
    const int is = img_start < lo ? lo : (img_start > hi ? hi : img_start);
    int ie = img_start + img_len;
    ie = ie < is ? is : (ie > hi ? hi : ie);
    SegPlan p;
    p.a = is - lo; p.b = ie - is; p.c = hi - ie;
    p.ga = (p.a + RCH - 1) / RCH; p.gb = (p.b + RCH - 1) / RCH; p.gc = (p.c + RCH - 1) / RCH;
    return p;
}

// slot of a frame -> (first row of the chunk inside the frame's range, rows of the chunk, image segment?)
__device__ __forceinline__ void decode_slot(const SegPlan& sp, int slot, int& i0, int& cnt, bool& img) {
    if (slot < sp.ga) { i0 = slot * RCH; cnt = sp.a - i0; img = false; }
    else if (slot < sp.ga + sp.gb) { const int s = slot - sp.ga; i0 = sp.a + s * RCH; cnt = sp.b - s * RCH; img = true; }
    else { const int s = slot - sp.ga - sp.gb; i0 = sp.a + sp.b + s * RCH; cnt = sp.c - s * RCH; img = false; }
    cnt = cnt < RCH ? cnt : RCH;
}

template <typename T> __device__ __forceinline__ float4 ld4(const T* p);
template <> __device__ __forceinline__ float4 ld4<float>(const float* p) { return *reinterpret_cast<const float4*>(p); }
template <> __device__ __forceinline__ float4 ld4<__nv_bfloat16>(const __nv_bfloat16* p) {
    const uint2 t = *reinterpret_cast<const uint2*>(p);
    return make_float4(__uint_as_float(t.x << 16), __uint_as_float(t.x & 0xffff0000u), __uint_as_float(t.y << 16), __uint_as_float(t.y & 0xffff0000u));
}
template <typename T> __device__ __forceinline__ void st4(T* p, float4 v);
template <> __device__ __forceinline__ void st4<float>(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
template <> __device__ __forceinline__ void st4<__nv_bfloat16>(__nv_bfloat16* p, float4 v) {
    *reinterpret_cast<uint2*>(p) = make_uint2(hv_pack_bf16(v.x, v.y), hv_pack_bf16(v.z, v.w));
}

// ---------------------------------------------------------------- gate backward
// forward: res[row] += gate[cond(frame), seg] * y[row]   (hidden = hidden + gate * attn / ff, modules.py:1094-1107)
// backward: dY[row] = gate * dres[row]  (dense, compute dtype), dgate[cond, seg] += sum_rows dres[row] * y[row].
// gate == NULL: plain residual branch (encoder blocks) -> dY = dres rows (gather + convert), no partial sums.
template <typename TY>
__global__ void __launch_bounds__(256) gate_bwd_kernel(const float* __restrict__ dres, long long frame_rows, int row_off, int L, int D,
                                                       const TY* __restrict__ y, const float* __restrict__ gate, long long gate_ld,
                                                       int col_mot, int col_img, int cond_div, int cond_mul, SegPlan sp,
                                                       TY* __restrict__ dY, float* __restrict__ partial) {
    __shared__ float4 sh[256];
    const int G = sp.ga + sp.gb + sp.gc;
    const int frame = blockIdx.x / G, slot = blockIdx.x - frame * G;
    int i0, cnt; bool img;
    decode_slot(sp, slot, i0, cnt, img);
    const int CT = D >> 2, RL = 256 / CT;            // column threads (one float4 each), row lanes
    const int ct = threadIdx.x % CT, rl = threadIdx.x / CT;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (rl < RL) {
        float4 g = make_float4(1.f, 1.f, 1.f, 1.f);
        if (gate) g = __ldg(reinterpret_cast<const float4*>(gate + (long long)((frame / cond_div) * cond_mul) * gate_ld + (img ? col_img : col_mot)) + ct);
        for (int i = rl; i < cnt; i += RL) {
            const long long rr = (long long)frame * frame_rows + row_off + i0 + i, rd = (long long)frame * L + i0 + i;
            const float4 d = *(reinterpret_cast<const float4*>(dres + rr * D) + ct);
            st4<TY>(dY + rd * D + 4 * ct, make_float4(d.x * g.x, d.y * g.y, d.z * g.z, d.w * g.w));
            if (gate) {
                const float4 yv = ld4<TY>(y + rd * D + 4 * ct);
                acc.x = fmaf(d.x, yv.x, acc.x); acc.y = fmaf(d.y, yv.y, acc.y); acc.z = fmaf(d.z, yv.z, acc.z); acc.w = fmaf(d.w, yv.w, acc.w);
            }
        }
    }
    if (!gate) return;
    sh[threadIdx.x] = acc;
    __syncthreads();
    if (threadIdx.x < CT) {
        float4 t = sh[threadIdx.x];
        for (int k = 1; k < RL; ++k) { const float4 u = sh[k * CT + threadIdx.x]; t.x += u.x; t.y += u.y; t.z += u.z; t.w += u.w; }
        reinterpret_cast<float4*>(partial + (long long)blockIdx.x * D)[threadIdx.x] = t;
    }
}

// dtable[group row, gate column of the segment + c] += sum over the group's frames and the segment's chunks
__global__ void __launch_bounds__(256) gate_bwd_final_kernel(const float* __restrict__ partial, long long frames, int D, SegPlan sp,
                                                             int cond_div, int cond_mul, int col_mot, int col_img,
                                                             float* __restrict__ dtable, long long ld) {
    const int c = blockIdx.x * 256 + threadIdx.x;
    if (c >= D) return;
    const int G = sp.ga + sp.gb + sp.gc;
    const long long g = blockIdx.y, f0 = g * cond_div, f1 = f0 + cond_div < frames ? f0 + cond_div : frames;
    float sm = 0.f, si = 0.f;
    for (long long f = f0; f < f1; ++f)
        for (int s = 0; s < G; ++s) {
            const float v = partial[(f * G + s) * D + c];
            if (s >= sp.ga && s < sp.ga + sp.gb) si += v; else sm += v;
        }
    float* row = dtable + g * cond_mul * ld;
    if (col_mot == col_img) row[col_mot + c] += sm + si;
    else {
        if (sp.ga + sp.gc) row[col_mot + c] += sm;
        if (sp.gb) row[col_img + c] += si;
    }
}

// ---------------------------------------------------------------- modulated LayerNorm backward
// forward: y = (xhat * w + b) * (1 + scale) + shift,  xhat = (x - mean) * rstd   (modules.py:881-928)
// With weff = w (1 + scale) (one set per CTA):  dxhat = dy * weff;  dx = rstd (dxhat - mean(dxhat) - xhat mean(dxhat xhat))
// is ACCUMULATED into the residual-stream gradient; per CTA the column sums A = sum dy xhat, B = sum dy go to
// partial[cta][2][D]; the finishing kernels turn them into dscale = A w + B b, dshift = B (per conditioning row) and
// dw = sum A (1 + scale), db = sum B (1 + scale).
template <int NV, typename TDY>
__global__ void __launch_bounds__(256, 2) ln_mod_bwd_kernel(const float* __restrict__ x, const TDY* __restrict__ dy, const float* __restrict__ w,
                                                         float eps, int D, float* __restrict__ dres, long long frame_rows, int row_off, int L,
                                                         const float* __restrict__ ada, long long ada_ld, int scale_col_mot, int scale_col_img,
                                                         int cond_div, int cond_mul, SegPlan sp, float* __restrict__ partial) {
    extern __shared__ __align__(16) float sh[];  // [8 warps][2][D] column-sum partials, then weff[D]
    float* const weff = sh + 16 * D;
    const int G = sp.ga + sp.gb + sp.gc;
    const int frame = blockIdx.x / G, slot = blockIdx.x - frame * G;
    int i0, cnt; bool img;
    decode_slot(sp, slot, i0, cnt, img);
    for (int i = threadIdx.x; i < 16 * D; i += 256) sh[i] = 0.f;
    {
        const float* sc = ada ? ada + (long long)((frame / cond_div) * cond_mul) * ada_ld + (img ? scale_col_img : scale_col_mot) : nullptr;
        for (int i = threadIdx.x; i < D; i += 256) weff[i] = (w ? w[i] : 1.f) * (1.f + (sc ? sc[i] : 0.f));
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float inv_d = 1.0f / (float)D;
    float4* const accw = reinterpret_cast<float4*>(sh + (size_t)warp * 2 * D) + lane;
    float4* const accb = reinterpret_cast<float4*>(sh + (size_t)warp * 2 * D + D) + lane;
    for (int i = warp; i < cnt; i += 8) {
        const long long rr = (long long)frame * frame_rows + row_off + i0 + i, rd = (long long)frame * L + i0 + i;
        float xv[NV * 4], gv[NV * 4];
        float4 dv4[NV];
        float s = 0.f;
#pragma unroll
        for (int j = 0; j < NV; ++j) dv4[j] = ld4<TDY>(dy + rd * D + 4 * (j * 32 + lane));
#pragma unroll
        for (int j = 0; j < NV; ++j) asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const float4*>(dres + rr * D) + j * 32 + lane));
#pragma unroll
        for (int j = 0; j < NV; ++j) {
            const float4 v = reinterpret_cast<const float4*>(x + rd * D)[j * 32 + lane];
            xv[4 * j] = v.x; xv[4 * j + 1] = v.y; xv[4 * j + 2] = v.z; xv[4 * j + 3] = v.w;
            s += (v.x + v.y) + (v.z + v.w);
        }
        const float mean = hv_warp_sum(s) * inv_d;
        float q = 0.f;
#pragma unroll
        for (int j = 0; j < NV * 4; ++j) { xv[j] -= mean; q += xv[j] * xv[j]; }
        const float rstd = rsqrtf(hv_warp_sum(q) * inv_d + eps);
        float c1 = 0.f, c2 = 0.f;
#pragma unroll
        for (int j = 0; j < NV; ++j) {
            const float dv[4] = {dv4[j].x, dv4[j].y, dv4[j].z, dv4[j].w};
            const float4 wv4 = reinterpret_cast<const float4*>(weff)[j * 32 + lane];
            const float wv[4] = {wv4.x, wv4.y, wv4.z, wv4.w};
            float dwv[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float xh = xv[4 * j + e] * rstd, g = dv[e] * wv[e];
                xv[4 * j + e] = xh; gv[4 * j + e] = g;
                c1 += g; c2 += g * xh;
                dwv[e] = dv[e] * xh;
            }
            float4 aw = accw[j * 32], ab = accb[j * 32];
            aw.x += dwv[0]; aw.y += dwv[1]; aw.z += dwv[2]; aw.w += dwv[3];
            ab.x += dv[0]; ab.y += dv[1]; ab.z += dv[2]; ab.w += dv[3];
            accw[j * 32] = aw; accb[j * 32] = ab;
        }
        c1 = hv_warp_sum(c1) * inv_d; c2 = hv_warp_sum(c2) * inv_d;
#pragma unroll
        for (int j = 0; j < NV; ++j) {
            float4* dst = reinterpret_cast<float4*>(dres + rr * D) + j * 32 + lane;
            float4 o = *dst;
            o.x += rstd * (gv[4 * j] - c1 - xv[4 * j] * c2); o.y += rstd * (gv[4 * j + 1] - c1 - xv[4 * j + 1] * c2);
            o.z += rstd * (gv[4 * j + 2] - c1 - xv[4 * j + 2] * c2); o.w += rstd * (gv[4 * j + 3] - c1 - xv[4 * j + 3] * c2);
            *dst = o;
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 2 * D; i += 256) {
        float t = 0.f;
#pragma unroll
        for (int wv = 0; wv < 8; ++wv) t += sh[(size_t)wv * 2 * D + i];
        partial[(long long)blockIdx.x * 2 * D + i] = t;
    }
}

// one thread per (conditioning group, column): sums the group's CTAs per segment, accumulates dscale / dshift into the
// table gradient and writes the group's contribution to dw / db to gpart[group][2][D]
__global__ void __launch_bounds__(128) ln_mod_bwd_final_kernel(const float* __restrict__ partial, long long frames, int D, SegPlan sp,
                                                               const float* __restrict__ w, const float* __restrict__ b,
                                                               const float* __restrict__ ada, float* __restrict__ dtable, long long ld,
                                                               int shift_col_mot, int scale_col_mot, int shift_col_img, int scale_col_img,
                                                               int cond_div, int cond_mul, float* __restrict__ gpart) {
    const int c = blockIdx.x * 128 + threadIdx.x;
    if (c >= D) return;
    const int G = sp.ga + sp.gb + sp.gc;
    const long long g = blockIdx.y, f0 = g * cond_div, f1 = f0 + cond_div < frames ? f0 + cond_div : frames;
    float Am = 0.f, Bm = 0.f, Ai = 0.f, Bi = 0.f;
    for (long long f = f0; f < f1; ++f)
        for (int s = 0; s < G; ++s) {
            const float* p = partial + (f * G + s) * 2 * D;
            const float a = p[c], bb = p[D + c];
            if (s >= sp.ga && s < sp.ga + sp.gb) { Ai += a; Bi += bb; } else { Am += a; Bm += bb; }
        }
    float dwg = Am + Ai, dbg = Bm + Bi;
    if (ada) {
        const long long ro = g * cond_mul * ld;
        const float wv = w ? w[c] : 1.f, bv = b ? b[c] : 0.f;
        const float sm = ada[ro + scale_col_mot + c], si = ada[ro + scale_col_img + c];
        if (scale_col_mot == scale_col_img) {
            dtable[ro + scale_col_mot + c] += (Am + Ai) * wv + (Bm + Bi) * bv;
            dtable[ro + shift_col_mot + c] += Bm + Bi;
        } else {
            if (sp.ga + sp.gc) { dtable[ro + scale_col_mot + c] += Am * wv + Bm * bv; dtable[ro + shift_col_mot + c] += Bm; }
            if (sp.gb) { dtable[ro + scale_col_img + c] += Ai * wv + Bi * bv; dtable[ro + shift_col_img + c] += Bi; }
        }
        dwg = Am * (1.f + sm) + Ai * (1.f + si);
        dbg = Bm * (1.f + sm) + Bi * (1.f + si);
    }
    gpart[g * 2 * D + c] = dwg;
    gpart[g * 2 * D + D + c] = dbg;
}

// out_a[n] += sum_s part[s][n] (n < split), out_b[n - split] += ... (n >= split): accumulating 8-wide column reduction
__global__ void __launch_bounds__(256) accum_final2_kernel(const float* __restrict__ part, long long S, int N, int split,
                                                           float* __restrict__ out_a, float* __restrict__ out_b) {
    __shared__ float sh[8][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5, n = blockIdx.x * 32 + tx;
    float a = 0.f;
    if (n < N)
        for (long long i = ty; i < S; i += 8) a += part[i * N + n];
    sh[ty][tx] = a;
    __syncthreads();
    if (ty == 0 && n < N) {
        float t = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) t += sh[i][tx];
        if (n < split) { if (out_a) out_a[n] += t; }
        else if (out_b) out_b[n - split] += t;
    }
}

// ---------------------------------------------------------------- forward of the gated residual (training forward)
template <typename TY>
__global__ void gate_residual_kernel(float* __restrict__ res, long long rows, HvRowMap map, int D, const TY* __restrict__ y,
                                     const float* __restrict__ gate, long long gate_ld, int col_mot, int col_img,
                                     int img_start, int img_len, int cond_div, int cond_mul) {
    const int d4 = D >> 2;
    const long long total = rows * d4;
    for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
        const long long r = idx / d4;
        const int c = (int)(idx - r * d4);
        int frame, abs_i;
        const long long rr = map.res_row(r, frame, abs_i);
        float4 g = make_float4(1.f, 1.f, 1.f, 1.f);
        if (gate) {
            const bool img = abs_i >= img_start && abs_i < img_start + img_len;
            g = __ldg(reinterpret_cast<const float4*>(gate + (long long)((frame / cond_div) * cond_mul) * gate_ld + (img ? col_img : col_mot)) + c);
        }
        const float4 yv = ld4<TY>(y + r * D + 4 * c);
        float4* o = reinterpret_cast<float4*>(res + rr * D) + c;
        float4 v = *o;
        v.x = fmaf(g.x, yv.x, v.x); v.y = fmaf(g.y, yv.y, v.y); v.z = fmaf(g.z, yv.z, v.z); v.w = fmaf(g.w, yv.w, v.w);
        *o = v;
    }
}

// ---------------------------------------------------------------- small elementwise / reduction helpers
template <typename TDY, typename TO>
__global__ void silu_bwd_kernel(const float* __restrict__ x, const TDY* __restrict__ dy, TO* __restrict__ dx, long long n) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float v = x[i], s = 1.f / (1.f + __expf(-v));
        hv_st(dx + i, hv_ld(dy + i) * s * (1.f + v * (1.f - s)));
    }
}

__global__ void mse_bwd_kernel(const float* __restrict__ pre, const float* __restrict__ gt, long long n, const float* __restrict__ gout,
                               float coef, float* __restrict__ dpre) {
    const float k = coef * (gout ? gout[0] : 1.f);
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        dpre[i] = k * (pre[i] - gt[i]);
}

// out[o][c] (+)= sum_{a < A} sum_{b < B} src[(base + o*so + a*sa + b*sb) * D + c]   (fixed order)
__global__ void rows_reduce_kernel(const float* __restrict__ src, int D, long long n_out, long long base, long long so, long long A,
                                   long long sa, long long B, long long sb, float* __restrict__ out, int accumulate) {
    const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (idx >= n_out * D) return;
    const long long o = idx / D;
    const int c = (int)(idx - o * D);
    float t = 0.f;
    for (long long a = 0; a < A; ++a)
        for (long long b = 0; b < B; ++b) t += src[(base + o * so + a * sa + b * sb) * D + c];
    out[idx] = accumulate ? out[idx] + t : t;
}

// dst rows picked by index: out[(f*keep + j), :] = src[(f*S + idx[f*keep + j]), :]  (random_masking's torch.gather,
// transformer.py:181-183); scatter != 0 is its backward: dsrc[(f*S + idx), :] = dout[(f*keep + j), :] (dsrc pre-zeroed)
__global__ void rows_index_kernel(const float* __restrict__ in, float* __restrict__ out, const int* __restrict__ idx, long long frames,
                                  int S, int keep, int D, long long in_fs, long long out_fs, int scatter) {
    const int d4 = D >> 2;
    const long long total = frames * keep * d4;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long r = i / d4;
        const int c = (int)(i - r * d4);
        const long long f = r / keep;
        const int j = (int)(r - f * keep), s = idx[r];
        if (!scatter) reinterpret_cast<float4*>(out + (f * out_fs + j) * D)[c] = reinterpret_cast<const float4*>(in + (f * in_fs + s) * D)[c];
        else reinterpret_cast<float4*>(out + (f * out_fs + s) * D)[c] = reinterpret_cast<const float4*>(in + (f * in_fs + j) * D)[c];
    }
}

// ---------------------------------------------------------------- camera_down backward (model_AMD.py:1788-1793)
// direct 3x3 convolution, zero padding 1: input gradient
__global__ void conv3x3_bwd_x_kernel(const float* __restrict__ dy, const float* __restrict__ w, float* __restrict__ dx, long long B,
                                     int Cin, int Cout, int H, int W, int stride, int Ho, int Wo) {
    const long long total = B * Cin * H * W;
    for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
        const int ix = (int)(idx % W);
        long long t = idx / W;
        const int iy = (int)(t % H); t /= H;
        const int ci = (int)(t % Cin);
        const long long b = t / Cin;
        float acc = 0.f;
        for (int co = 0; co < Cout; ++co) {
            const float* dp = dy + (b * Cout + co) * (long long)Ho * Wo;
            const float* wp = w + ((long long)co * Cin + ci) * 9;
            for (int ky = 0; ky < 3; ++ky) {
                const int ty = iy + 1 - ky;
                if (ty < 0 || ty % stride) continue;
                const int oy = ty / stride;
                if (oy >= Ho) continue;
                for (int kx = 0; kx < 3; ++kx) {
                    const int tx = ix + 1 - kx;
                    if (tx < 0 || tx % stride) continue;
                    const int ox = tx / stride;
                    if (ox >= Wo) continue;
                    acc = fmaf(wp[ky * 3 + kx], dp[(long long)oy * Wo + ox], acc);
                }
            }
        }
        dx[idx] = acc;
    }
}
// weight / bias gradient: one CTA per weight element (co, ci, ky, kx), and Cout extra CTAs for the bias; accumulating
__global__ void __launch_bounds__(256) conv3x3_bwd_w_kernel(const float* __restrict__ x, const float* __restrict__ dy, float* __restrict__ dw,
                                                            float* __restrict__ db, long long B, int Cin, int Cout, int H, int W,
                                                            int stride, int Ho, int Wo) {
    __shared__ float sh[256];
    const int nW = Cout * Cin * 9;
    const bool bias = (int)blockIdx.x >= nW;
    int co, ci = 0, ky = 0, kx = 0;
    if (bias) co = blockIdx.x - nW;
    else { int t = blockIdx.x; kx = t % 3; t /= 3; ky = t % 3; t /= 3; ci = t % Cin; co = t / Cin; }
    const long long total = B * Ho * Wo;
    float acc = 0.f;
    for (long long i = threadIdx.x; i < total; i += 256) {
        const int ox = (int)(i % Wo);
        long long t = i / Wo;
        const int oy = (int)(t % Ho);
        const long long b = t / Ho;
        const float d = dy[((b * Cout + co) * Ho + oy) * (long long)Wo + ox];
        if (bias) acc += d;
        else {
            const int iy = oy * stride - 1 + ky, ix = ox * stride - 1 + kx;
            if (iy >= 0 && iy < H && ix >= 0 && ix < W) acc = fmaf(d, x[((b * Cin + ci) * H + iy) * (long long)W + ix], acc);
        }
    }
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if ((int)threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        if (bias) { if (db) db[co] += sh[0]; }
        else if (dw) dw[blockIdx.x] += sh[0];
    }
}
// MaxPool2d(2) backward: the gradient goes to the first maximum in (row, column) scan order, as ATen records it
__global__ void maxpool2_bwd_kernel(const float* __restrict__ x, const float* __restrict__ dy, float* __restrict__ dx, long long planes, int H, int W) {
    const int Ho = H / 2, Wo = W / 2;
    const long long total = planes * Ho * Wo;
    for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
        const int ox = (int)(idx % Wo);
        long long t = idx / Wo;
        const int oy = (int)(t % Ho);
        const long long pl = t / Ho;
        const long long o = (pl * H + 2 * oy) * (long long)W + 2 * ox;
        const float v[4] = {x[o], x[o + 1], x[o + W], x[o + W + 1]};
        int best = 0;
#pragma unroll
        for (int k = 1; k < 4; ++k) if (v[k] > v[best]) best = k;
        const float d = dy[idx];
        dx[o] = best == 0 ? d : 0.f; dx[o + 1] = best == 1 ? d : 0.f; dx[o + W] = best == 2 ? d : 0.f; dx[o + W + 1] = best == 3 ? d : 0.f;
    }
}

int grid_for(long long total) {
    long long g = (total + 255) / 256;
    const long long cap = (long long)hv_num_sms() * 16;
    return (int)(g > cap ? cap : (g < 1 ? 1 : g));
}

}  // namespace

extern "C" int hv_train_chunk_rows(void) { return RCH; }

extern "C" int hv_gate_bwd(const float* dres, int64_t frames, int64_t frame_rows, int row_off, int L, int D, const void* y,
                           const float* gate, int64_t gate_ld, int gate_col_mot, int gate_col_img, int img_start, int img_len,
                           int cond_div, int cond_mul, void* dY, int dtype, float* dtable, float* workspace, int64_t ws_floats,
                           void* stream) {
    HV_REQUIRE(D > 0 && D % 4 == 0 && D <= 1024, "hv_gate_bwd: D=%d must be a multiple of 4 and <= 1024", D);
    HV_REQUIRE(frames >= 0 && L >= 0 && row_off >= 0 && row_off + L <= frame_rows, "hv_gate_bwd: bad row range");
    HV_REQUIRE(dres && dY && (dtype == HV_F32 || dtype == HV_BF16), "hv_gate_bwd: bad arguments");
    if (frames == 0 || L == 0) return HV_OK;
    if (!gate) { img_start = 0; img_len = 0; cond_div = cond_mul = 1; }
    HV_REQUIRE(cond_div > 0, "hv_gate_bwd: cond_div must be > 0");
    const SegPlan sp = make_plan(row_off, L, img_start, img_len);
    const int G = sp.ga + sp.gb + sp.gc;
    if (gate) {
        HV_REQUIRE(y && dtable && workspace && ws_floats >= frames * G * D, "hv_gate_bwd: workspace of %lld floats needed", (long long)frames * G * D);
        HV_REQUIRE(gate_ld % 4 == 0 && ((gate_col_mot | gate_col_img) & 3) == 0 && (reinterpret_cast<uintptr_t>(gate) & 15) == 0, "hv_gate_bwd: gate table must allow 16-byte accesses");
    }
    cudaStream_t st = hv_stream(stream);
    const unsigned ctas = (unsigned)(frames * G);
    if (dtype == HV_F32) gate_bwd_kernel<float><<<ctas, 256, 0, st>>>(dres, frame_rows, row_off, L, D, (const float*)y, gate, gate_ld, gate_col_mot, gate_col_img, cond_div, cond_mul, sp, (float*)dY, workspace);
    else gate_bwd_kernel<__nv_bfloat16><<<ctas, 256, 0, st>>>(dres, frame_rows, row_off, L, D, (const __nv_bfloat16*)y, gate, gate_ld, gate_col_mot, gate_col_img, cond_div, cond_mul, sp, (__nv_bfloat16*)dY, workspace);
    HV_LAUNCH_CHECK();
    if (gate) {
        const long long groups = (frames + cond_div - 1) / cond_div;
        gate_bwd_final_kernel<<<dim3((D + 255) / 256, (unsigned)groups), 256, 0, st>>>(workspace, frames, D, sp, cond_div, cond_mul, gate_col_mot, gate_col_img, dtable, gate_ld);
        HV_LAUNCH_CHECK();
    }
    return HV_OK;
}

extern "C" int hv_ln_modulate_bwd(const float* x, const void* dy, int dy_dtype, const float* ln_w, const float* ln_b, float eps,
                                  float* dres, int64_t frames, int64_t frame_rows, int row_off, int L, int D, const float* ada,
                                  int64_t ada_ld, int shift_col_mot, int scale_col_mot, int shift_col_img, int scale_col_img,
                                  int img_start, int img_len, int cond_div, int cond_mul, float* dtable, float* dw, float* db,
                                  float* workspace, int64_t ws_floats, void* stream) {
    HV_REQUIRE(D > 0 && D % 128 == 0 && D <= 1024, "hv_ln_modulate_bwd: D=%d must be a multiple of 128 and <= 1024", D);
    HV_REQUIRE(frames >= 0 && L >= 0 && row_off >= 0 && row_off + L <= frame_rows, "hv_ln_modulate_bwd: bad row range");
    HV_REQUIRE(x && dy && dres && workspace && (dy_dtype == HV_F32 || dy_dtype == HV_BF16), "hv_ln_modulate_bwd: bad arguments");
    if (frames == 0 || L == 0) return HV_OK;
    if (!ada) { img_start = 0; img_len = 0; cond_div = 8; cond_mul = 1; }
    HV_REQUIRE(cond_div > 0 && (!ada || dtable), "hv_ln_modulate_bwd: cond_div must be > 0 and dtable given with ada");
    const SegPlan sp = make_plan(row_off, L, img_start, img_len);
    const int G = sp.ga + sp.gb + sp.gc;
    const long long groups = (frames + cond_div - 1) / cond_div;
    HV_REQUIRE(ws_floats >= (frames * G + groups) * 2 * D, "hv_ln_modulate_bwd: workspace of %lld floats needed", (long long)(frames * G + groups) * 2 * D);
    float* gpart = workspace + frames * G * 2 * D;
    cudaStream_t st = hv_stream(stream);
    const size_t smem = (size_t)17 * D * sizeof(float);
    const unsigned ctas = (unsigned)(frames * G);
#define LMB(NV, T) do { HV_CUDA(cudaFuncSetAttribute(ln_mod_bwd_kernel<NV, T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        ln_mod_bwd_kernel<NV, T><<<ctas, 256, smem, st>>>(x, (const T*)dy, ln_w, eps, D, dres, frame_rows, row_off, L, ada, ada_ld, scale_col_mot, scale_col_img, cond_div, cond_mul, sp, workspace); } while (0)
#define LMB_T(NV) do { if (dy_dtype == HV_F32) LMB(NV, float); else LMB(NV, __nv_bfloat16); } while (0)
    switch (D / 128) {
        case 1: LMB_T(1); break; case 2: LMB_T(2); break; case 3: LMB_T(3); break; case 4: LMB_T(4); break;
        case 5: LMB_T(5); break; case 6: LMB_T(6); break; case 7: LMB_T(7); break; default: LMB_T(8); break;
    }
#undef LMB_T
#undef LMB
    HV_LAUNCH_CHECK();
    if (ada || dw || db) {
        ln_mod_bwd_final_kernel<<<dim3((D + 127) / 128, (unsigned)groups), 128, 0, st>>>(workspace, frames, D, sp, ln_w, ln_b, ada, dtable, ada_ld,
            shift_col_mot, scale_col_mot, shift_col_img, scale_col_img, cond_div, cond_mul, gpart);
        HV_LAUNCH_CHECK();
    }
    if (dw || db) {
        accum_final2_kernel<<<(2 * D + 31) / 32, 256, 0, st>>>(gpart, groups, 2 * D, D, dw, db);
        HV_LAUNCH_CHECK();
    }
    return HV_OK;
}

extern "C" int hv_gate_residual(float* res, int64_t frames, int64_t frame_rows, int row_off, int L, int D, const void* y, int dtype,
                                const float* gate, int64_t gate_ld, int gate_col_mot, int gate_col_img, int img_start, int img_len,
                                int cond_div, int cond_mul, void* stream) {
    HV_REQUIRE(D > 0 && D % 4 == 0, "hv_gate_residual: D must be a multiple of 4");
    HV_REQUIRE(frames >= 0 && L >= 0 && row_off >= 0 && row_off + L <= frame_rows, "hv_gate_residual: bad row range");
    HV_REQUIRE(res && y && (dtype == HV_F32 || dtype == HV_BF16) && (!gate || cond_div > 0), "hv_gate_residual: bad arguments");
    const long long rows = (long long)frames * L;
    if (rows == 0) return HV_OK;
    HvRowMap map{L, (long long)frame_rows, row_off};
    const int grid = grid_for(rows * (D / 4));
    if (dtype == HV_F32) gate_residual_kernel<float><<<grid, 256, 0, hv_stream(stream)>>>(res, rows, map, D, (const float*)y, gate, gate_ld, gate_col_mot, gate_col_img, img_start, img_len, cond_div, cond_mul);
    else gate_residual_kernel<__nv_bfloat16><<<grid, 256, 0, hv_stream(stream)>>>(res, rows, map, D, (const __nv_bfloat16*)y, gate, gate_ld, gate_col_mot, gate_col_img, img_start, img_len, cond_div, cond_mul);
    HV_LAUNCH_CHECK();
    return HV_OK;
}

extern "C" int hv_silu_bwd(const float* x, const void* dy, int dy_dtype, void* dx, int dx_dtype, int64_t n, void* stream) {
    HV_REQUIRE(n >= 0 && x && dy && dx, "hv_silu_bwd: bad arguments");
    if (n == 0) return HV_OK;
    const int grid = grid_for(n);
    cudaStream_t st = hv_stream(stream);
    if (dy_dtype == HV_F32 && dx_dtype == HV_F32) silu_bwd_kernel<float, float><<<grid, 256, 0, st>>>(x, (const float*)dy, (float*)dx, n);
    else if (dy_dtype == HV_F32 && dx_dtype == HV_BF16) silu_bwd_kernel<float, __nv_bfloat16><<<grid, 256, 0, st>>>(x, (const float*)dy, (__nv_bfloat16*)dx, n);
    else if (dy_dtype == HV_BF16 && dx_dtype == HV_F32) silu_bwd_kernel<__nv_bfloat16, float><<<grid, 256, 0, st>>>(x, (const __nv_bfloat16*)dy, (float*)dx, n);
    else if (dy_dtype == HV_BF16 && dx_dtype == HV_BF16) silu_bwd_kernel<__nv_bfloat16, __nv_bfloat16><<<grid, 256, 0, st>>>(x, (const __nv_bfloat16*)dy, (__nv_bfloat16*)dx, n);
    else HV_REQUIRE(false, "hv_silu_bwd: bad dtype");
    HV_LAUNCH_CHECK();
    return HV_OK;
}

extern "C" int hv_mse_bwd(const float* pre, const float* gt, int64_t n, const float* gout, float* dpre, void* stream) {
    HV_REQUIRE(n >= 0 && pre && gt && dpre, "hv_mse_bwd: bad arguments");
    if (n == 0) return HV_OK;
    mse_bwd_kernel<<<grid_for(n), 256, 0, hv_stream(stream)>>>(pre, gt, n, gout, 2.0f / (float)n, dpre);
    HV_LAUNCH_CHECK();
    return HV_OK;
}

extern "C" int hv_rows_reduce(const float* src, int D, int64_t n_out, int64_t base, int64_t stride_out, int64_t A, int64_t stride_a,
                              int64_t B, int64_t stride_b, float* out, int accumulate, void* stream) {
    HV_REQUIRE(D > 0 && n_out >= 0 && A >= 0 && B >= 0 && src && out, "hv_rows_reduce: bad arguments");
    if (n_out == 0) return HV_OK;
    const long long total = n_out * D;
    rows_reduce_kernel<<<(unsigned)((total + 255) / 256), 256, 0, hv_stream(stream)>>>(src, D, n_out, base, stride_out, A, stride_a, B, stride_b, out, accumulate);
    HV_LAUNCH_CHECK();
    return HV_OK;
}

extern "C" int hv_rows_index(const float* in, float* out, const int* idx, int64_t frames, int S, int keep, int D, int64_t in_frame_rows,
                             int64_t out_frame_rows, int scatter, void* stream) {
    HV_REQUIRE(D > 0 && D % 4 == 0 && frames >= 0 && keep >= 0 && keep <= S && in && out && idx, "hv_rows_index: bad arguments");
    if (frames == 0 || keep == 0) return HV_OK;
    rows_index_kernel<<<grid_for(frames * keep * (D / 4)), 256, 0, hv_stream(stream)>>>(in, out, idx, frames, S, keep, D, in_frame_rows, out_frame_rows, scatter);
    HV_LAUNCH_CHECK();
    return HV_OK;
}

extern "C" int hv_conv3x3_bwd(const float* x, const float* w, const float* dy, float* dx, float* dw, float* db, int64_t B, int Cin,
                              int Cout, int H, int W, int stride, void* stream) {
    HV_REQUIRE(stride == 1 || stride == 2, "hv_conv3x3_bwd: stride must be 1 or 2");
    HV_REQUIRE(B >= 0 && Cin > 0 && Cout > 0 && H > 0 && W > 0 && x && w && dy, "hv_conv3x3_bwd: bad arguments");
    if (B == 0) return HV_OK;
    const int Ho = (H + 2 - 3) / stride + 1, Wo = (W + 2 - 3) / stride + 1;
    cudaStream_t st = hv_stream(stream);
    if (dx) {
        conv3x3_bwd_x_kernel<<<grid_for(B * Cin * H * W), 256, 0, st>>>(dy, w, dx, B, Cin, Cout, H, W, stride, Ho, Wo);
        HV_LAUNCH_CHECK();
    }
    if (dw || db) {
        conv3x3_bwd_w_kernel<<<Cout * Cin * 9 + Cout, 256, 0, st>>>(x, dy, dw, db, B, Cin, Cout, H, W, stride, Ho, Wo);
        HV_LAUNCH_CHECK();
    }
    return HV_OK;
}

extern "C" int hv_maxpool2_bwd(const float* x, const float* dy, float* dx, int64_t planes, int H, int W, void* stream) {
    HV_REQUIRE(planes >= 0 && H >= 2 && W >= 2 && H % 2 == 0 && W % 2 == 0 && x && dy && dx, "hv_maxpool2_bwd: bad arguments (even H, W)");
    if (planes == 0) return HV_OK;
    maxpool2_bwd_kernel<<<grid_for(planes * (H / 2) * (W / 2)), 256, 0, hv_stream(stream)>>>(x, dy, dx, planes, H, W);
    HV_LAUNCH_CHECK();
    return HV_OK;
}
