// Bandwidth-bound row kernels: LayerNorm/adaLN modulate, timestep sinusoid, elementwise helpers,
// residual-frame scatter/gather.  One warp per token row, 128-bit accesses, shuffle reductions.
#include "hv_common.cuh"

// ---------------------------------------------------------------------------------------------
// y = (LN(x) * w + b) * (1 + scale) + shift          (model/modules.py:881-928, 697, 710)
template <int MAXNV, typename OutT>
__global__ void __launch_bounds__(256) ln_modulate_kernel(
    const float* __restrict__ res, long long rows, HvRowMap map, int D,
    const float* __restrict__ ln_w, const float* __restrict__ ln_b, float eps,
    const float* __restrict__ ada, long long ada_ld, int shift_col_mot, int scale_col_mot,
    int shift_col_img, int scale_col_img, int img_start, int img_len, int cond_div, int cond_mul,
    OutT* __restrict__ out, float* __restrict__ save) {
    const int lane = threadIdx.x & 31;
    const long long r = blockIdx.x * (long long)(blockDim.x >> 5) + (threadIdx.x >> 5);
    if (r >= rows) return;
    const int nv = D >> 7;  // float4 per lane
    int frame, abs_i;
    const long long rr = map.res_row(r, frame, abs_i);
    const float4* xp = reinterpret_cast<const float4*>(res + rr * D);
    float4 v[MAXNV];
    float sum = 0.f;
#pragma unroll
    for (int j = 0; j < MAXNV; ++j)
        if (j < nv) {
            v[j] = xp[j * 32 + lane];
            if (save) reinterpret_cast<float4*>(save + r * D)[j * 32 + lane] = v[j];  // training forward: the rows entering the norm
            sum += (v[j].x + v[j].y) + (v[j].z + v[j].w);
        }
    const float mean = hv_warp_sum(sum) / (float)D;
    float sq = 0.f;
#pragma unroll
    for (int j = 0; j < MAXNV; ++j)
        if (j < nv) {
            float a = v[j].x - mean, b = v[j].y - mean, c = v[j].z - mean, d = v[j].w - mean;
            sq += (a * a + b * b) + (c * c + d * d);
        }
    const float rstd = rsqrtf(hv_warp_sum(sq) / (float)D + eps);
    const float4 *sh = nullptr, *sc = nullptr;
    if (ada != nullptr) {
        const bool img = (abs_i >= img_start) && (abs_i < img_start + img_len);
        const float* arow = ada + (long long)((frame / cond_div) * cond_mul) * ada_ld;
        sh = reinterpret_cast<const float4*>(arow + (img ? shift_col_img : shift_col_mot));
        sc = reinterpret_cast<const float4*>(arow + (img ? scale_col_img : scale_col_mot));
    }
    const float4* wp = reinterpret_cast<const float4*>(ln_w);
    const float4* bp = reinterpret_cast<const float4*>(ln_b);
#pragma unroll
    for (int j = 0; j < MAXNV; ++j)
        if (j < nv) {
            const int c4 = j * 32 + lane;
            float4 y;
            y.x = (v[j].x - mean) * rstd; y.y = (v[j].y - mean) * rstd;
            y.z = (v[j].z - mean) * rstd; y.w = (v[j].w - mean) * rstd;
            if (wp) { float4 w = wp[c4]; y.x *= w.x; y.y *= w.y; y.z *= w.z; y.w *= w.w; }
            if (bp) { float4 b = bp[c4]; y.x += b.x; y.y += b.y; y.z += b.z; y.w += b.w; }
            if (sh) {
                float4 s = sc[c4], t = sh[c4];
                y.x = y.x * (1.f + s.x) + t.x; y.y = y.y * (1.f + s.y) + t.y;
                y.z = y.z * (1.f + s.z) + t.z; y.w = y.w * (1.f + s.w) + t.w;
            }
            if constexpr (sizeof(OutT) == 4) {
                reinterpret_cast<float4*>(out + r * D)[c4] = y;
            } else {
                uint2 pk = make_uint2(hv_pack_bf16(y.x, y.y), hv_pack_bf16(y.z, y.w));
                reinterpret_cast<uint2*>(out + r * D)[c4] = pk;
            }
        }
}

// Two rows per warp, up to 16 consecutive rows per CTA.  With adaLN the CTAs are aligned to (frame, segment):
// a frame's row range splits into [motion | image | motion] segments (any of them may be empty) and a CTA never
// crosses a segment, so its rows share ONE conditioning set.  The CTA folds the LayerNorm affine and the adaLN
// shift / scale once into A = w (1 + scale), B = b (1 + scale) + shift in shared memory and every row is
// y = xhat * A + B: 16 KB of parameter reads per CTA instead of per row (x itself is only 4 KB per row).
// Without adaLN the rows are simply taken 16 at a time (A = w, B = b).
// The row loads are issued first so that their DRAM round trip overlaps the parameter fold.
struct LnSeg {
    int ga, gb, gc;  // CTAs per frame for the three segments (0 when ada == nullptr: flat mapping)
    int a, b, c;     // segment lengths: rows [row_off, img_start), [img_start, img_start+img_len), rest
};

template <int NV, typename OutT>
__global__ void __launch_bounds__(256, 2) ln_modulate2_kernel(
    const float* __restrict__ res, long long rows, HvRowMap map, int D,
    const float* __restrict__ ln_w, const float* __restrict__ ln_b, float eps,
    const float* __restrict__ ada, long long ada_ld, int shift_col_mot, int scale_col_mot,
    int shift_col_img, int scale_col_img, int img_start, int img_len, int cond_div, int cond_mul,
    LnSeg sg, OutT* __restrict__ out, float* __restrict__ save) {
    __shared__ float4 sA[NV * 32], sB[NV * 32];
    asm volatile("griddepcontrol.wait;" ::: "memory");  // PDL (hv_launch_pdl): launched while the previous kernel drains; no early trigger,
                                                        // the grid is larger than one wave (hv_common.cuh / tc_ptx.cuh)
    const int lane = threadIdx.x & 31, k0 = 2 * (threadIdx.x >> 5);
    // ---- which rows does this CTA own?  (dense output row, residual-buffer row, count)
    long long rd, rr;   // first dense row / first residual row of the CTA
    int cnt;            // rows of the CTA (<= 16)
    bool img = false;
    int frame = 0;
    if (ada != nullptr) {
        const int G = sg.ga + sg.gb + sg.gc;
        frame = blockIdx.x / G;
        int g = blockIdx.x - frame * G, abs0, len;
        if (g < sg.ga) { abs0 = map.row_off; len = sg.a; }
        else if (g < sg.ga + sg.gb) { g -= sg.ga; abs0 = img_start; len = sg.b; img = true; }
        else { g -= sg.ga + sg.gb; abs0 = img_start + img_len; len = sg.c; }
        const int i0 = g * 16;
        cnt = len - i0 < 16 ? len - i0 : 16;
        rr = (long long)frame * map.frame_rows + abs0 + i0;
        rd = (long long)frame * map.L + (abs0 - map.row_off) + i0;
    } else {
        rd = blockIdx.x * 16ll;
        cnt = rows - rd < 16 ? (int)(rows - rd) : 16;
        int f, a;
        rr = map.res_row(rd, f, a);
        if (a + cnt > map.row_off + map.L) rr = -1;  // the 16 rows cross a frame of a non-dense range: per-row mapping below
    }
    const bool valid = k0 < cnt, two = k0 + 1 < cnt;
    float4 u[NV], v[NV];
    if (valid) {
        long long ra = rr + k0, rb2 = rr + k0 + 1;
        if (rr < 0) { int f, a; ra = map.res_row(rd + k0, f, a); rb2 = two ? map.res_row(rd + k0 + 1, f, a) : ra; }
        const float4* x0 = reinterpret_cast<const float4*>(res + ra * D);
        const float4* x1 = two ? reinterpret_cast<const float4*>(res + rb2 * D) : x0;
#pragma unroll
        for (int j = 0; j < NV; ++j) { u[j] = x0[j * 32 + lane]; v[j] = x1[j * 32 + lane]; }
    } else {
#pragma unroll
        for (int j = 0; j < NV; ++j) u[j] = v[j] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    {
        const float4* wp = reinterpret_cast<const float4*>(ln_w);
        const float4* bp = reinterpret_cast<const float4*>(ln_b);
        const float4 *shf = nullptr, *scf = nullptr;
        if (ada != nullptr) {
            const float* rowp = ada + (long long)((frame / cond_div) * cond_mul) * ada_ld;
            shf = reinterpret_cast<const float4*>(rowp + (img ? shift_col_img : shift_col_mot));
            scf = reinterpret_cast<const float4*>(rowp + (img ? scale_col_img : scale_col_mot));
        }
        for (int c4 = threadIdx.x; c4 < NV * 32; c4 += 256) {
            const float4 w = wp ? __ldg(wp + c4) : make_float4(1.f, 1.f, 1.f, 1.f);
            const float4 b = bp ? __ldg(bp + c4) : make_float4(0.f, 0.f, 0.f, 0.f);
            const float4 s = scf ? __ldg(scf + c4) : make_float4(0.f, 0.f, 0.f, 0.f), t = shf ? __ldg(shf + c4) : make_float4(0.f, 0.f, 0.f, 0.f);
            sA[c4] = make_float4(w.x * (1.f + s.x), w.y * (1.f + s.y), w.z * (1.f + s.z), w.w * (1.f + s.w));
            sB[c4] = make_float4(fmaf(b.x, 1.f + s.x, t.x), fmaf(b.y, 1.f + s.y, t.y), fmaf(b.z, 1.f + s.z, t.z), fmaf(b.w, 1.f + s.w, t.w));
        }
    }
    const float inv_d = 1.0f / (float)D;
    float su = 0.f, sv = 0.f;
#pragma unroll
    for (int j = 0; j < NV; ++j) { su += (u[j].x + u[j].y) + (u[j].z + u[j].w); sv += (v[j].x + v[j].y) + (v[j].z + v[j].w); }
    const float mu = hv_warp_sum(su) * inv_d, mv = hv_warp_sum(sv) * inv_d;
    float qu = 0.f, qv = 0.f;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
        float a = u[j].x - mu, b = u[j].y - mu, c = u[j].z - mu, d = u[j].w - mu;
        qu += (a * a + b * b) + (c * c + d * d);
        a = v[j].x - mv; b = v[j].y - mv; c = v[j].z - mv; d = v[j].w - mv;
        qv += (a * a + b * b) + (c * c + d * d);
    }
    const float ru = rsqrtf(hv_warp_sum(qu) * inv_d + eps), rv = rsqrtf(hv_warp_sum(qv) * inv_d + eps);
    __syncthreads();  // sA / sB complete (every thread of the CTA reaches this point)
    if (!valid) return;
    auto emit = [&](OutT* o, int c4, float4 y) {
        if constexpr (sizeof(OutT) == 4) reinterpret_cast<float4*>(o)[c4] = y;
        else reinterpret_cast<uint2*>(o)[c4] = make_uint2(hv_pack_bf16(y.x, y.y), hv_pack_bf16(y.z, y.w));
    };
    OutT* o0 = out + (rd + k0) * D;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
        const int c4 = j * 32 + lane;
        const float4 A = sA[c4], B = sB[c4];
        float4 y, z;
        y.x = fmaf((u[j].x - mu) * ru, A.x, B.x); y.y = fmaf((u[j].y - mu) * ru, A.y, B.y);
        y.z = fmaf((u[j].z - mu) * ru, A.z, B.z); y.w = fmaf((u[j].w - mu) * ru, A.w, B.w);
        z.x = fmaf((v[j].x - mv) * rv, A.x, B.x); z.y = fmaf((v[j].y - mv) * rv, A.y, B.y);
        z.z = fmaf((v[j].z - mv) * rv, A.z, B.z); z.w = fmaf((v[j].w - mv) * rv, A.w, B.w);
        emit(o0, c4, y);
        if (two) emit(o0 + D, c4, z);
        if (save) {  // training forward: keep the fp32 rows entering the norm (the backward's x), dense
            reinterpret_cast<float4*>(save + (rd + k0) * D)[c4] = u[j];
            if (two) reinterpret_cast<float4*>(save + (rd + k0 + 1) * D)[c4] = v[j];
        }
    }
}

static int ln_modulate_impl(const float* res, int64_t frames, int64_t frame_rows, int row_off, int L, int D,
                            const float* ln_w, const float* ln_b, float eps,
                            const float* ada, int64_t ada_ld, int shift_col_mot, int scale_col_mot,
                            int shift_col_img, int scale_col_img, int img_start, int img_len,
                            int cond_div, int cond_mul, void* out, int out_dtype, float* save, void* stream) {
    HV_REQUIRE(D > 0 && D % 128 == 0 && D <= 2048, "hv_ln_modulate: D=%d must be a multiple of 128 and <= 2048", D);
    HV_REQUIRE(frames >= 0 && L >= 0 && row_off >= 0 && row_off + L <= frame_rows, "hv_ln_modulate: bad row range");
    HV_REQUIRE(out_dtype == HV_F32 || out_dtype == HV_BF16, "hv_ln_modulate: bad out dtype");
    if (ada) {
        HV_REQUIRE(cond_div > 0, "hv_ln_modulate: cond_div must be > 0");
        HV_REQUIRE(((shift_col_mot | scale_col_mot | shift_col_img | scale_col_img) & 3) == 0 && ada_ld % 4 == 0,
                   "hv_ln_modulate: adaLN column offsets / ld must be multiples of 4");
    }
    const long long rows = (long long)frames * L;
    if (rows == 0) return HV_OK;
    HvRowMap map{L, (long long)frame_rows, row_off};
    const int nv = D / 128;
    cudaStream_t st = hv_stream(stream);
#define LN_ARGS res, rows, map, D, ln_w, ln_b, eps, ada, ada_ld, shift_col_mot, scale_col_mot, shift_col_img, scale_col_img, \
                img_start, img_len, cond_div, cond_mul, sg
    if (nv == 8 || nv == 4) {  // D = 1024 / 512 (the DiT and the encoders): two rows per warp
        LnSeg sg{0, 0, 0, 0, 0, 0};
        long long ctas = (rows + 15) / 16;
        if (ada) {  // segment-aligned CTAs: [row_off, img_start) | [img_start, img_start + img_len) | rest, clipped to the range
            const int lo = row_off, hi = row_off + L;
            const int is = img_start < lo ? lo : (img_start > hi ? hi : img_start);
            const int ie = img_start + img_len < is ? is : (img_start + img_len > hi ? hi : img_start + img_len);
            sg.a = is - lo; sg.b = ie - is; sg.c = hi - ie;
            sg.ga = (sg.a + 15) / 16; sg.gb = (sg.b + 15) / 16; sg.gc = (sg.c + 15) / 16;
            ctas = (long long)frames * (sg.ga + sg.gb + sg.gc);
            img_start = is; img_len = sg.b;  // the kernel derives the segment starts from these
        }
        dim3 grid((unsigned)ctas), block(256);
        if (out_dtype == HV_F32) { if (nv == 8) HV_CUDA(hv_launch_pdl(ln_modulate2_kernel<8, float>, dim3(grid), dim3(block), 0, st, LN_ARGS, (float*)out, save)); else HV_CUDA(hv_launch_pdl(ln_modulate2_kernel<4, float>, dim3(grid), dim3(block), 0, st, LN_ARGS, (float*)out, save)); }
        else { if (nv == 8) HV_CUDA(hv_launch_pdl(ln_modulate2_kernel<8, __nv_bfloat16>, dim3(grid), dim3(block), 0, st, LN_ARGS, (__nv_bfloat16*)out, save)); else HV_CUDA(hv_launch_pdl(ln_modulate2_kernel<4, __nv_bfloat16>, dim3(grid), dim3(block), 0, st, LN_ARGS, (__nv_bfloat16*)out, save)); }
        HV_LAUNCH_CHECK();
        return HV_OK;
    }
    dim3 grid((unsigned)((rows + 7) / 8)), block(256);
#define LN_LAUNCH(MAXNV, T)                                                                                   \
    ln_modulate_kernel<MAXNV, T><<<grid, block, 0, st>>>(res, rows, map, D, ln_w, ln_b, eps, ada, ada_ld,     \
        shift_col_mot, scale_col_mot, shift_col_img, scale_col_img, img_start, img_len, cond_div, cond_mul, (T*)out, save)
    if (out_dtype == HV_F32) {
        if (nv <= 4) LN_LAUNCH(4, float); else if (nv <= 8) LN_LAUNCH(8, float); else LN_LAUNCH(16, float);
    } else {
        if (nv <= 4) LN_LAUNCH(4, __nv_bfloat16); else if (nv <= 8) LN_LAUNCH(8, __nv_bfloat16); else LN_LAUNCH(16, __nv_bfloat16);
    }
#undef LN_LAUNCH
    HV_LAUNCH_CHECK();
    return HV_OK;
}

extern "C" int hv_ln_modulate(const float* res, int64_t frames, int64_t frame_rows, int row_off, int L, int D,
                              const float* ln_w, const float* ln_b, float eps,
                              const float* ada, int64_t ada_ld, int shift_col_mot, int scale_col_mot,
                              int shift_col_img, int scale_col_img, int img_start, int img_len,
                              int cond_div, int cond_mul, void* out, int out_dtype, void* stream) {
    return ln_modulate_impl(res, frames, frame_rows, row_off, L, D, ln_w, ln_b, eps, ada, ada_ld, shift_col_mot, scale_col_mot,
                            shift_col_img, scale_col_img, img_start, img_len, cond_div, cond_mul, out, out_dtype, nullptr, stream);
}
// training forward: the same kernel also writes the fp32 rows it normalises to `save` (dense, frames * L x D) — what the
// backward of the norm needs — instead of a separate gather pass over the residual stream
extern "C" int hv_ln_modulate_save(const float* res, int64_t frames, int64_t frame_rows, int row_off, int L, int D,
                                   const float* ln_w, const float* ln_b, float eps,
                                   const float* ada, int64_t ada_ld, int shift_col_mot, int scale_col_mot,
                                   int shift_col_img, int scale_col_img, int img_start, int img_len,
                                   int cond_div, int cond_mul, void* out, int out_dtype, float* save, void* stream) {
    HV_REQUIRE(save != nullptr, "hv_ln_modulate_save: save is NULL");
    return ln_modulate_impl(res, frames, frame_rows, row_off, L, D, ln_w, ln_b, eps, ada, ada_ld, shift_col_mot, scale_col_mot,
                            shift_col_img, scale_col_img, img_start, img_len, cond_div, cond_mul, out, out_dtype, save, stream);
}

// ---------------------------------------------------------------------------------------------
// Timesteps(dim, flip_sin_to_cos=True, freq_shift=0): [cos | sin], freq_i = exp(-ln(1e4) i / half)
template <typename OutT>
__global__ void timestep_embed_kernel(const float* __restrict__ t, long long rows, int dim, OutT* __restrict__ out) {
    const int half = dim >> 1;
    const long long total = rows * half;
    for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
        const long long r = idx / half;
        const int i = (int)(idx - r * half);
        const float f = expf(-9.210340371976184f * (float)i / (float)half);
        const float a = t[r] * f;
        float s, c;
        sincosf(a, &s, &c);
        hv_st(out + r * dim + i, c);
        hv_st(out + r * dim + half + i, s);
    }
}

extern "C" int hv_timestep_embed(const float* t, int64_t rows, int dim, void* out, int out_dtype, void* stream) {
    HV_REQUIRE(dim > 0 && dim % 2 == 0 && rows >= 0, "hv_timestep_embed: bad sizes");
    if (rows == 0) return HV_OK;
    long long total = rows * (dim / 2);
    int grid = (int)((total + 255) / 256);
    if (grid > 148 * 8) grid = 148 * 8;
    if (out_dtype == HV_F32) timestep_embed_kernel<float><<<grid, 256, 0, hv_stream(stream)>>>(t, rows, dim, (float*)out);
    else if (out_dtype == HV_BF16) timestep_embed_kernel<__nv_bfloat16><<<grid, 256, 0, hv_stream(stream)>>>(t, rows, dim, (__nv_bfloat16*)out);
    else HV_REQUIRE(false, "hv_timestep_embed: bad dtype");
    HV_LAUNCH_CHECK();
    return HV_OK;
}

// ---------------------------------------------------------------------------------------------
template <typename InT, typename OutT, int OP>  // OP 0: copy/convert, 1: silu
__global__ void unary_kernel(const InT* __restrict__ x, OutT* __restrict__ y, long long n) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        float v = hv_ld(x + i);
        if (OP == 1) v = hv_silu_f(v);
        hv_st(y + i, v);
    }
}

// 8 elements per thread, 16-byte accesses on the wider side (n % 8 == 0, 16-byte aligned bases): the weight refresh after
// an optimizer step converts every fp32 master weight to bf16 (4.2 GB of traffic for the default model)
template <typename InT, typename OutT, int OP>
__global__ void unary_vec8_kernel(const InT* __restrict__ x, OutT* __restrict__ y, long long n8) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
        float v[8];
        if constexpr (sizeof(InT) == 4) {
            const float4 a = reinterpret_cast<const float4*>(x)[2 * i], b = reinterpret_cast<const float4*>(x)[2 * i + 1];
            v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
        } else {
            const uint4 a = reinterpret_cast<const uint4*>(x)[i];
            const uint32_t w[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) { v[2 * j] = __uint_as_float(w[j] << 16); v[2 * j + 1] = __uint_as_float(w[j] & 0xffff0000u); }
        }
        if (OP == 1) {
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = hv_silu_f(v[j]);
        }
        if constexpr (sizeof(OutT) == 4) {
            reinterpret_cast<float4*>(y)[2 * i] = make_float4(v[0], v[1], v[2], v[3]);
            reinterpret_cast<float4*>(y)[2 * i + 1] = make_float4(v[4], v[5], v[6], v[7]);
        } else {
            reinterpret_cast<uint4*>(y)[i] = make_uint4(hv_pack_bf16(v[0], v[1]), hv_pack_bf16(v[2], v[3]), hv_pack_bf16(v[4], v[5]), hv_pack_bf16(v[6], v[7]));
        }
    }
}

template <int OP>
static int unary_launch(const void* x, int in_dtype, void* y, int out_dtype, int64_t n, void* stream) {
    HV_REQUIRE(n >= 0, "unary: negative size");
    HV_REQUIRE((in_dtype == HV_F32 || in_dtype == HV_BF16) && (out_dtype == HV_F32 || out_dtype == HV_BF16), "unary: bad dtype");
    if (n == 0) return HV_OK;
    cudaStream_t st = hv_stream(stream);
    const bool vec = n % 8 == 0 && ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y)) & 15) == 0;
    long long work = vec ? n / 8 : n;
    int grid = (int)((work + 255) / 256);
    if (grid > hv_num_sms() * 16) grid = hv_num_sms() * 16;
#define UNARY(K, ...) do { \
        if (in_dtype == HV_F32 && out_dtype == HV_F32) K<float, float, OP><<<grid, 256, 0, st>>>((const float*)x, (float*)y, __VA_ARGS__); \
        else if (in_dtype == HV_F32) K<float, __nv_bfloat16, OP><<<grid, 256, 0, st>>>((const float*)x, (__nv_bfloat16*)y, __VA_ARGS__); \
        else if (out_dtype == HV_F32) K<__nv_bfloat16, float, OP><<<grid, 256, 0, st>>>((const __nv_bfloat16*)x, (float*)y, __VA_ARGS__); \
        else K<__nv_bfloat16, __nv_bfloat16, OP><<<grid, 256, 0, st>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, __VA_ARGS__); } while (0)
    if (vec) UNARY(unary_vec8_kernel, work); else UNARY(unary_kernel, n);
#undef UNARY
    HV_LAUNCH_CHECK();
    return HV_OK;
}
extern "C" int hv_silu(const void* x, int in_dtype, void* y, int out_dtype, int64_t n, void* stream) {
    return unary_launch<1>(x, in_dtype, y, out_dtype, n, stream);
}
extern "C" int hv_convert(const void* x, int in_dtype, void* y, int out_dtype, int64_t n, void* stream) {
    return unary_launch<0>(x, in_dtype, y, out_dtype, n, stream);
}

__global__ void axpy_kernel(float a, const float* __restrict__ x, float* __restrict__ y, long long n) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        y[i] = fmaf(a, x[i], y[i]);
}
extern "C" int hv_axpy(float a, const float* x, float* y, int64_t n, void* stream) {
    HV_REQUIRE(n >= 0, "hv_axpy: negative size");
    if (n == 0) return HV_OK;
    int grid = (int)((n + 255) / 256);
    if (grid > 148 * 16) grid = 148 * 16;
    axpy_kernel<<<grid, 256, 0, hv_stream(stream)>>>(a, x, y, n);
    HV_LAUNCH_CHECK();
    return HV_OK;
}

// ---------------------------------------------------------------------------------------------
// res[f*frame_rows + row_off + i, :] = src[f*src_stride_f + i*src_stride_i, :] (+ add[i, :])
// (row strides in rows of D floats; src_stride_f = 0 broadcasts one token set to every frame)
__global__ void rows_scatter_kernel(const float* __restrict__ src, long long sf, long long si, const float* __restrict__ add,
                                    float* __restrict__ res, long long rows, HvRowMap map, int D) {
    const int d4 = D >> 2;
    const long long total = rows * d4;
    for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
        const long long r = idx / d4;
        const int c = (int)(idx - r * d4);
        int frame, abs_i;
        const long long rr = map.res_row(r, frame, abs_i);
        const int i = abs_i - map.row_off;
        float4 v = reinterpret_cast<const float4*>(src + ((long long)frame * sf + (long long)i * si) * D)[c];
        if (add) {
            float4 a = reinterpret_cast<const float4*>(add + (long long)i * D)[c];
            v.x += a.x; v.y += a.y; v.z += a.z; v.w += a.w;
        }
        reinterpret_cast<float4*>(res + rr * D)[c] = v;
    }
}
extern "C" int hv_rows_scatter(const float* src, int64_t src_stride_f, int64_t src_stride_i, const float* add,
                               float* res, int64_t frames, int64_t frame_rows, int row_off, int L, int D, void* stream) {
    HV_REQUIRE(D > 0 && D % 4 == 0, "hv_rows_scatter: D must be a multiple of 4");
    HV_REQUIRE(frames >= 0 && L >= 0 && row_off >= 0 && row_off + L <= frame_rows, "hv_rows_scatter: bad row range");
    const long long rows = (long long)frames * L;
    if (rows == 0) return HV_OK;
    long long total = rows * (D / 4);
    int grid = (int)((total + 255) / 256);
    if (grid > 148 * 16) grid = 148 * 16;
    HvRowMap map{L, (long long)frame_rows, row_off};
    rows_scatter_kernel<<<grid, 256, 0, hv_stream(stream)>>>(src, src_stride_f, src_stride_i, add, res, rows, map, D);
    HV_LAUNCH_CHECK();
    return HV_OK;
}

__global__ void rows_gather_kernel(const float* __restrict__ res, long long rows, HvRowMap map, int D, float* __restrict__ out) {
    const int d4 = D >> 2;
    const long long total = rows * d4;
    for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
        const long long r = idx / d4;
        const int c = (int)(idx - r * d4);
        int frame, abs_i;
        const long long rr = map.res_row(r, frame, abs_i);
        reinterpret_cast<float4*>(out + r * D)[c] = reinterpret_cast<const float4*>(res + rr * D)[c];
    }
}
extern "C" int hv_rows_gather(const float* res, int64_t frames, int64_t frame_rows, int row_off, int L, int D,
                              float* out, void* stream) {
    HV_REQUIRE(D > 0 && D % 4 == 0, "hv_rows_gather: D must be a multiple of 4");
    HV_REQUIRE(frames >= 0 && L >= 0 && row_off >= 0 && row_off + L <= frame_rows, "hv_rows_gather: bad row range");
    const long long rows = (long long)frames * L;
    if (rows == 0) return HV_OK;
    long long total = rows * (D / 4);
    int grid = (int)((total + 255) / 256);
    if (grid > 148 * 16) grid = 148 * 16;
    HvRowMap map{L, (long long)frame_rows, row_off};
    rows_gather_kernel<<<grid, 256, 0, hv_stream(stream)>>>(res, rows, map, D, out);
    HV_LAUNCH_CHECK();
    return HV_OK;
}
