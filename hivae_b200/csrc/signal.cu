// Haar DWT / IDWT (model/wavelet.py) and the 3-D Gaussian frequency split
// (model/frequency_utils.py) — HBM-bound, 128-bit vectorised kernels.
#include <cstdlib>
#include <type_traits>
#include "hv_common.cuh"

// ---------------------------------------------------------------------------------------------
// The reference evaluates every add/sub in the tensor's own dtype; for bf16 that means a
// rounding after each op.  R<T>() reproduces that so results are bit-exact in both dtypes.
template <typename T> __device__ __forceinline__ float hv_rnd(float v);
template <> __device__ __forceinline__ float hv_rnd<float>(float v) { return v; }
template <> __device__ __forceinline__ float hv_rnd<__nv_bfloat16>(float v) { return __bfloat162float(__float2bfloat16(v)); }

template <typename T> struct Vec16 { static constexpr int N = 16 / sizeof(T); };

template <typename T>
__device__ __forceinline__ void load16(const T* p, float (&v)[16 / sizeof(T)]) {
    uint4 raw = *reinterpret_cast<const uint4*>(p);
    if constexpr (sizeof(T) == 4) {
        v[0] = __uint_as_float(raw.x); v[1] = __uint_as_float(raw.y);
        v[2] = __uint_as_float(raw.z); v[3] = __uint_as_float(raw.w);
    } else {
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
        for (int i = 0; i < 4; ++i) { float2 f = __bfloat1622float2(h[i]); v[2 * i] = f.x; v[2 * i + 1] = f.y; }
    }
}
template <typename T>
__device__ __forceinline__ void store16(T* p, const float (&v)[16 / sizeof(T)]) {
    uint4 raw;
    if constexpr (sizeof(T) == 4) {
        raw.x = __float_as_uint(v[0]); raw.y = __float_as_uint(v[1]);
        raw.z = __float_as_uint(v[2]); raw.w = __float_as_uint(v[3]);
    } else {
        raw.x = hv_pack_bf16(v[0], v[1]); raw.y = hv_pack_bf16(v[2], v[3]);
        raw.z = hv_pack_bf16(v[4], v[5]); raw.w = hv_pack_bf16(v[6], v[7]);
    }
    *reinterpret_cast<uint4*>(p) = raw;
}

// one 2x2 Haar butterfly, op order identical to wavelet.py:15-24
template <typename T>
__device__ __forceinline__ void haar_fwd(float p00, float p10, float p01, float p11,
                                         float& ll, float& hl, float& lh, float& hh) {
    float x1 = p00 * 0.5f, x2 = p10 * 0.5f, x3 = p01 * 0.5f, x4 = p11 * 0.5f;  // exact in both dtypes
    ll = hv_rnd<T>(hv_rnd<T>(hv_rnd<T>(x1 + x2) + x3) + x4);
    hl = hv_rnd<T>(hv_rnd<T>(hv_rnd<T>(-x1 - x2) + x3) + x4);
    lh = hv_rnd<T>(hv_rnd<T>(hv_rnd<T>(-x1 + x2) - x3) + x4);
    hh = hv_rnd<T>(hv_rnd<T>(hv_rnd<T>(x1 - x2) - x3) + x4);
}

// Vector path: each thread produces VEC output columns of one output row (2 rows x 2*VEC inputs).
template <typename T>
__global__ void __launch_bounds__(256) dwt2_haar_vec_kernel(const T* __restrict__ x, T* __restrict__ ll, T* __restrict__ hl,
                                                            T* __restrict__ lh, T* __restrict__ hh,
                                                            long long planes, int H, int W) {
    constexpr int VEC = Vec16<T>::N;
    const int h2 = H >> 1, w2 = W >> 1, wv = w2 / VEC;
    const long long total = planes * h2 * wv;
    for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        int cv = (int)(idx % wv);
        long long t = idx / wv;
        int i = (int)(t % h2);
        long long pl = t / h2;
        const T* r0 = x + (pl * H + 2 * i) * (long long)W + 2 * cv * VEC;
        const T* r1 = r0 + W;
        float a0[VEC], a1[VEC], b0[VEC], b1[VEC];
        load16<T>(r0, a0); load16<T>(r0 + VEC, a1);
        load16<T>(r1, b0); load16<T>(r1 + VEC, b1);
        float oll[VEC], ohl[VEC], olh[VEC], ohh[VEC];
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            // input column 2j / 2j+1 of the 2*VEC-wide strip
            float p00 = (2 * j < VEC) ? a0[(2 * j) % VEC] : a1[(2 * j) % VEC];
            float p01 = (2 * j + 1 < VEC) ? a0[(2 * j + 1) % VEC] : a1[(2 * j + 1) % VEC];
            float p10 = (2 * j < VEC) ? b0[(2 * j) % VEC] : b1[(2 * j) % VEC];
            float p11 = (2 * j + 1 < VEC) ? b0[(2 * j + 1) % VEC] : b1[(2 * j + 1) % VEC];
            haar_fwd<T>(p00, p10, p01, p11, oll[j], ohl[j], olh[j], ohh[j]);
        }
        long long o = (pl * h2 + i) * (long long)w2 + (long long)cv * VEC;
        store16<T>(ll + o, oll); store16<T>(hl + o, ohl); store16<T>(lh + o, olh); store16<T>(hh + o, ohh);
    }
}

// Scalar path for widths that are not a multiple of 2*VEC (ragged inputs).
template <typename T>
__global__ void __launch_bounds__(256) dwt2_haar_scalar_kernel(const T* __restrict__ x, T* __restrict__ ll, T* __restrict__ hl,
                                                               T* __restrict__ lh, T* __restrict__ hh,
                                                               long long planes, int H, int W) {
    const int h2 = H >> 1, w2 = W >> 1;
    const long long total = planes * h2 * w2;
    for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        int j = (int)(idx % w2);
        long long t = idx / w2;
        int i = (int)(t % h2);
        long long pl = t / h2;
        const T* r0 = x + (pl * H + 2 * i) * (long long)W + 2 * j;
        float a, b, c, d;
        haar_fwd<T>(hv_ld(r0), hv_ld(r0 + W), hv_ld(r0 + 1), hv_ld(r0 + W + 1), a, b, c, d);
        hv_st(ll + idx, a); hv_st(hl + idx, b); hv_st(lh + idx, c); hv_st(hh + idx, d);
    }
}

// ---------------------------------------------------------------------------------------------
// Fused multi-level analysis (the caller-side recursion on LL of frenquence_wavelet.py:86-89): one thread
// owns a 2^L x 8 input patch, keeps every intermediate LL in registers (rounded to the tensor dtype exactly
// where the per-level path would store it) and writes the 3 detail bands of every level plus the last LL.
// x is read once and no intermediate LL ever touches HBM: 1 launch instead of L, (1 + 1) N bytes instead of
// 2 N (1 + 1/4 + ...).
struct PyrOut {
    void* hl[3];
    void* lh[3];
    void* hh[3];
    void* ll;
};

template <typename T, int N>
__device__ __forceinline__ void store_n(T* p, const float* v) {  // N in {1, 2, 4} consecutive elements, N*sizeof(T)-aligned
    if constexpr (sizeof(T) == 4) {
        if constexpr (N == 4) *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
        else if constexpr (N == 2) *reinterpret_cast<float2*>(p) = make_float2(v[0], v[1]);
        else *p = v[0];
    } else {
        if constexpr (N == 4) *reinterpret_cast<uint2*>(p) = make_uint2(hv_pack_bf16(v[0], v[1]), hv_pack_bf16(v[2], v[3]));
        else if constexpr (N == 2) *reinterpret_cast<uint32_t*>(p) = hv_pack_bf16(v[0], v[1]);
        else *p = __float2bfloat16(v[0]);
    }
}

// one level on a register patch: in[R][C] -> ll[R/2][C/2]; bands stored at (row0 + i, col0 .. col0 + C/2) of a plane of width w
template <typename T, int R, int C>
__device__ __forceinline__ void pyr_level(const float (&in)[R][C], float (&ll)[R / 2][C / 2], T* hl, T* lh, T* hh,
                                          long long plane_off, int row0, int col0, int w) {
#pragma unroll
    for (int i = 0; i < R / 2; ++i) {
        float bhl[C / 2], blh[C / 2], bhh[C / 2];
#pragma unroll
        for (int j = 0; j < C / 2; ++j)
            haar_fwd<T>(in[2 * i][2 * j], in[2 * i + 1][2 * j], in[2 * i][2 * j + 1], in[2 * i + 1][2 * j + 1], ll[i][j], bhl[j], blh[j], bhh[j]);
        const long long o = plane_off + (long long)(row0 + i) * w + col0;
        store_n<T, C / 2>(hl + o, bhl); store_n<T, C / 2>(lh + o, blh); store_n<T, C / 2>(hh + o, bhh);
    }
}

template <typename T, int LV>
__global__ void __launch_bounds__(256) dwt2_haar_pyramid_kernel(const T* __restrict__ x, PyrOut out, long long planes, int H, int W) {
    constexpr int B = 1 << LV;
    const int tw = W >> 3, th = H / B;
    const long long total = planes * th * tw;
    for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
        const int cx = (int)(idx % tw);
        const long long t = idx / tw;
        const int by = (int)(t % th);
        const long long pl = t / th;
        float a[B][8];
        const T* src = x + (pl * H + (long long)by * B) * W + cx * 8;
#pragma unroll
        for (int r = 0; r < B; ++r) {
            if constexpr (sizeof(T) == 4) {
                float lo[4], hi[4];
                load16<T>(src + (long long)r * W, lo); load16<T>(src + (long long)r * W + 4, hi);
#pragma unroll
                for (int j = 0; j < 4; ++j) { a[r][j] = lo[j]; a[r][4 + j] = hi[j]; }
            } else {
                load16<T>(src + (long long)r * W, a[r]);
            }
        }
        float l1[B / 2][4];
        pyr_level<T, B, 8>(a, l1, (T*)out.hl[0], (T*)out.lh[0], (T*)out.hh[0], pl * (H / 2) * (W / 2), by * (B / 2), cx * 4, W / 2);
        if constexpr (LV == 1) {
#pragma unroll
            for (int i = 0; i < B / 2; ++i) store_n<T, 4>((T*)out.ll + pl * (H / 2) * (W / 2) + (long long)(by * (B / 2) + i) * (W / 2) + cx * 4, l1[i]);
        } else {
            float l2[B / 4][2];
            pyr_level<T, B / 2, 4>(l1, l2, (T*)out.hl[1], (T*)out.lh[1], (T*)out.hh[1], pl * (H / 4) * (W / 4), by * (B / 4), cx * 2, W / 4);
            if constexpr (LV == 2) {
#pragma unroll
                for (int i = 0; i < B / 4; ++i) store_n<T, 2>((T*)out.ll + pl * (H / 4) * (W / 4) + (long long)(by * (B / 4) + i) * (W / 4) + cx * 2, l2[i]);
            } else {
                float l3[B / 8][1];
                pyr_level<T, B / 4, 2>(l2, l3, (T*)out.hl[2], (T*)out.lh[2], (T*)out.hh[2], pl * (H / 8) * (W / 8), by * (B / 8), cx, W / 8);
                store_n<T, 1>((T*)out.ll + pl * (H / 8) * (W / 8) + (long long)by * (W / 8) + cx, l3[0]);
            }
        }
    }
}

template <typename T>
__device__ __forceinline__ void haar_inv(float a, float b, float c, float d, float& o00, float& o10, float& o01, float& o11) {
    float x1 = a * 0.5f, x2 = b * 0.5f, x3 = c * 0.5f, x4 = d * 0.5f;
    o00 = hv_rnd<T>(hv_rnd<T>(hv_rnd<T>(x1 - x2) - x3) + x4);  // h[0::2,0::2]  wavelet.py:43
    o10 = hv_rnd<T>(hv_rnd<T>(hv_rnd<T>(x1 - x2) + x3) - x4);  // h[1::2,0::2]
    o01 = hv_rnd<T>(hv_rnd<T>(hv_rnd<T>(x1 + x2) - x3) - x4);  // h[0::2,1::2]
    o11 = hv_rnd<T>(hv_rnd<T>(hv_rnd<T>(x1 + x2) + x3) + x4);  // h[1::2,1::2]
}

// VEC consecutive elements of type T -> floats (16-byte load for 4 x fp32 / 8 x bf16, 8-byte load for 4 x bf16)
template <typename T, int VEC>
__device__ __forceinline__ void load_vec(const T* p, float (&v)[VEC]) {
    if constexpr (sizeof(T) * VEC == 16) {
        load16<T>(p, v);
    } else {
        static_assert(sizeof(T) == 2 && VEC == 4, "load_vec: unsupported combination");
        uint2 raw = *reinterpret_cast<const uint2*>(p);
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
        float2 f0 = __bfloat1622float2(h[0]), f1 = __bfloat1622float2(h[1]);
        v[0] = f0.x; v[1] = f0.y; v[2] = f1.x; v[3] = f1.y;
    }
}

// TL = dtype of LL, TB = dtype of the three detail bands.  The reference concatenates the four bands
// (torch.cat promotes a mixed fp32/bf16 list to fp32), so the arithmetic type R is bf16 only when both are.
template <typename TL, typename TB>
__global__ void __launch_bounds__(256) idwt2_haar_vec_kernel(const TL* __restrict__ ll, const TB* __restrict__ hl,
                                                             const TB* __restrict__ lh, const TB* __restrict__ hh,
                                                             float* __restrict__ out, long long planes, int h, int w) {
    using R = typename std::conditional<sizeof(TL) == 2 && sizeof(TB) == 2, __nv_bfloat16, float>::type;
    constexpr int VEC = 16 / (sizeof(TL) > sizeof(TB) ? sizeof(TL) : sizeof(TB));
    const int wv = w / VEC;
    const long long total = planes * h * wv;
    for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        int cv = (int)(idx % wv);
        long long t = idx / wv;
        int i = (int)(t % h);
        long long pl = t / h;
        long long o = (pl * h + i) * (long long)w + (long long)cv * VEC;
        float a[VEC], b[VEC], c[VEC], d[VEC];
        load_vec<TL, VEC>(ll + o, a); load_vec<TB, VEC>(hl + o, b); load_vec<TB, VEC>(lh + o, c); load_vec<TB, VEC>(hh + o, d);
        float r0[2 * VEC], r1[2 * VEC];
#pragma unroll
        for (int j = 0; j < VEC; ++j) haar_inv<R>(a[j], b[j], c[j], d[j], r0[2 * j], r1[2 * j], r0[2 * j + 1], r1[2 * j + 1]);
        float* p0 = out + (pl * 2 * h + 2 * i) * (long long)(2 * w) + (long long)cv * 2 * VEC;
        float* p1 = p0 + 2 * w;
#pragma unroll
        for (int q = 0; q < 2 * VEC / 4; ++q) {
            *reinterpret_cast<float4*>(p0 + 4 * q) = make_float4(r0[4 * q], r0[4 * q + 1], r0[4 * q + 2], r0[4 * q + 3]);
            *reinterpret_cast<float4*>(p1 + 4 * q) = make_float4(r1[4 * q], r1[4 * q + 1], r1[4 * q + 2], r1[4 * q + 3]);
        }
    }
}

template <typename TL, typename TB>
__global__ void __launch_bounds__(256) idwt2_haar_scalar_kernel(const TL* __restrict__ ll, const TB* __restrict__ hl,
                                                                const TB* __restrict__ lh, const TB* __restrict__ hh,
                                                                float* __restrict__ out, long long planes, int h, int w) {
    using R = typename std::conditional<sizeof(TL) == 2 && sizeof(TB) == 2, __nv_bfloat16, float>::type;
    const long long total = planes * h * w;
    for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        int j = (int)(idx % w);
        long long t = idx / w;
        int i = (int)(t % h);
        long long pl = t / h;
        float o00, o10, o01, o11;
        haar_inv<R>(hv_ld(ll + idx), hv_ld(hl + idx), hv_ld(lh + idx), hv_ld(hh + idx), o00, o10, o01, o11);
        float* p0 = out + (pl * 2 * h + 2 * i) * (long long)(2 * w) + 2 * j;
        p0[0] = o00; p0[1] = o01; p0[2 * w] = o10; p0[2 * w + 1] = o11;
    }
}

// ---------------------------------------------------------------------------------------------
// Fused multi-level synthesis: the mirror of dwt2_haar_pyramid_kernel.  One thread rebuilds a 2^L x 8 output
// patch from 1 x (8 >> L) coarse LL values and the matching detail-band patches of every level; no
// intermediate image touches HBM.  Rounding follows the per-level path: the coarsest level computes in the
// band dtype (four same-dtype bands), every later level in fp32 (iwt_init returned fp32, torch.cat promoted).
template <typename T, int N>
__device__ __forceinline__ void load_n(const T* p, float* v) {  // N in {1, 2, 4}
    if constexpr (sizeof(T) == 4) {
        if constexpr (N == 4) { float4 t = *reinterpret_cast<const float4*>(p); v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w; }
        else if constexpr (N == 2) { float2 t = *reinterpret_cast<const float2*>(p); v[0] = t.x; v[1] = t.y; }
        else v[0] = *p;
    } else {
        if constexpr (N == 4) {
            uint2 raw = *reinterpret_cast<const uint2*>(p);
            const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
            float2 f0 = __bfloat1622float2(h[0]), f1 = __bfloat1622float2(h[1]);
            v[0] = f0.x; v[1] = f0.y; v[2] = f1.x; v[3] = f1.y;
        } else if constexpr (N == 2) {
            float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(p));
            v[0] = f.x; v[1] = f.y;
        } else v[0] = __bfloat162float(*p);
    }
}

// cur[R][C] (LL) + bands at (row0 + i, col0 .. col0 + C) of planes of width w  ->  next[2R][2C]
template <typename TB, typename RT, int R, int C>
__device__ __forceinline__ void pyr_inv_level(const float (&cur)[R][C], float (&next)[2 * R][2 * C], const TB* hl, const TB* lh,
                                              const TB* hh, long long plane_off, int row0, int col0, int w) {
#pragma unroll
    for (int i = 0; i < R; ++i) {
        float b[C], c[C], d[C];
        const long long o = plane_off + (long long)(row0 + i) * w + col0;
        load_n<TB, C>(hl + o, b); load_n<TB, C>(lh + o, c); load_n<TB, C>(hh + o, d);
#pragma unroll
        for (int j = 0; j < C; ++j)
            haar_inv<RT>(cur[i][j], b[j], c[j], d[j], next[2 * i][2 * j], next[2 * i + 1][2 * j], next[2 * i][2 * j + 1], next[2 * i + 1][2 * j + 1]);
    }
}

template <typename T, int LV>
__global__ void __launch_bounds__(256) idwt2_haar_pyramid_kernel(PyrOut in, float* __restrict__ out, long long planes, int H, int W) {
    constexpr int B = 1 << LV;  // H, W: full-resolution output size
    const int tw = W >> 3, th = H / B;
    const long long total = planes * th * tw;
    for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
        const int cx = (int)(idx % tw);
        const long long t = idx / tw;
        const int by = (int)(t % th);
        const long long pl = t / th;
        float full[B][8];
        constexpr int CL = 8 >> LV;  // coarse LL values per thread
        float top[1][CL];
        load_n<T, CL>((const T*)in.ll + pl * (H >> LV) * (W >> LV) + (long long)by * (W >> LV) + cx * CL, top[0]);
        if constexpr (LV == 1) {
            pyr_inv_level<T, T, 1, 4>(top, full, (const T*)in.hl[0], (const T*)in.lh[0], (const T*)in.hh[0], pl * (H / 2) * (W / 2), by, cx * 4, W / 2);
        } else if constexpr (LV == 2) {
            float m1[2][4];
            pyr_inv_level<T, T, 1, 2>(top, m1, (const T*)in.hl[1], (const T*)in.lh[1], (const T*)in.hh[1], pl * (H / 4) * (W / 4), by, cx * 2, W / 4);
            pyr_inv_level<T, float, 2, 4>(m1, full, (const T*)in.hl[0], (const T*)in.lh[0], (const T*)in.hh[0], pl * (H / 2) * (W / 2), by * 2, cx * 4, W / 2);
        } else {
            float m2[2][2], m1[4][4];
            pyr_inv_level<T, T, 1, 1>(top, m2, (const T*)in.hl[2], (const T*)in.lh[2], (const T*)in.hh[2], pl * (H / 8) * (W / 8), by, cx, W / 8);
            pyr_inv_level<T, float, 2, 2>(m2, m1, (const T*)in.hl[1], (const T*)in.lh[1], (const T*)in.hh[1], pl * (H / 4) * (W / 4), by * 2, cx * 2, W / 4);
            pyr_inv_level<T, float, 4, 4>(m1, full, (const T*)in.hl[0], (const T*)in.lh[0], (const T*)in.hh[0], pl * (H / 2) * (W / 2), by * 4, cx * 4, W / 2);
        }
        float* dst = out + (pl * H + (long long)by * B) * W + cx * 8;
#pragma unroll
        for (int r = 0; r < B; ++r) {
            *reinterpret_cast<float4*>(dst + (long long)r * W) = make_float4(full[r][0], full[r][1], full[r][2], full[r][3]);
            *reinterpret_cast<float4*>(dst + (long long)r * W + 4) = make_float4(full[r][4], full[r][5], full[r][6], full[r][7]);
        }
    }
}

static int grid_for(long long total, int block) {
    long long need = (total + block - 1) / block;
    long long cap = (long long)hv_num_sms() * 8;  // 8 resident 256-thread CTAs per SM
    if (need < 1) need = 1;
    return (int)(need < cap ? need : cap);
}

template <typename T>
static int dwt_launch(const void* x, void* ll, void* hl, void* lh, void* hh, int64_t planes, int H, int W, cudaStream_t st) {
    constexpr int VEC = Vec16<T>::N;
    const bool vec_ok = (W % (2 * VEC) == 0) && ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(ll) |
                        reinterpret_cast<uintptr_t>(hl) | reinterpret_cast<uintptr_t>(lh) | reinterpret_cast<uintptr_t>(hh)) % 16 == 0);
    if (vec_ok) {
        long long total = planes * (H / 2) * (W / 2 / VEC);
        dwt2_haar_vec_kernel<T><<<grid_for(total, 256), 256, 0, st>>>((const T*)x, (T*)ll, (T*)hl, (T*)lh, (T*)hh, planes, H, W);
    } else {
        long long total = planes * (H / 2) * (W / 2);
        dwt2_haar_scalar_kernel<T><<<grid_for(total, 256), 256, 0, st>>>((const T*)x, (T*)ll, (T*)hl, (T*)lh, (T*)hh, planes, H, W);
    }
    HV_LAUNCH_CHECK();
    return HV_OK;
}

extern "C" int hv_dwt2_haar(const void* x, void* ll, void* hl, void* lh, void* hh,
                            int64_t planes, int H, int W, int dtype, void* stream) {
    HV_REQUIRE(planes >= 0 && H >= 0 && W >= 0, "hv_dwt2_haar: negative size");
    HV_REQUIRE(H % 2 == 0 && W % 2 == 0, "hv_dwt2_haar: H and W must be even (got %d x %d)", H, W);
    if (planes == 0 || H == 0 || W == 0) return HV_OK;
    if (dtype == HV_F32) return dwt_launch<float>(x, ll, hl, lh, hh, planes, H, W, hv_stream(stream));
    if (dtype == HV_BF16) return dwt_launch<__nv_bfloat16>(x, ll, hl, lh, hh, planes, H, W, hv_stream(stream));
    HV_REQUIRE(false, "hv_dwt2_haar: bad dtype %d", dtype);
}

template <typename T>
static int pyramid_launch(const void* x, const PyrOut& o, int levels, int64_t planes, int H, int W, cudaStream_t st) {
    const long long total = planes * (H >> levels) * (W >> 3);
    const int grid = grid_for(total, 256);
    if (levels == 1) dwt2_haar_pyramid_kernel<T, 1><<<grid, 256, 0, st>>>((const T*)x, o, planes, H, W);
    else if (levels == 2) dwt2_haar_pyramid_kernel<T, 2><<<grid, 256, 0, st>>>((const T*)x, o, planes, H, W);
    else dwt2_haar_pyramid_kernel<T, 3><<<grid, 256, 0, st>>>((const T*)x, o, planes, H, W);
    HV_LAUNCH_CHECK();
    return HV_OK;
}

// levels in 1..3; hl/lh/hh: arrays of `levels` device pointers (level k: planes x H/2^(k+1) x W/2^(k+1)); ll: last level.
// Needs H % 2^levels == 0, W % 8 == 0 and 16-byte aligned x (the caller falls back to per-level hv_dwt2_haar otherwise).
extern "C" int hv_dwt2_haar_pyramid(const void* x, int levels, void* const* hl, void* const* lh, void* const* hh, void* ll,
                                    int64_t planes, int H, int W, int dtype, void* stream) {
    HV_REQUIRE(levels >= 1 && levels <= 3, "hv_dwt2_haar_pyramid: levels must be 1..3 (got %d)", levels);
    HV_REQUIRE(planes >= 0 && H >= 0 && W >= 0, "hv_dwt2_haar_pyramid: negative size");
    HV_REQUIRE(H % (1 << levels) == 0 && W % 8 == 0, "hv_dwt2_haar_pyramid: needs H %% 2^levels == 0 and W %% 8 == 0 (got %d x %d)", H, W);
    HV_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15) == 0, "hv_dwt2_haar_pyramid: x must be 16-byte aligned");
    HV_REQUIRE(dtype == HV_F32 || dtype == HV_BF16, "hv_dwt2_haar_pyramid: bad dtype %d", dtype);
    if (planes == 0 || H == 0 || W == 0) return HV_OK;
    PyrOut o;
    for (int k = 0; k < 3; ++k) { o.hl[k] = k < levels ? hl[k] : nullptr; o.lh[k] = k < levels ? lh[k] : nullptr; o.hh[k] = k < levels ? hh[k] : nullptr; }
    o.ll = ll;
    const int es = dtype == HV_F32 ? 4 : 2;
    for (int k = 0; k < levels; ++k) {  // store_n alignment: N = 4 >> k elements per store
        const uintptr_t al = (uintptr_t)((4 >> k) * es - 1);
        HV_REQUIRE(((reinterpret_cast<uintptr_t>(o.hl[k]) | reinterpret_cast<uintptr_t>(o.lh[k]) | reinterpret_cast<uintptr_t>(o.hh[k])) & al) == 0,
                   "hv_dwt2_haar_pyramid: level %d bands are misaligned", k);
    }
    if (dtype == HV_F32) return pyramid_launch<float>(x, o, levels, planes, H, W, hv_stream(stream));
    return pyramid_launch<__nv_bfloat16>(x, o, levels, planes, H, W, hv_stream(stream));
}

// inverse of hv_dwt2_haar_pyramid: H x W is the full-resolution OUTPUT size; out is fp32 (wavelet.py:40-41)
extern "C" int hv_idwt2_haar_pyramid(const void* ll, int levels, const void* const* hl, const void* const* lh, const void* const* hh,
                                     float* out, int64_t planes, int H, int W, int dtype, void* stream) {
    HV_REQUIRE(levels >= 1 && levels <= 3, "hv_idwt2_haar_pyramid: levels must be 1..3 (got %d)", levels);
    HV_REQUIRE(planes >= 0 && H >= 0 && W >= 0, "hv_idwt2_haar_pyramid: negative size");
    HV_REQUIRE(H % (1 << levels) == 0 && W % 8 == 0, "hv_idwt2_haar_pyramid: needs H %% 2^levels == 0 and W %% 8 == 0 (got %d x %d)", H, W);
    HV_REQUIRE((reinterpret_cast<uintptr_t>(out) & 15) == 0, "hv_idwt2_haar_pyramid: out must be 16-byte aligned");
    HV_REQUIRE(dtype == HV_F32 || dtype == HV_BF16, "hv_idwt2_haar_pyramid: bad dtype %d", dtype);
    if (planes == 0 || H == 0 || W == 0) return HV_OK;
    PyrOut o;
    for (int k = 0; k < 3; ++k) {
        o.hl[k] = k < levels ? const_cast<void*>(hl[k]) : nullptr; o.lh[k] = k < levels ? const_cast<void*>(lh[k]) : nullptr;
        o.hh[k] = k < levels ? const_cast<void*>(hh[k]) : nullptr;
    }
    o.ll = const_cast<void*>(ll);
    const int es = dtype == HV_F32 ? 4 : 2;
    for (int k = 0; k < levels; ++k) {
        const uintptr_t al = (uintptr_t)((4 >> k) * es - 1);
        HV_REQUIRE(((reinterpret_cast<uintptr_t>(o.hl[k]) | reinterpret_cast<uintptr_t>(o.lh[k]) | reinterpret_cast<uintptr_t>(o.hh[k])) & al) == 0,
                   "hv_idwt2_haar_pyramid: level %d bands are misaligned", k);
    }
    HV_REQUIRE((reinterpret_cast<uintptr_t>(ll) & (uintptr_t)((8 >> levels) * es - 1)) == 0, "hv_idwt2_haar_pyramid: ll is misaligned");
    const long long total = planes * (H >> levels) * (W >> 3);
    const int grid = grid_for(total, 256);
    cudaStream_t st = hv_stream(stream);
#define IPYR(T, L) idwt2_haar_pyramid_kernel<T, L><<<grid, 256, 0, st>>>(o, out, planes, H, W)
    if (dtype == HV_F32) { if (levels == 1) IPYR(float, 1); else if (levels == 2) IPYR(float, 2); else IPYR(float, 3); }
    else { if (levels == 1) IPYR(__nv_bfloat16, 1); else if (levels == 2) IPYR(__nv_bfloat16, 2); else IPYR(__nv_bfloat16, 3); }
#undef IPYR
    HV_LAUNCH_CHECK();
    return HV_OK;
}

template <typename TL, typename TB>
static int idwt_launch(const void* ll, const void* hl, const void* lh, const void* hh, float* out,
                       int64_t planes, int h, int w, cudaStream_t st) {
    constexpr int VEC = 16 / (sizeof(TL) > sizeof(TB) ? sizeof(TL) : sizeof(TB));
    const bool vec_ok = (w % VEC == 0) && ((reinterpret_cast<uintptr_t>(out) | reinterpret_cast<uintptr_t>(ll) |
                        reinterpret_cast<uintptr_t>(hl) | reinterpret_cast<uintptr_t>(lh) | reinterpret_cast<uintptr_t>(hh)) % 16 == 0);
    if (vec_ok) {
        long long total = planes * h * (w / VEC);
        idwt2_haar_vec_kernel<TL, TB><<<grid_for(total, 256), 256, 0, st>>>((const TL*)ll, (const TB*)hl, (const TB*)lh, (const TB*)hh, out, planes, h, w);
    } else {
        long long total = planes * h * w;
        idwt2_haar_scalar_kernel<TL, TB><<<grid_for(total, 256), 256, 0, st>>>((const TL*)ll, (const TB*)hl, (const TB*)lh, (const TB*)hh, out, planes, h, w);
    }
    HV_LAUNCH_CHECK();
    return HV_OK;
}

// dtype_in = dtype of all four bands, or (dtype_ll | dtype_bands << 8) for the mixed case that a bf16 pyramid
// produces from its second inverse level on (LL is the fp32 output of the previous level, wavelet.py:40-41).
extern "C" int hv_idwt2_haar(const void* ll, const void* hl, const void* lh, const void* hh, float* out,
                             int64_t planes, int h, int w, int dtype_in, void* stream) {
    HV_REQUIRE(planes >= 0 && h >= 0 && w >= 0, "hv_idwt2_haar: negative size");
    if (planes == 0 || h == 0 || w == 0) return HV_OK;
    const int dl = dtype_in & 0xff, db = (dtype_in >> 8) ? ((dtype_in >> 8) & 0xff) - 1 : dl;
    if (dl == HV_F32 && db == HV_F32) return idwt_launch<float, float>(ll, hl, lh, hh, out, planes, h, w, hv_stream(stream));
    if (dl == HV_BF16 && db == HV_BF16) return idwt_launch<__nv_bfloat16, __nv_bfloat16>(ll, hl, lh, hh, out, planes, h, w, hv_stream(stream));
    if (dl == HV_F32 && db == HV_BF16) return idwt_launch<float, __nv_bfloat16>(ll, hl, lh, hh, out, planes, h, w, hv_stream(stream));
    if (dl == HV_BF16 && db == HV_F32) return idwt_launch<__nv_bfloat16, float>(ll, hl, lh, hh, out, planes, h, w, hv_stream(stream));
    HV_REQUIRE(false, "hv_idwt2_haar: bad dtype code %d", dtype_in);
}

// ---------------------------------------------------------------------------------------------
// 3-D Gaussian low/high split.  One CTA owns one T x H x W volume in shared memory (row pitch
// W+1 floats -> conflict-free line access along every axis) and applies the three circulant
// matrices in place, one axis at a time; each thread owns whole lines, held in registers.
template <int MAXN>
__device__ __forceinline__ void circ_line(float* line, int stride, int N, const float* __restrict__ C) {
    float xr[MAXN];
#pragma unroll
    for (int k = 0; k < MAXN; ++k) xr[k] = (k < N) ? line[k * stride] : 0.0f;
    for (int i = 0; i < N; ++i) {
        const float4* row = reinterpret_cast<const float4*>(C + i * MAXN);
        float acc = 0.0f;
#pragma unroll
        for (int k4 = 0; k4 < MAXN / 4; ++k4) {
            float4 c = row[k4];  // broadcast read, zero-padded beyond N
            acc = fmaf(c.x, xr[4 * k4], acc); acc = fmaf(c.y, xr[4 * k4 + 1], acc);
            acc = fmaf(c.z, xr[4 * k4 + 2], acc); acc = fmaf(c.w, xr[4 * k4 + 3], acc);
        }
        line[i * stride] = acc;
    }
}

template <int MAXN>
__global__ void __launch_bounds__(256) freq_split3d_kernel(const float* __restrict__ x, const float* __restrict__ ct,
                                                           const float* __restrict__ ch, const float* __restrict__ cw,
                                                           float* __restrict__ lo, float* __restrict__ hi,
                                                           int T, int H, int W) {
    extern __shared__ __align__(16) float smem[];
    float* sct = smem;                 // T rows x MAXN
    float* sch = sct + MAXN * MAXN;
    float* scw = sch + MAXN * MAXN;
    float* vol = scw + MAXN * MAXN;    // T*H rows, pitch W+1
    const int P = W + 1;
    const long long nvol = (long long)T * H * W;
    const float* xv = x + blockIdx.x * nvol;
    for (int i = threadIdx.x; i < 3 * MAXN * MAXN; i += blockDim.x) smem[i] = 0.0f;
    __syncthreads();
    for (int i = threadIdx.x; i < T * T; i += blockDim.x) sct[(i / T) * MAXN + i % T] = ct[i];
    for (int i = threadIdx.x; i < H * H; i += blockDim.x) sch[(i / H) * MAXN + i % H] = ch[i];
    for (int i = threadIdx.x; i < W * W; i += blockDim.x) scw[(i / W) * MAXN + i % W] = cw[i];
    for (long long i = threadIdx.x; i < nvol; i += blockDim.x) vol[(i / W) * P + i % W] = xv[i];
    __syncthreads();
    // W axis: line = (t,h), element stride 1, consecutive threads -> pitch P apart (conflict-free)
    for (int l = threadIdx.x; l < T * H; l += blockDim.x) circ_line<MAXN>(vol + l * P, 1, W, scw);
    __syncthreads();
    // H axis: line = (t,w), stride P
    for (int l = threadIdx.x; l < T * W; l += blockDim.x) circ_line<MAXN>(vol + (l / W) * H * P + l % W, P, H, sch);
    __syncthreads();
    // T axis: line = (h,w), stride H*P
    for (int l = threadIdx.x; l < H * W; l += blockDim.x) circ_line<MAXN>(vol + (l / W) * P + l % W, H * P, T, sct);
    __syncthreads();
    float* lov = lo + blockIdx.x * nvol;
    float* hiv = hi + blockIdx.x * nvol;
    for (long long i = threadIdx.x; i < nvol; i += blockDim.x) {
        float l = vol[(i / W) * P + i % W];
        lov[i] = l;
        hiv[i] = xv[i] - l;
    }
}

// Slab variant for axis lengths <= 32 (every shape on the path): a volume is split into T/TC slabs of TC
// consecutive frames, one CTA per slab, so that 32 volumes already give 128 CTAs and several CTAs fit on an SM
// (46 KB instead of 152 KB of shared memory).  The T axis goes FIRST, straight from global memory: thread (h, w)
// reads its T-long column (coalesced across w), keeps it in registers and produces only the slab's TC output
// frames — a CTA reads the whole volume (L2 hits after the first slab) but never needs another CTA's data.
// The H and W passes then run inside the slab in shared memory as above.  Four accumulators per output break
// the 32-long dependent FMA chain.
__device__ __forceinline__ float circ_dot32(const float* __restrict__ crow, const float (&xr)[32]) {
    const float4* row = reinterpret_cast<const float4*>(crow);
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
    for (int k4 = 0; k4 < 8; ++k4) {
        const float4 c = row[k4];  // broadcast read, zero-padded beyond the axis length
        a0 = fmaf(c.x, xr[4 * k4], a0); a1 = fmaf(c.y, xr[4 * k4 + 1], a1);
        a2 = fmaf(c.z, xr[4 * k4 + 2], a2); a3 = fmaf(c.w, xr[4 * k4 + 3], a3);
    }
    return (a0 + a1) + (a2 + a3);
}

__device__ __forceinline__ void circ_line32(float* line, int stride, int N, const float* __restrict__ C) {
    float xr[32];
#pragma unroll
    for (int k = 0; k < 32; ++k) xr[k] = (k < N) ? line[k * stride] : 0.0f;
    for (int i = 0; i < N; ++i) line[i * stride] = circ_dot32(C + i * 32, xr);
}

__global__ void __launch_bounds__(256, 3) freq_split3d_slab_kernel(const float* __restrict__ x, const float* __restrict__ ct,
                                                                const float* __restrict__ ch, const float* __restrict__ cw,
                                                                float* __restrict__ lo, float* __restrict__ hi,
                                                                int T, int H, int W, int TC) {
    extern __shared__ __align__(16) float smem[];
    float* sct = smem;                 // TC rows x 32 (rows t0 .. t0+TC-1 of the T circulant)
    float* sch = sct + 32 * 32;
    float* scw = sch + 32 * 32;
    float* slab = scw + 32 * 32;       // TC*H rows, pitch W+1
    const int P = W + 1, HW = H * W;
    const int slabs = T / TC;
    const long long v = blockIdx.x / slabs;
    const int t0 = (int)(blockIdx.x - v * slabs) * TC;
    const float* xv = x + v * (long long)T * HW;
    for (int i = threadIdx.x; i < 3 * 32 * 32; i += blockDim.x) smem[i] = 0.0f;
    __syncthreads();
    for (int i = threadIdx.x; i < TC * T; i += blockDim.x) sct[(i / T) * 32 + i % T] = ct[(t0 + i / T) * T + i % T];
    for (int i = threadIdx.x; i < H * H; i += blockDim.x) sch[(i / H) * 32 + i % H] = ch[i];
    for (int i = threadIdx.x; i < W * W; i += blockDim.x) scw[(i / W) * 32 + i % W] = cw[i];
    __syncthreads();
    // T axis from global memory: column (h, w), element stride H*W
    for (int col = threadIdx.x; col < HW; col += blockDim.x) {
        float xr[32];
#pragma unroll
        for (int k = 0; k < 32; ++k) xr[k] = (k < T) ? __ldg(xv + (long long)k * HW + col) : 0.0f;
        float* dst = slab + (col / W) * P + col % W;
        for (int tt = 0; tt < TC; ++tt) dst[tt * H * P] = circ_dot32(sct + tt * 32, xr);
    }
    __syncthreads();
    // H axis: line = (tt, w), stride P
    for (int l = threadIdx.x; l < TC * W; l += blockDim.x) circ_line32(slab + (l / W) * H * P + l % W, P, H, sch);
    __syncthreads();
    // W axis: line = (tt, h), stride 1, consecutive threads a pitch apart (conflict-free)
    for (int l = threadIdx.x; l < TC * H; l += blockDim.x) circ_line32(slab + l * P, 1, W, scw);
    __syncthreads();
    const long long off = v * (long long)T * HW + (long long)t0 * HW;
    for (int i = threadIdx.x; i < TC * HW; i += blockDim.x) {
        const float l = slab[(i / W) * P + i % W];
        lo[off + i] = l;
        hi[off + i] = __ldg(x + off + i) - l;
    }
}

// 32 x 32 x 32 volumes (the 2t = 32 frame clip of the default model): the three matrices are circulant
// (C[i][k] = c[(k - i) mod 32], a circular convolution), so a thread keeps the 32 taps AND its 32-long line in
// registers and every output is 32 FMAs with compile-time register indices — no shared-memory coefficient reads
// at all (the generic kernel issues one LDS.128 per four FMAs, which is what bounds it).
__device__ __forceinline__ void circ32_line(float* line, int stride, const float (&c)[32]) {
    float xr[32];
#pragma unroll
    for (int k = 0; k < 32; ++k) xr[k] = line[k * stride];
#pragma unroll
    for (int i = 0; i < 32; ++i) {
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
        for (int k = 0; k < 32; k += 4) {
            a0 = fmaf(c[(k - i) & 31], xr[k], a0); a1 = fmaf(c[(k + 1 - i) & 31], xr[k + 1], a1);
            a2 = fmaf(c[(k + 2 - i) & 31], xr[k + 2], a2); a3 = fmaf(c[(k + 3 - i) & 31], xr[k + 3], a3);
        }
        line[i * stride] = (a0 + a1) + (a2 + a3);
    }
}

__global__ void __launch_bounds__(256, 2) freq_split3d_slab32_kernel(const float* __restrict__ x, const float* __restrict__ ct,
                                                                     const float* __restrict__ ch, const float* __restrict__ cw,
                                                                     float* __restrict__ lo, float* __restrict__ hi) {
    constexpr int N = 32, TC = 8, P = N + 1, HW = N * N;
    __shared__ float slab[TC * N * P];
    const long long v = blockIdx.x >> 2;
    const int t0 = (int)(blockIdx.x & 3) * TC;
    const float* xv = x + v * (long long)N * HW;
    float c[32];
    // T axis from global memory; the column is loaded rotated by t0 so that the tap indices stay compile-time constants
#pragma unroll
    for (int k = 0; k < 32; ++k) c[k] = __ldg(ct + k);
    for (int col = threadIdx.x; col < HW; col += 256) {
        float xr[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) xr[j] = __ldg(xv + (long long)((j + t0) & 31) * HW + col);
        float* dst = slab + (col >> 5) * P + (col & 31);
#pragma unroll
        for (int tt = 0; tt < TC; ++tt) {
            float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
                a0 = fmaf(c[(j - tt) & 31], xr[j], a0); a1 = fmaf(c[(j + 1 - tt) & 31], xr[j + 1], a1);
                a2 = fmaf(c[(j + 2 - tt) & 31], xr[j + 2], a2); a3 = fmaf(c[(j + 3 - tt) & 31], xr[j + 3], a3);
            }
            dst[tt * N * P] = (a0 + a1) + (a2 + a3);
        }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 32; ++k) c[k] = __ldg(ch + k);
    circ32_line(slab + (threadIdx.x >> 5) * N * P + (threadIdx.x & 31), P, c);   // H axis: line (tt, w), 256 lines
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 32; ++k) c[k] = __ldg(cw + k);
    circ32_line(slab + threadIdx.x * P, 1, c);                                    // W axis: line (tt, h)
    __syncthreads();
    const long long off = v * (long long)N * HW + (long long)t0 * HW;
    for (int i = threadIdx.x; i < TC * HW; i += 256) {
        const float l = slab[(i >> 5) * P + (i & 31)];
        lo[off + i] = l;
        hi[off + i] = __ldg(x + off + i) - l;
    }
}

// Tensor-core variant for the shapes of the path (T in {8, 16, 32}, H = W in {16, 32}).  The SIMT kernels above are bound by
// the FMA pipe: 96 FMAs per element against 12 B of HBM traffic is 55 % of the HBM roofline at best, 19 % measured.  Each of
// the three circulant passes is a (lines x n) . (n x n) matrix product, so it goes to the tensor cores as m16n8k8 TF32
// `mma.sync` tiles with the 3xTF32 split (x = hi + lo, both TF32; hi.lo + lo.hi + hi.hi accumulated in fp32: ~1e-6 relative,
// the golden test's bound is 1e-5).  This is the legacy tensor path on purpose: the operands are 32 x 32 fp32 matrices that
// change layout between the passes, the op stays HBM / L2 bound, and tcgen05 would need split operands staged in swizzled
// shared memory three times over.
//   CTA = (volume, slab of TC = 16 output frames), 8 warps, the slab lives in shared memory between the passes:
//   T pass: out[tt][col] = sum_k Ct[t0 + tt][k] x[k][col]      A = Ct rows (registers), B = x straight from global / L2
//   H pass: out[h'][w]   = sum_h Ch[h'][h] s[tt][h][w]         A = Ch (registers), B = slab columns; in place, a warp owns
//                                                               whole 8-column blocks
//   W pass: out[h][w']   = sum_w s[tt][h][w] Cw[w'][w]         A = slab rows (float2 loads, k-permuted), B = Cw (registers);
//                                                               in place, a warp owns whole 16-row blocks
//   last:   low = slab, high = x - low                          128-bit coalesced loads and stores
// `slab_idx` (XOR swizzle for N = 32, padded rows for N = 16) keeps every fragment load / store bank-conflict free.
// Measured (profiles/r02_x_ncu_freq_split3d.md, r02_y_bench_freq.txt): 2511 GB/s of algorithmic traffic at 512 volumes of
// 32^3 = 38 % of HBM (SIMT kernels 1742), legacy tensor pipe 47 % active; m16n8k8 TF32 issues at 0.5 / clk / SM, so the
// formulation tops out at 86 % of the HBM roofline.  The T pass (waiting for x) is the largest phase; tried and kept, in
// order: XOR-swizzled slab, two-instruction split, n-permuted 128-bit loads, warp-owned blocks instead of pair barriers,
// cp.async staging, requests hoisted above the CTA barriers.  tools/emulate_freq_mma.py replays the index arithmetic on the host.
__device__ __forceinline__ void split_tf32(float v, uint32_t& hi, uint32_t& lo) {
    // hi = v rounded to TF32 (nearest, ties away: add half an ulp of the 10-bit mantissa, clear the 13 low bits), lo = v - hi
    // exactly; the tensor core ignores the 13 low mantissa bits of lo (|lo| <= 2^-11 |v|, so that costs 2^-21 |v|, either sign).
    // `cvt.rna.tf32.f32` would do the same with inf / nan handling, at five instructions instead of two per conversion.
    hi = (__float_as_uint(v) + 0x1000u) & 0xffffe000u;
    lo = __float_as_uint(v - __uint_as_float(hi));
}
__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// d += a . b with both operands split: the two small cross terms first, the big term last
__device__ __forceinline__ void mma_3xtf32(float (&d)[4], const uint32_t (&ah)[4], const uint32_t (&al)[4],
                                           uint32_t bh0, uint32_t bh1, uint32_t bl0, uint32_t bl1) {
    mma_tf32(d, al, bh0, bh1);
    mma_tf32(d, ah, bl0, bl1);
    mma_tf32(d, ah, bh0, bh1);
}

__device__ __forceinline__ void cp_async16(float* smem_dst, const float* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
}

// slab addressing: element (h, w) of a frame.  N = 32: rows of 32 floats with the 8-float column blocks XOR-swizzled by h & 3
// (no row padding: a 16-frame slab is 66 KB, three CTAs per SM); N = 16: row pitch 24.  Either way the fragment accesses of the
// three passes ((k = t rows) x (n = g columns) loads, (m = g rows) x float2 stores / loads) hit 32 distinct banks.
template <int N>
__device__ __forceinline__ int slab_idx(int h, int w) {
    return N == 32 ? h * 32 + (w ^ ((h & 3) << 3)) : h * 24 + w;
}

template <int T, int N>
__global__ void __launch_bounds__(256, 2) freq_split3d_mma_kernel(const float* __restrict__ x, const float* __restrict__ ct,
                                                                  const float* __restrict__ ch, const float* __restrict__ cw,
                                                                  float* __restrict__ lo, float* __restrict__ hi) {
    constexpr int TC = T >= 16 ? 16 : 8, SLABS = T / TC, HW = N * N, FP = (N == 32 ? N * 32 : N * 24) + 4;
    constexpr int KT = T / 8, KN = N / 8, MT = N / 16;
    extern __shared__ __align__(16) float S[];     // TC frames, frame pitch FP
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const long long v = blockIdx.x / SLABS;
    const int t0 = (int)(blockIdx.x % SLABS) * TC;
    const float* xv = x + v * (long long)T * HW;
    // ---- T pass: A = rows t0 .. t0 + 15 of Ct (rows 8 .. 15 are zero when the slab has 8 frames).  A warp takes 32 columns
    // at a time as four n-tiles whose n index is permuted (tile nt, index g <-> column 4 g + nt): one 128-bit load per k row
    // feeds all four tiles, and a lane's outputs are eight consecutive columns (two 128-bit stores per frame).
    {
        // x reaches the warp through its own staging buffer (T rows of 32 columns, row pitch 40 floats: conflict-free 128-bit
        // fragment reads): `cp.async` of the next 32 columns is in flight while these 32 are multiplied
        float* stg = S + TC * FP + warp * (T * 40);
        auto stage = [&](int j) {
#pragma unroll
            for (int i = 0; i < T / 4; ++i) {
                const int c = lane + 32 * i, row = c >> 3, ch = c & 7;
                cp_async16(stg + row * 40 + 4 * ch, xv + (long long)row * HW + 32 * j + 4 * ch);
            }
        };
        stage(warp);
        uint32_t ah[KT][4], al[KT][4];
#pragma unroll
        for (int ks = 0; ks < KT; ++ks) {
            const float* r0 = ct + (t0 + g) * T + 8 * ks + t;
            const float* r1 = ct + (t0 + (TC == 16 ? g + 8 : g)) * T + 8 * ks + t;
            split_tf32(__ldg(r0), ah[ks][0], al[ks][0]);
            split_tf32(__ldg(r0 + 4), ah[ks][2], al[ks][2]);
            split_tf32(TC == 16 ? __ldg(r1) : 0.f, ah[ks][1], al[ks][1]);
            split_tf32(TC == 16 ? __ldg(r1 + 4) : 0.f, ah[ks][3], al[ks][3]);
        }
        for (int j = warp; j < HW / 32; j += 8) {
            cp_async_wait_all();
            __syncwarp();
            float4 b[KT][2];
#pragma unroll
            for (int ks = 0; ks < KT; ++ks) {
                b[ks][0] = *reinterpret_cast<const float4*>(stg + (8 * ks + t) * 40 + 4 * g);
                b[ks][1] = *reinterpret_cast<const float4*>(stg + (8 * ks + t + 4) * 40 + 4 * g);
            }
            __syncwarp();
            if (j + 8 < HW / 32) stage(j + 8);
            float acc[4][4];
#pragma unroll
            for (int nt = 0; nt < 4; ++nt) acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.f;
#pragma unroll
            for (int ks = 0; ks < KT; ++ks) {
                const float b0[4] = {b[ks][0].x, b[ks][0].y, b[ks][0].z, b[ks][0].w};
                const float b1[4] = {b[ks][1].x, b[ks][1].y, b[ks][1].z, b[ks][1].w};
#pragma unroll
                for (int nt = 0; nt < 4; ++nt) {
                    uint32_t bh0, bl0, bh1, bl1;
                    split_tf32(b0[nt], bh0, bl0);
                    split_tf32(b1[nt], bh1, bl1);
                    mma_3xtf32(acc[nt], ah[ks], al[ks], bh0, bh1, bl0, bl1);
                }
            }
            const int col = 32 * j + 8 * t;
            float* dst = S + g * FP + slab_idx<N>(col / N, col % N);
            *reinterpret_cast<float4*>(dst) = make_float4(acc[0][0], acc[1][0], acc[2][0], acc[3][0]);
            *reinterpret_cast<float4*>(dst + 4) = make_float4(acc[0][1], acc[1][1], acc[2][1], acc[3][1]);
            if (TC == 16) {
                *reinterpret_cast<float4*>(dst + 8 * FP) = make_float4(acc[0][2], acc[1][2], acc[2][2], acc[3][2]);
                *reinterpret_cast<float4*>(dst + 8 * FP + 4) = make_float4(acc[0][3], acc[1][3], acc[2][3], acc[3][3]);
            }
        }
    }
    // the constants of the next pass are requested before the barrier, so that their L2 latency overlaps the wait
    float chr_[MT][KN][4];
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
        for (int ks = 0; ks < KN; ++ks) {
            const float* r0 = ch + (16 * mt + g) * N + 8 * ks + t;
            chr_[mt][ks][0] = __ldg(r0);
            chr_[mt][ks][1] = __ldg(r0 + 8 * N);
            chr_[mt][ks][2] = __ldg(r0 + 4);
            chr_[mt][ks][3] = __ldg(r0 + 8 * N + 4);
        }
    // the staging buffers are free now: each warp fills its own with the next piece of the slab's x (the epilogue's high = x - low)
    constexpr int XP = (8 * T * 40 < TC * HW ? 8 * T * 40 : TC * HW), PW = XP / 8;   // floats of x staged: in all, per warp
    {
        const long long off = v * (long long)T * HW + (long long)t0 * HW;
        float* stg = S + TC * FP + warp * (T * 40);
        for (int c = lane; c < PW / 4; c += 32) cp_async16(stg + 4 * c, x + off + warp * PW + 4 * c);
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    __syncthreads();
    // ---- H pass, in place: a warp owns whole column blocks (frame tt, 8 columns): it reads the block (all rows), multiplies
    // it by every row tile of Ch (registers) and writes it back, so that no other warp ever touches the block.  The loads of
    // the next block are issued before the MMAs of this one.
    {
        constexpr int ITEMS = TC * KN;
        uint32_t ah[MT][KN][4], al[MT][KN][4];
#pragma unroll
        for (int mt = 0; mt < MT; ++mt)
#pragma unroll
            for (int ks = 0; ks < KN; ++ks) {
                split_tf32(chr_[mt][ks][0], ah[mt][ks][0], al[mt][ks][0]);
                split_tf32(chr_[mt][ks][1], ah[mt][ks][1], al[mt][ks][1]);
                split_tf32(chr_[mt][ks][2], ah[mt][ks][2], al[mt][ks][2]);
                split_tf32(chr_[mt][ks][3], ah[mt][ks][3], al[mt][ks][3]);
            }
        float b[KN][2], bn[KN][2];
        auto fetch = [&](float (&d)[KN][2], int it) {
            const float* fr = S + (it / KN) * FP;
            const int w = 8 * (it % KN) + g;
#pragma unroll
            for (int ks = 0; ks < KN; ++ks) {
                d[ks][0] = fr[slab_idx<N>(8 * ks + t, w)];
                d[ks][1] = fr[slab_idx<N>(8 * ks + t + 4, w)];
            }
        };
        fetch(b, warp);
        for (int it = warp; it < ITEMS; it += 8) {
            __syncwarp();
            if (it + 8 < ITEMS) fetch(bn, it + 8);
            float acc[MT][2][4];                    // per row tile two chains: even / odd k-steps
#pragma unroll
            for (int mt = 0; mt < MT; ++mt)
#pragma unroll
                for (int c = 0; c < 2; ++c) acc[mt][c][0] = acc[mt][c][1] = acc[mt][c][2] = acc[mt][c][3] = 0.f;
#pragma unroll
            for (int ks = 0; ks < KN; ++ks) {
                uint32_t bh0, bl0, bh1, bl1;
                split_tf32(b[ks][0], bh0, bl0);
                split_tf32(b[ks][1], bh1, bl1);
#pragma unroll
                for (int mt = 0; mt < MT; ++mt) mma_3xtf32(acc[mt][ks & 1], ah[mt][ks], al[mt][ks], bh0, bh1, bl0, bl1);
            }
            float* fr = S + (it / KN) * FP;
            const int w = 8 * (it % KN) + 2 * t;
#pragma unroll
            for (int mt = 0; mt < MT; ++mt) {
                *reinterpret_cast<float2*>(fr + slab_idx<N>(16 * mt + g, w)) =
                    make_float2(acc[mt][0][0] + acc[mt][1][0], acc[mt][0][1] + acc[mt][1][1]);
                *reinterpret_cast<float2*>(fr + slab_idx<N>(16 * mt + g + 8, w)) =
                    make_float2(acc[mt][0][2] + acc[mt][1][2], acc[mt][0][3] + acc[mt][1][3]);
            }
#pragma unroll
            for (int ks = 0; ks < KN; ++ks) { b[ks][0] = bn[ks][0]; b[ks][1] = bn[ks][1]; }
        }
    }
    float2 cwr_[KN][KN];                            // Cw^T fragments, requested before the barrier (8-byte aligned pairs)
#pragma unroll
    for (int nt = 0; nt < KN; ++nt)
#pragma unroll
        for (int ks = 0; ks < KN; ++ks)
            cwr_[nt][ks] = __ldg(reinterpret_cast<const float2*>(cw + (8 * nt + g) * N + 8 * ks + 2 * t));
    __syncthreads();
    // ---- W pass, in place as well: a warp owns 16 rows of a frame: A = those rows, B = Cw^T (all of it, registers).  Within
    // a k-step the MMA's k index t <-> w = 2t and t + 4 <-> w = 2t + 1, so that a lane's two A values of a row are one
    // float2 (conflict-free, half the loads).
    {
        constexpr int ITEMS = TC * MT;
        uint32_t bh[KN][KN][2], bl[KN][KN][2];      // [n-tile][k-step]
#pragma unroll
        for (int nt = 0; nt < KN; ++nt)
#pragma unroll
            for (int ks = 0; ks < KN; ++ks) {
                split_tf32(cwr_[nt][ks].x, bh[nt][ks][0], bl[nt][ks][0]);
                split_tf32(cwr_[nt][ks].y, bh[nt][ks][1], bl[nt][ks][1]);
            }
        for (int it = warp; it < ITEMS; it += 8) {
            float* fr = S + (it / MT) * FP;
            const int r0 = 16 * (it % MT) + g;
            float2 a[KN][2];
#pragma unroll
            for (int ks = 0; ks < KN; ++ks) {
                a[ks][0] = *reinterpret_cast<const float2*>(fr + slab_idx<N>(r0, 8 * ks + 2 * t));
                a[ks][1] = *reinterpret_cast<const float2*>(fr + slab_idx<N>(r0 + 8, 8 * ks + 2 * t));
            }
            float acc[KN][4];
#pragma unroll
            for (int nt = 0; nt < KN; ++nt) acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.f;
#pragma unroll
            for (int ks = 0; ks < KN; ++ks) {
                uint32_t ah[4], al[4];
                split_tf32(a[ks][0].x, ah[0], al[0]);
                split_tf32(a[ks][1].x, ah[1], al[1]);
                split_tf32(a[ks][0].y, ah[2], al[2]);
                split_tf32(a[ks][1].y, ah[3], al[3]);
#pragma unroll
                for (int nt = 0; nt < KN; ++nt)    // the n-tiles are independent accumulation chains
                    mma_3xtf32(acc[nt], ah, al, bh[nt][ks][0], bh[nt][ks][1], bl[nt][ks][0], bl[nt][ks][1]);
            }
            __syncwarp();
#pragma unroll
            for (int nt = 0; nt < KN; ++nt) {
                const int w = 8 * nt + 2 * t;
                *reinterpret_cast<float2*>(fr + slab_idx<N>(r0, w)) = make_float2(acc[nt][0], acc[nt][1]);
                *reinterpret_cast<float2*>(fr + slab_idx<N>(r0 + 8, w)) = make_float2(acc[nt][2], acc[nt][3]);
            }
        }
    }
    // the part of x that is not staged is requested before the barrier too
    constexpr int Q = TC * HW / 4, Q0 = XP / 4, NG = (Q - Q0 + 255) / 256;   // float4 per slab; staged; per thread from global
    const long long xoff = v * (long long)T * HW + (long long)t0 * HW;
    float4 xg[NG > 0 ? NG : 1];
#pragma unroll
    for (int u = 0; u < NG; ++u)
        if (Q0 + threadIdx.x + 256 * u < Q) xg[u] = __ldg(reinterpret_cast<const float4*>(x + xoff) + Q0 + threadIdx.x + 256 * u);
    cp_async_wait_all();                            // this thread's share of the staged x has landed
    __syncthreads();
    // ---- low = slab, high = x - low: 128-bit accesses, whole lines; the head of x comes from the staging area
    {
        const float* xst = S + TC * FP;
        auto emit = [&](int i, const float4& xx) {
            const int e = 4 * i, tt = e / HW, rem = e % HW;
            const float4 l = *reinterpret_cast<const float4*>(S + tt * FP + slab_idx<N>(rem / N, rem % N));
            reinterpret_cast<float4*>(lo + xoff)[i] = l;
            reinterpret_cast<float4*>(hi + xoff)[i] = make_float4(xx.x - l.x, xx.y - l.y, xx.z - l.z, xx.w - l.w);
        };
        for (int i = threadIdx.x; i < Q0; i += 256) {
            const int w = (4 * i) / PW;                   // the warp that staged this piece, in its own buffer
            emit(i, *reinterpret_cast<const float4*>(xst + w * (T * 40) + (4 * i - w * PW)));
        }
#pragma unroll
        for (int u = 0; u < NG; ++u)
            if (Q0 + threadIdx.x + 256 * u < Q) emit(Q0 + threadIdx.x + 256 * u, xg[u]);
    }
}

template <int T, int N>
static int freq_split3d_mma_launch(const float* x, const float* ct, const float* ch, const float* cw, float* lo, float* hi,
                                   int64_t vols, cudaStream_t st) {
    constexpr int TC = T >= 16 ? 16 : 8;
    constexpr size_t smem = ((size_t)TC * ((N == 32 ? N * 32 : N * 24) + 4) + 8 * T * 40) * sizeof(float);   // slab + staging
    HV_CUDA(cudaFuncSetAttribute(freq_split3d_mma_kernel<T, N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    freq_split3d_mma_kernel<T, N><<<(unsigned)(vols * (T / TC)), 256, smem, st>>>(x, ct, ch, cw, lo, hi);
    HV_LAUNCH_CHECK();
    return HV_OK;
}

extern "C" int hv_freq_split3d(const float* x, const float* ct, const float* ch, const float* cw,
                               float* lo, float* hi, int64_t vols, int T, int H, int W, void* stream) {
    HV_REQUIRE(vols >= 0 && T > 0 && H > 0 && W > 0, "hv_freq_split3d: bad sizes");
    if (vols == 0) return HV_OK;
    int maxn = T > H ? T : H;
    maxn = maxn > W ? maxn : W;
    HV_REQUIRE(maxn <= 64, "hv_freq_split3d: axis length %d > 64 unsupported", maxn);
    const int MAXN = maxn <= 32 ? 32 : 64;
    size_t smem = (size_t)(3 * MAXN * MAXN + (size_t)T * H * (W + 1)) * sizeof(float);
    HV_REQUIRE(smem <= 227 * 1024, "hv_freq_split3d: volume %dx%dx%d needs %zu B of shared memory (> 227 KB)", T, H, W, smem);
    cudaStream_t st = hv_stream(stream);
    // tensor-core path (HV_FREQ_TC=0 keeps the SIMT kernels, =1 forces the tensor kernel, for A/B runs); 128-bit accesses need 16-byte aligned volumes
    const char* tc_env = getenv("HV_FREQ_TC");
    const bool aligned16 = ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(lo) | reinterpret_cast<uintptr_t>(hi)) & 15) == 0;
    // a 32^3 launch that cannot fill the SMs with 16-frame slabs stays on the SIMT kernel (8-frame slabs, twice the CTAs):
    // 1040 against 770 GB/s at 32 volumes, 1457 against 1913 at 128 (`profiles/r02_x_bench_freq.txt`)
    const bool too_small = T == 32 && H == 32 && vols * 2 < hv_num_sms() && !(tc_env && tc_env[0] == '1');
    if (H == W && aligned16 && !too_small && !(tc_env && tc_env[0] == '0')) {
#define HV_FREQ_CASE(TT, NN) if (T == TT && H == NN) return freq_split3d_mma_launch<TT, NN>(x, ct, ch, cw, lo, hi, vols, st);
        HV_FREQ_CASE(32, 32) HV_FREQ_CASE(16, 32) HV_FREQ_CASE(8, 32) HV_FREQ_CASE(32, 16) HV_FREQ_CASE(16, 16) HV_FREQ_CASE(8, 16)
#undef HV_FREQ_CASE
    }
    if (T == 32 && H == 32 && W == 32) {  // circulant taps in registers (row 0 of each matrix)
        freq_split3d_slab32_kernel<<<(unsigned)(vols * 4), 256, 0, st>>>(x, ct, ch, cw, lo, hi);
        HV_LAUNCH_CHECK();
        return HV_OK;
    }
    if (MAXN == 32) {
        int TC = T;  // frames per slab: at least 4, at most 8 slabs' worth of parallelism per volume
        for (int d = 8; d >= 2; d >>= 1) if (T % d == 0 && T / d >= 4) { TC = T / d; break; }
        const size_t sm = (size_t)(3 * 32 * 32 + (size_t)TC * H * (W + 1)) * sizeof(float);
        HV_CUDA(cudaFuncSetAttribute(freq_split3d_slab_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
        freq_split3d_slab_kernel<<<(unsigned)(vols * (T / TC)), 256, sm, st>>>(x, ct, ch, cw, lo, hi, T, H, W, TC);
        HV_LAUNCH_CHECK();
        return HV_OK;
    }
    if (MAXN == 32) {
        HV_CUDA(cudaFuncSetAttribute(freq_split3d_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        freq_split3d_kernel<32><<<(unsigned)vols, 256, smem, st>>>(x, ct, ch, cw, lo, hi, T, H, W);
    } else {
        HV_CUDA(cudaFuncSetAttribute(freq_split3d_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        freq_split3d_kernel<64><<<(unsigned)vols, 256, smem, st>>>(x, ct, ch, cw, lo, hi, T, H, W);
    }
    HV_LAUNCH_CHECK();
    return HV_OK;
}
