// Building blocks of the backward pass of the blocks (`accelerator.backward(loss)` at
// train_amd.py:489 reaches, through autograd, the backward of nn.Linear / LayerNorm / GELU(tanh) of
// model/modules.py:651-1188).  The gradient GEMMs reuse the tcgen05 kernel of gemm_tc.cu (C = A W^T):
//   dX[M,K] = dY[M,N]  W[N,K]      -> A = dY,   W := W^T  (K x N, from hv_transpose2d)
//   dW[N,K] = dY^T[N,M] X[M,K]     -> A = dY^T, W := X^T  (both from hv_transpose2d), fp32 output
//   db[N]   = column sums of dY    -> hv_colsum
// Everything here is HBM-bound streaming work: 128-byte coalesced rows, fp32 accumulation, deterministic two-stage
// reductions (fixed grid, fixed order).
#include "hv_common.cuh"
#include <type_traits>

namespace {

template <typename T> __device__ __forceinline__ float ldf(const T* p, long long i);
template <> __device__ __forceinline__ float ldf<float>(const float* p, long long i) { return p[i]; }
template <> __device__ __forceinline__ float ldf<__nv_bfloat16>(const __nv_bfloat16* p, long long i) { return __bfloat162float(p[i]); }
template <typename T> __device__ __forceinline__ void stf(T* p, long long i, float v);
template <> __device__ __forceinline__ void stf<float>(float* p, long long i, float v) { p[i] = v; }
template <> __device__ __forceinline__ void stf<__nv_bfloat16>(__nv_bfloat16* p, long long i, float v) { p[i] = __float2bfloat16_rn(v); }

// ---------------------------------------------------------------- out[c][r] = in[r][c]
// 64 x 64 tiles through shared memory (pitch 65): reads and writes are both row-contiguous.
// bf16 fast path (even sizes / strides): 4-byte accesses on both sides — a warp reads 128 B of an input row and writes
// 128 B of an output row per instruction; the tile is stored transposed ([col][row], pitch 66) so that the write side
// reads row pairs as one word.
__global__ void __launch_bounds__(256) transpose_bf16x2_kernel(const __nv_bfloat16* __restrict__ in, __nv_bfloat16* __restrict__ out,
                                                               long long rows, long long cols, long long ld_in, long long ld_out) {
    __shared__ __align__(4) __nv_bfloat16 tile[64][66];
    const long long r0 = blockIdx.y * 64ll, c0 = blockIdx.x * 64ll;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
    for (int i = ty; i < 64; i += 8) {
        if (r0 + i < rows && c0 + 2 * tx < cols) {  // cols even: the pair is inside
            const __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(in + (r0 + i) * ld_in + c0 + 2 * tx);
            tile[2 * tx][i] = v.x; tile[2 * tx + 1][i] = v.y;
        }
    }
    __syncthreads();
    for (int j = ty; j < 64; j += 8)
        if (c0 + j < cols && r0 + 2 * tx < rows)
            *reinterpret_cast<__nv_bfloat162*>(out + (c0 + j) * ld_out + r0 + 2 * tx) = *reinterpret_cast<const __nv_bfloat162*>(&tile[j][2 * tx]);
}

// bf16 with sizes / strides that are multiples of 8 and 16-byte aligned bases: 16-byte global accesses on both sides.
// Load: lane = (row, 16-byte chunk) so that 8 lanes cover 128 B of an input row; the eight values are scattered into the
// transposed tile [col][row] (pitch 66: 2-way conflicts on these 2-byte stores).  Store: lane = (column, 8-row chunk),
// four conflict-free word reads of the transposed tile make one 16-byte store; 8 lanes cover 128 B of an output row.
// COLSUM: the eight values a lane is about to store are eight rows of one input column, so the column sum over the tile's
// 64 rows is one 8-lane reduction away; it goes to partial[row tile][column] (summed by colsum_final_kernel).
template <bool COLSUM>
__global__ void __launch_bounds__(256) transpose_bf16x8_kernel(const __nv_bfloat16* __restrict__ in, __nv_bfloat16* __restrict__ out,
                                                               long long rows, long long cols, long long ld_in, long long ld_out,
                                                               float* __restrict__ partial) {
    __shared__ __align__(4) __nv_bfloat16 tile[64][66];
    const long long r0 = blockIdx.y * 64ll, c0 = blockIdx.x * 64ll;
    const int ch = threadIdx.x & 7, rr = threadIdx.x >> 3;  // 8 chunks x 32 rows
    uint4 v[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const long long r = r0 + rr + 32 * h, c = c0 + 8 * ch;
        v[h] = (r < rows && c < cols) ? __ldg(reinterpret_cast<const uint4*>(in + r * ld_in + c)) : make_uint4(0u, 0u, 0u, 0u);
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const uint32_t w[4] = {v[h].x, v[h].y, v[h].z, v[h].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const __nv_bfloat162 p = *reinterpret_cast<const __nv_bfloat162*>(&w[j]);
            tile[8 * ch + 2 * j][rr + 32 * h] = p.x;
            tile[8 * ch + 2 * j + 1][rr + 32 * h] = p.y;
        }
    }
    __syncthreads();
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int cc = rr + 32 * h;  // tile column = output row
        const long long c = c0 + cc, r = r0 + 8 * ch;
        const uint32_t* src = reinterpret_cast<const uint32_t*>(&tile[cc][8 * ch]);
        const bool inside = c < cols && r < rows;
        if (inside) *reinterpret_cast<uint4*>(out + c * ld_out + r) = make_uint4(src[0], src[1], src[2], src[3]);
        if (COLSUM) {
            float t = 0.f;
            if (inside) {
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&src[k]));
                    t += f.x + f.y;
                }
            }
            t += __shfl_xor_sync(0xffffffffu, t, 1); t += __shfl_xor_sync(0xffffffffu, t, 2); t += __shfl_xor_sync(0xffffffffu, t, 4);
            if (ch == 0 && c < cols) partial[(long long)blockIdx.y * cols + c] = t;
        }
    }
}

template <typename T>
__global__ void __launch_bounds__(256) transpose_kernel(const T* __restrict__ in, T* __restrict__ out, long long rows, long long cols,
                                                        long long ld_in, long long ld_out) {
    __shared__ T tile[64][65];
    const long long r0 = blockIdx.y * 64ll, c0 = blockIdx.x * 64ll;
    const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;  // 64 x 4
    for (int i = ty; i < 64; i += 4)
        if (r0 + i < rows && c0 + tx < cols) tile[i][tx] = in[(r0 + i) * ld_in + c0 + tx];
    __syncthreads();
    for (int i = ty; i < 64; i += 4)
        if (c0 + i < cols && r0 + tx < rows) out[(c0 + i) * ld_out + r0 + tx] = tile[tx][i];
}

// ---------------------------------------------------------------- column sums, stage 1: partial[s][n] over a row slice
template <typename T>
__global__ void __launch_bounds__(256) colsum_partial_kernel(const T* __restrict__ x, long long M, int N, long long ld,
                                                             float* __restrict__ partial) {
    __shared__ float sh[8][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int n = blockIdx.x * 32 + tx;
    float acc = 0.f;
    if (n < N)
        for (long long m = blockIdx.y * 8ll + ty; m < M; m += (long long)gridDim.y * 8) acc += ldf<T>(x, m * ld + n);
    sh[ty][tx] = acc;
    __syncthreads();
    if (ty == 0 && n < N) {
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) s += sh[i][tx];
        partial[(long long)blockIdx.y * N + n] = s;
    }
}
// bf16, two columns per thread (a warp reads 128 B of a row); N and ld even
__global__ void __launch_bounds__(256) colsum_partial_bf16x2_kernel(const __nv_bfloat16* __restrict__ x, long long M, int N, long long ld,
                                                                    float* __restrict__ partial) {
    __shared__ float sh[8][65];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int n = blockIdx.x * 64 + 2 * tx;
    float a0 = 0.f, a1 = 0.f;
    if (n < N)
        for (long long m = blockIdx.y * 8ll + ty; m < M; m += (long long)gridDim.y * 8) {
            const float2 v = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(x + m * ld + n));
            a0 += v.x; a1 += v.y;
        }
    sh[ty][2 * tx] = a0; sh[ty][2 * tx + 1] = a1;
    __syncthreads();
    if (ty < 2 && n < N) {
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) s += sh[i][2 * tx + ty];
        partial[(long long)blockIdx.y * N + n + ty] = s;
    }
}
// bf16, eight columns per thread (16-byte loads, a warp reads 512 B of a row), eight rows in flight per thread;
// N, ld multiples of 8 and a 16-byte aligned base
__global__ void __launch_bounds__(256) colsum_partial_bf16x8_kernel(const __nv_bfloat16* __restrict__ x, long long M, int N, long long ld,
                                                                    float* __restrict__ partial) {
    __shared__ float sh[8][256 + 8];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int n = blockIdx.x * 256 + 8 * tx;
    float acc[8] = {};
    if (n < N) {
        const long long step = (long long)gridDim.y * 8;
        long long m = blockIdx.y * 8ll + ty;
        const __nv_bfloat16* p = x + n;
        for (; m + 7 * step < M; m += 8 * step) {  // eight 16-byte loads in flight per thread (64 KB per SM at 2 CTAs)
            uint4 v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) v[u] = __ldg(reinterpret_cast<const uint4*>(p + (m + u * step) * ld));
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const uint32_t w[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[j]));
                    acc[2 * j] += f.x; acc[2 * j + 1] += f.y;
                }
            }
        }
        for (; m < M; m += step) {
            const uint4 v = __ldg(reinterpret_cast<const uint4*>(p + m * ld));
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[j]));
                acc[2 * j] += f.x; acc[2 * j + 1] += f.y;
            }
        }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) sh[ty][8 * tx + j] = acc[j];
    __syncthreads();
    const int c = threadIdx.x, nn = blockIdx.x * 256 + c;
    if (nn < N) {
        float t = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) t += sh[i][c];
        partial[(long long)blockIdx.y * N + nn] = t;
    }
}
__global__ void colsum_final_kernel(const float* __restrict__ partial, int S, int N, float* __restrict__ out) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    float s = 0.f;
    for (int i = 0; i < S; ++i) s += partial[(long long)i * N + n];
    out[n] = s;
}

// colsum_final_kernel with up to three destination ranges (columns [0, n0) -> o0, [n0, n0 + n1) -> o1, the rest -> o2) and an
// optional += : bias gradients go straight into their slots of the gradient arena (q / k / v biases are separate tensors)
__global__ void colsum_final_acc_kernel(const float* __restrict__ partial, int S, int N, float* __restrict__ o0, int n0,
                                        float* __restrict__ o1, int n1, float* __restrict__ o2, int accumulate) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    float s = 0.f;
    for (int i = 0; i < S; ++i) s += partial[(long long)i * N + n];
    float* dst = n < n0 ? o0 + n : (n < n0 + n1 ? o1 + (n - n0) : o2 + (n - n0 - n1));
    *dst = accumulate ? *dst + s : s;
}

// the same reduction for many slices (S in the hundreds): 32 columns x 8 row groups per CTA, fixed summation order
__global__ void __launch_bounds__(256) colsum_final_wide_kernel(const float* __restrict__ partial, int S, int N, float* __restrict__ out) {
    __shared__ float sh[8][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5, n = blockIdx.x * 32 + tx;
    float a = 0.f;
    if (n < N)
        for (int i = ty; i < S; i += 8) a += partial[(long long)i * N + n];
    sh[ty][tx] = a;
    __syncthreads();
    if (ty == 0 && n < N) {
        float t = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) t += sh[i][tx];
        out[n] = t;
    }
}

// same wide reduction, the first `split` sums going to out_a and the rest to out_b (dw / db of a norm; either may be NULL)
__global__ void __launch_bounds__(256) colsum_final2_kernel(const float* __restrict__ partial, int S, int N, int split,
                                                            float* __restrict__ out_a, float* __restrict__ out_b, int accumulate = 0) {
    __shared__ float sh[8][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5, n = blockIdx.x * 32 + tx;
    float a = 0.f;
    if (n < N)
        for (int i = ty; i < S; i += 8) a += partial[(long long)i * N + n];
    sh[ty][tx] = a;
    __syncthreads();
    if (ty == 0 && n < N) {
        float t = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) t += sh[i][tx];
        if (n < split) { if (out_a) out_a[n] = accumulate ? out_a[n] + t : t; }
        else if (out_b) out_b[n - split] = accumulate ? out_b[n - split] + t : t;
    }
}

// ---------------------------------------------------------------- dx = dy * d/dx gelu_tanh(x)
__device__ __forceinline__ float gelu_tanh_grad(float x) {
    const float k0 = 0.7978845608028654f, k1 = 0.044715f;
    const float x2 = x * x, u = k0 * (x + k1 * x2 * x), t = tanhf(u);
    return 0.5f * (1.0f + t) + 0.5f * x * (1.0f - t * t) * k0 * (1.0f + 3.0f * k1 * x2);
}
template <typename TI, typename TO>
__global__ void gelu_bwd_kernel(const TI* __restrict__ pre, const TI* __restrict__ dy, TO* __restrict__ dx, long long n) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        stf<TO>(dx, i, ldf<TI>(dy, i) * gelu_tanh_grad(ldf<TI>(pre, i)));
}

// bf16 -> bf16, 8 elements (16 bytes) per thread
__global__ void gelu_bwd_bf16x8_kernel(const uint4* __restrict__ pre, const uint4* __restrict__ dy, uint4* __restrict__ dx, long long n8) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
        const uint4 p = pre[i], d = dy[i];
        const uint32_t pw[4] = {p.x, p.y, p.z, p.w}, dw[4] = {d.x, d.y, d.z, d.w};
        uint32_t o[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float p0 = __uint_as_float(pw[k] << 16), p1 = __uint_as_float(pw[k] & 0xffff0000u);
            const float d0 = __uint_as_float(dw[k] << 16), d1 = __uint_as_float(dw[k] & 0xffff0000u);
            o[k] = hv_pack_bf16(d0 * gelu_tanh_grad(p0), d1 * gelu_tanh_grad(p1));
        }
        dx[i] = make_uint4(o[0], o[1], o[2], o[3]);
    }
}

// ---------------------------------------------------------------- LayerNorm backward (rows of D <= 2048, D % 128 == 0)
// y = xhat * w + b, xhat = (x - mean) * rstd.  One warp per row, the row in registers:
//   g = dy * w;  dx = rstd * (g - mean(g) - xhat * mean(g * xhat))      (written or accumulated into dx)
//   dw += dy * xhat, db += dy   -> per-CTA partial sums in shared memory -> partial[cta][2][D] -> hv_colsum-style final
template <int NV, typename TDY>
__global__ void __launch_bounds__(256, 2) ln_bwd_kernel(const float* __restrict__ x, const TDY* __restrict__ dy, const float* __restrict__ w,
                                                     float eps, long long rows, int D, float* __restrict__ dx, int accumulate,
                                                     float* __restrict__ partial) {
    extern __shared__ __align__(16) float sh[];  // [8 warps][2][D] dw / db partials: each lane owns its columns of its warp's rows,
                                                 // which keeps 2 x NV x 4 accumulators out of the register file (2 CTAs per SM)
    for (int i = threadIdx.x; i < 16 * D; i += 256) sh[i] = 0.f;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float inv_d = 1.0f / (float)D;
    float4* const accw = reinterpret_cast<float4*>(sh + (size_t)warp * 2 * D) + lane;
    float4* const accb = reinterpret_cast<float4*>(sh + (size_t)warp * 2 * D + D) + lane;
    for (long long r = blockIdx.x * 8ll + warp; r < rows; r += (long long)gridDim.x * 8) {
        float xv[NV * 4], gv[NV * 4];
        float s = 0.f;
        // every global read of the row is issued before the first reduction: dy stays raw in registers, the dx row that an
        // accumulating call reads back at the end is pulled into L2 now
        using DyVec = typename std::conditional<sizeof(TDY) == 4, float4, uint2>::type;
        DyVec dyraw[NV];
#pragma unroll
        for (int j = 0; j < NV; ++j) dyraw[j] = reinterpret_cast<const DyVec*>(dy + r * D)[j * 32 + lane];
        if (accumulate) {
#pragma unroll
            for (int j = 0; j < NV; ++j) asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const float4*>(dx + r * D) + j * 32 + lane));
        }
#pragma unroll
        for (int j = 0; j < NV; ++j) {
            const float4 v = reinterpret_cast<const float4*>(x + r * D)[j * 32 + lane];
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
            // four consecutive columns per lane: one 16-byte (fp32) or 8-byte (bf16) load of dy, one 16-byte load of w
            float dv[4];
            if constexpr (sizeof(TDY) == 4) {
                const float4 t = *reinterpret_cast<const float4*>(&dyraw[j]);
                dv[0] = t.x; dv[1] = t.y; dv[2] = t.z; dv[3] = t.w;
            } else {
                const uint2 t = *reinterpret_cast<const uint2*>(&dyraw[j]);
                dv[0] = __uint_as_float(t.x << 16); dv[1] = __uint_as_float(t.x & 0xffff0000u);
                dv[2] = __uint_as_float(t.y << 16); dv[3] = __uint_as_float(t.y & 0xffff0000u);
            }
            const float4 wv4 = w ? __ldg(reinterpret_cast<const float4*>(w) + j * 32 + lane) : make_float4(1.f, 1.f, 1.f, 1.f);
            const float wv[4] = {wv4.x, wv4.y, wv4.z, wv4.w};
            float dwv[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float d = dv[e];
                const float xh = xv[4 * j + e] * rstd;
                const float g = d * wv[e];
                xv[4 * j + e] = xh; gv[4 * j + e] = g;
                c1 += g; c2 += g * xh;
                dwv[e] = d * xh;
            }
            float4 aw = accw[j * 32], ab = accb[j * 32];
            aw.x += dwv[0]; aw.y += dwv[1]; aw.z += dwv[2]; aw.w += dwv[3];
            ab.x += dv[0]; ab.y += dv[1]; ab.z += dv[2]; ab.w += dv[3];
            accw[j * 32] = aw; accb[j * 32] = ab;
        }
        c1 = hv_warp_sum(c1) * inv_d; c2 = hv_warp_sum(c2) * inv_d;
#pragma unroll
        for (int j = 0; j < NV; ++j) {
            float4 o;
            o.x = rstd * (gv[4 * j] - c1 - xv[4 * j] * c2); o.y = rstd * (gv[4 * j + 1] - c1 - xv[4 * j + 1] * c2);
            o.z = rstd * (gv[4 * j + 2] - c1 - xv[4 * j + 2] * c2); o.w = rstd * (gv[4 * j + 3] - c1 - xv[4 * j + 3] * c2);
            float4* dst = reinterpret_cast<float4*>(dx + r * D) + j * 32 + lane;
            if (accumulate) { const float4 p = *dst; o.x += p.x; o.y += p.y; o.z += p.z; o.w += p.w; }
            *dst = o;
        }
    }
    // the eight warps' partials are added in a fixed order (deterministic)
    __syncthreads();
    for (int i = threadIdx.x; i < 2 * D; i += 256) {
        float t = 0.f;
#pragma unroll
        for (int wv = 0; wv < 8; ++wv) t += sh[(size_t)wv * 2 * D + i];
        partial[(long long)blockIdx.x * 2 * D + i] = t;
    }
}

// ---------------------------------------------------------------- per-head LayerNorm(64) backward (QK-norm)
// q (or k) = LayerNorm_64(q_raw) * w + b per (token, head) — diffusers Attention(qk_norm="layer_norm"), eps 1e-6
// (modules.py:699-707).  Row r = (token, head) lives at x + token*ld + head*64; one warp per row, two columns per lane.
template <typename T>
__global__ void __launch_bounds__(256) headnorm_bwd_kernel(const T* __restrict__ x, long long ldx, const T* __restrict__ dy, long long lddy,
                                                           const float* __restrict__ w, float eps, long long tokens, int heads,
                                                           T* __restrict__ dx, long long lddx, float* __restrict__ partial) {
    __shared__ float sh[8][128];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float w0 = w ? w[2 * lane] : 1.f, w1 = w ? w[2 * lane + 1] : 1.f;
    float dw0 = 0.f, dw1 = 0.f, db0 = 0.f, db1 = 0.f;
    const long long rows = tokens * heads;
    for (long long r = blockIdx.x * 8ll + warp; r < rows; r += (long long)gridDim.x * 8) {
        const long long t = r / heads;
        const int c = (int)(r - t * heads) * 64 + 2 * lane;
        float x0 = ldf<T>(x, t * ldx + c), x1 = ldf<T>(x, t * ldx + c + 1);
        const float d0 = ldf<T>(dy, t * lddy + c), d1 = ldf<T>(dy, t * lddy + c + 1);
        const float mean = hv_warp_sum(x0 + x1) * (1.f / 64.f);
        x0 -= mean; x1 -= mean;
        const float rstd = rsqrtf(hv_warp_sum(x0 * x0 + x1 * x1) * (1.f / 64.f) + eps);
        x0 *= rstd; x1 *= rstd;
        const float g0 = d0 * w0, g1 = d1 * w1;
        const float c1 = hv_warp_sum(g0 + g1) * (1.f / 64.f), c2 = hv_warp_sum(g0 * x0 + g1 * x1) * (1.f / 64.f);
        stf<T>(dx, t * lddx + c, rstd * (g0 - c1 - x0 * c2));
        stf<T>(dx, t * lddx + c + 1, rstd * (g1 - c1 - x1 * c2));
        dw0 += d0 * x0; dw1 += d1 * x1; db0 += d0; db1 += d1;
    }
    sh[warp][2 * lane] = dw0; sh[warp][2 * lane + 1] = dw1; sh[warp][64 + 2 * lane] = db0; sh[warp][64 + 2 * lane + 1] = db1;
    __syncthreads();
    if (threadIdx.x < 128) {
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) s += sh[i][threadIdx.x];
        partial[(long long)blockIdx.x * 128 + threadIdx.x] = s;
    }
}

// bf16 with 16-byte alignment: eight lanes per row (one 16-byte load of x and of dy per lane), four rows per warp, so the
// four row reductions cost three shuffles each for four rows at once
__global__ void __launch_bounds__(512, 2) headnorm_bwd_bf16x8_kernel(const __nv_bfloat16* __restrict__ x, long long ldx,
                                                                  const __nv_bfloat16* __restrict__ dy, long long lddy,
                                                                  const float* __restrict__ w, float eps, long long tokens, int heads,
                                                                  __nv_bfloat16* __restrict__ dx, long long lddx, float* __restrict__ partial) {
    __shared__ float sh[16][128];   // 16 warps: 32 warps per SM at two CTAs (16 at 256 threads left the loads latency-bound)
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, sub = lane & 7, rg = lane >> 3;
    float wv[8], dwa[8] = {}, dba[8] = {};
#pragma unroll
    for (int j = 0; j < 8; ++j) wv[j] = w ? w[8 * sub + j] : 1.f;
    const long long rows = tokens * heads;
    auto red8 = [](float v) {
        v += __shfl_xor_sync(0xffffffffu, v, 1); v += __shfl_xor_sync(0xffffffffu, v, 2); v += __shfl_xor_sync(0xffffffffu, v, 4);
        return v;
    };
    // two row groups per trip, the loads of both issued before any arithmetic: one 16-byte load of x and of dy per thread
    // and twelve dependent shuffles per group left too few bytes in flight (49 % of the HBM copy rate)
    struct Grp { bool valid; long long t; int c; uint4 a, b; };
    auto fetch = [&](long long r0) {
        Grp g;
        const long long r = r0 + rg;
        g.valid = r < rows;
        g.t = g.valid ? r / heads : 0;
        g.c = g.valid ? (int)(r - g.t * heads) * 64 + 8 * sub : 0;
        g.a = g.b = make_uint4(0u, 0u, 0u, 0u);
        if (g.valid) { g.a = *reinterpret_cast<const uint4*>(x + g.t * ldx + g.c); g.b = *reinterpret_cast<const uint4*>(dy + g.t * lddy + g.c); }
        return g;
    };
    auto process = [&](const Grp& gp) {
        float xv[8], dv[8];
        const uint32_t aw[4] = {gp.a.x, gp.a.y, gp.a.z, gp.a.w}, bw[4] = {gp.b.x, gp.b.y, gp.b.z, gp.b.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float2 fa = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&aw[j]));
            const float2 fb = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&bw[j]));
            xv[2 * j] = fa.x; xv[2 * j + 1] = fa.y; dv[2 * j] = fb.x; dv[2 * j + 1] = fb.y;
        }
        // the four row sums in ONE round of shuffles (three dependent steps instead of twelve): sum x, sum x^2, sum g, sum g x
        // (g = dy w); var = E[x^2] - mean^2 is accurate enough for a bf16 result (the fp32 kernel keeps the two-pass form)
        float g[8], s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) { g[j] = dv[j] * wv[j]; s0 += xv[j]; s1 = fmaf(xv[j], xv[j], s1); s2 += g[j]; s3 = fmaf(g[j], xv[j], s3); }
#pragma unroll
        for (int o = 1; o < 8; o <<= 1) {
            s0 += __shfl_xor_sync(0xffffffffu, s0, o); s1 += __shfl_xor_sync(0xffffffffu, s1, o);
            s2 += __shfl_xor_sync(0xffffffffu, s2, o); s3 += __shfl_xor_sync(0xffffffffu, s3, o);
        }
        const float mean = s0 * (1.f / 64.f);
        const float rstd = rsqrtf(fmaxf(s1 * (1.f / 64.f) - mean * mean, 0.f) + eps);
        const float c1 = s2 * (1.f / 64.f), c2 = rstd * (s3 - mean * s2) * (1.f / 64.f);
#pragma unroll
        for (int j = 0; j < 8; ++j) xv[j] = (xv[j] - mean) * rstd;
        if (gp.valid) {
            uint32_t o[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const __nv_bfloat162 v = __floats2bfloat162_rn(rstd * (g[2 * j] - c1 - xv[2 * j] * c2), rstd * (g[2 * j + 1] - c1 - xv[2 * j + 1] * c2));
                o[j] = *reinterpret_cast<const uint32_t*>(&v);
            }
            *reinterpret_cast<uint4*>(dx + gp.t * lddx + gp.c) = make_uint4(o[0], o[1], o[2], o[3]);
#pragma unroll
            for (int j = 0; j < 8; ++j) { dwa[j] += dv[j] * xv[j]; dba[j] += dv[j]; }
        }
    };
    const long long stride = (long long)gridDim.x * 64;
    for (long long r0 = (blockIdx.x * 16ll + warp) * 4; r0 < rows; r0 += 2 * stride) {   // r0 is warp-uniform
        const Grp g0 = fetch(r0), g1 = fetch(r0 + stride);   // rows past the end: valid = false, zeros
        process(g0);
        process(g1);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        dwa[j] += __shfl_xor_sync(0xffffffffu, dwa[j], 8); dwa[j] += __shfl_xor_sync(0xffffffffu, dwa[j], 16);
        dba[j] += __shfl_xor_sync(0xffffffffu, dba[j], 8); dba[j] += __shfl_xor_sync(0xffffffffu, dba[j], 16);
    }
    if (rg == 0) {
#pragma unroll
        for (int j = 0; j < 8; ++j) { sh[warp][8 * sub + j] = dwa[j]; sh[warp][64 + 8 * sub + j] = dba[j]; }
    }
    __syncthreads();
    if (threadIdx.x < 128) {
        float t = 0.f;
#pragma unroll
        for (int i = 0; i < 16; ++i) t += sh[i][threadIdx.x];
        partial[(long long)blockIdx.x * 128 + threadIdx.x] = t;
    }
}

}  // namespace

// backward of the per-head LayerNorm(64) of q / k (QK-norm): x, dy, dx are (tokens, heads*64) views with their own leading
// dimensions and one dtype; dw, db (64 each, fp32) are summed over tokens AND heads (the affine is shared by all heads).
// workspace = (hv_layernorm_bwd_ctas() + 1) * 128 floats.
static int headnorm64_bwd_impl(const void* x, int64_t ldx, const void* dy, int64_t lddy, const float* w, float eps, int64_t tokens,
                               int heads, void* dx, int64_t lddx, int dtype, float* workspace, float* dw, float* db, int accumulate, void* stream) {
    HV_REQUIRE(tokens >= 0 && heads > 0 && x && dy && dx && workspace && (dtype == HV_F32 || dtype == HV_BF16), "hv_headnorm64_bwd: bad arguments");
    if (tokens == 0) return HV_OK;
    int ctas = hv_layernorm_bwd_ctas();
    const long long rows = tokens * heads;
    auto al16 = [](const void* p, int64_t ld) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0 && ld % 8 == 0; };
    const bool x8 = dtype == HV_BF16 && al16(x, ldx) && al16(dy, lddy) && al16(dx, lddx);
    const long long per_cta = x8 ? 64 : 8;
    if (ctas > (rows + per_cta - 1) / per_cta) ctas = (int)((rows + per_cta - 1) / per_cta);
    cudaStream_t st = hv_stream(stream);
    if (x8) headnorm_bwd_bf16x8_kernel<<<ctas, 512, 0, st>>>((const __nv_bfloat16*)x, ldx, (const __nv_bfloat16*)dy, lddy, w, eps, tokens, heads, (__nv_bfloat16*)dx, lddx, workspace);
    else if (dtype == HV_F32) headnorm_bwd_kernel<float><<<ctas, 256, 0, st>>>((const float*)x, ldx, (const float*)dy, lddy, w, eps, tokens, heads, (float*)dx, lddx, workspace);
    else headnorm_bwd_kernel<__nv_bfloat16><<<ctas, 256, 0, st>>>((const __nv_bfloat16*)x, ldx, (const __nv_bfloat16*)dy, lddy, w, eps, tokens, heads, (__nv_bfloat16*)dx, lddx, workspace);
    HV_LAUNCH_CHECK();
    if (dw || db) {
        colsum_final2_kernel<<<4, 256, 0, st>>>(workspace, ctas, 128, 64, dw, db, accumulate);
        HV_LAUNCH_CHECK();
    }
    return HV_OK;
}

extern "C" int hv_headnorm64_bwd(const void* x, int64_t ldx, const void* dy, int64_t lddy, const float* w, float eps, int64_t tokens,
                                 int heads, void* dx, int64_t lddx, int dtype, float* workspace, float* dw, float* db, void* stream) {
    return headnorm64_bwd_impl(x, ldx, dy, lddy, w, eps, tokens, heads, dx, lddx, dtype, workspace, dw, db, 0, stream);
}
// hv_headnorm64_bwd with dw / db ADDED to their destinations (accumulate != 0): the gradient arena slots of norm_q / norm_k
extern "C" int hv_headnorm64_bwd_acc(const void* x, int64_t ldx, const void* dy, int64_t lddy, const float* w, float eps, int64_t tokens,
                                     int heads, void* dx, int64_t lddx, int dtype, float* workspace, float* dw, float* db, int accumulate,
                                     void* stream) {
    return headnorm64_bwd_impl(x, ldx, dy, lddy, w, eps, tokens, heads, dx, lddx, dtype, workspace, dw, db, accumulate, stream);
}

extern "C" int hv_transpose2d(const void* in, void* out, int64_t rows, int64_t cols, int64_t ld_in, int64_t ld_out, int dtype, void* stream) {
    HV_REQUIRE(rows >= 0 && cols >= 0 && ld_in >= cols && ld_out >= rows && (dtype == HV_F32 || dtype == HV_BF16), "hv_transpose2d: bad arguments");
    if (rows == 0 || cols == 0) return HV_OK;
    dim3 grid((unsigned)((cols + 63) / 64), (unsigned)((rows + 63) / 64));
    if (dtype == HV_F32) transpose_kernel<float><<<grid, 256, 0, hv_stream(stream)>>>((const float*)in, (float*)out, rows, cols, ld_in, ld_out);
    else if (((rows | cols | ld_in | ld_out) & 7) == 0 && ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15) == 0)
        transpose_bf16x8_kernel<false><<<grid, 256, 0, hv_stream(stream)>>>((const __nv_bfloat16*)in, (__nv_bfloat16*)out, rows, cols, ld_in, ld_out, nullptr);
    else if (((rows | cols | ld_in | ld_out) & 1) == 0 && ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 3) == 0)
        transpose_bf16x2_kernel<<<grid, 256, 0, hv_stream(stream)>>>((const __nv_bfloat16*)in, (__nv_bfloat16*)out, rows, cols, ld_in, ld_out);
    else transpose_kernel<__nv_bfloat16><<<grid, 256, 0, hv_stream(stream)>>>((const __nv_bfloat16*)in, (__nv_bfloat16*)out, rows, cols, ld_in, ld_out);
    HV_LAUNCH_CHECK();
    return HV_OK;
}

// out = in^T and colsum[c] = sum_r in[r][c] in one pass over `in` (dY^T and the bias gradient of linear_backward).
// bf16 only, rows / cols / strides multiples of 8, 16-byte aligned bases (HV_EINVAL otherwise: the caller then uses
// hv_transpose2d + hv_colsum).  workspace: ceil(rows / 64) * cols floats.
extern "C" int hv_transpose2d_colsum(const void* in, void* out, int64_t rows, int64_t cols, int64_t ld_in, int64_t ld_out, int dtype,
                                     float* workspace, float* colsum, void* stream) {
    HV_REQUIRE(rows > 0 && cols > 0 && ld_in >= cols && ld_out >= rows && dtype == HV_BF16 && workspace && colsum, "hv_transpose2d_colsum: bad arguments");
    HV_REQUIRE(((rows | cols | ld_in | ld_out) & 7) == 0 && ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15) == 0,
               "hv_transpose2d_colsum: sizes / strides must be multiples of 8 and the bases 16-byte aligned");
    HV_REQUIRE(cols < (1ll << 31) && (rows + 63) / 64 < 65536, "hv_transpose2d_colsum: too large");
    dim3 grid((unsigned)((cols + 63) / 64), (unsigned)((rows + 63) / 64));
    cudaStream_t st = hv_stream(stream);
    transpose_bf16x8_kernel<true><<<grid, 256, 0, st>>>((const __nv_bfloat16*)in, (__nv_bfloat16*)out, rows, cols, ld_in, ld_out, workspace);
    HV_LAUNCH_CHECK();
    colsum_final_wide_kernel<<<(unsigned)((cols + 31) / 32), 256, 0, st>>>(workspace, (int)grid.y, (int)cols, colsum);
    HV_LAUNCH_CHECK();
    return HV_OK;
}

// number of row slices (= rows of fp32 workspace of width N) the column-sum entry points use
extern "C" int hv_colsum_slices() { return hv_num_sms() * 2; }

// out[n] = sum_m x[m][n]  (bias gradient of nn.Linear); workspace: hv_colsum_slices() * N floats
static int colsum_impl(const void* x, int64_t M, int N, int64_t ld, int dtype, float* workspace, float* out, int n0, float* out1, int n1,
                       float* out2, int accumulate, void* stream) {
    HV_REQUIRE(M >= 0 && N > 0 && ld >= N && workspace && out && (dtype == HV_F32 || dtype == HV_BF16), "hv_colsum: bad arguments");
    HV_REQUIRE(n0 >= 0 && n1 >= 0 && n0 + n1 <= N && (n0 >= N || out1) && (n0 + n1 >= N || out2), "hv_colsum_acc: destination ranges do not cover the columns");
    int S = hv_colsum_slices();
    const bool x2 = dtype == HV_BF16 && (N % 2 == 0) && (ld % 2 == 0) && (reinterpret_cast<uintptr_t>(x) & 3) == 0;
    const bool x8 = dtype == HV_BF16 && (N % 8 == 0) && (ld % 8 == 0) && (reinterpret_cast<uintptr_t>(x) & 15) == 0 && M >= 2048;
    const long long strips = x8 ? (N + 255) / 256 : x2 ? (N + 63) / 64 : (N + 31) / 32;
    long long sy = (long long)S / strips;  // keep the grid near 2 CTAs per SM overall
    if (sy < 1) sy = 1;
    if (sy > (M + 7) / 8) sy = (M + 7) / 8 > 0 ? (M + 7) / 8 : 1;
    S = (int)sy;
    dim3 grid((unsigned)strips, (unsigned)S);
    cudaStream_t st = hv_stream(stream);
    if (dtype == HV_F32) colsum_partial_kernel<float><<<grid, 256, 0, st>>>((const float*)x, M, N, ld, workspace);
    else if (x8) colsum_partial_bf16x8_kernel<<<grid, 256, 0, st>>>((const __nv_bfloat16*)x, M, N, ld, workspace);
    else if (x2) colsum_partial_bf16x2_kernel<<<grid, 256, 0, st>>>((const __nv_bfloat16*)x, M, N, ld, workspace);
    else colsum_partial_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)x, M, N, ld, workspace);
    HV_LAUNCH_CHECK();
    if (n0 >= N && !accumulate) colsum_final_kernel<<<(N + 255) / 256, 256, 0, st>>>(workspace, S, N, out);
    else colsum_final_acc_kernel<<<(N + 255) / 256, 256, 0, st>>>(workspace, S, N, out, n0, out1, n1, out2, accumulate);
    HV_LAUNCH_CHECK();
    return HV_OK;
}

extern "C" int hv_colsum(const void* x, int64_t M, int N, int64_t ld, int dtype, float* workspace, float* out, void* stream) {
    return colsum_impl(x, M, N, ld, dtype, workspace, out, N, nullptr, 0, nullptr, 0, stream);
}
// hv_colsum into up to three destinations (columns [0, n0) -> out0, [n0, n0 + n1) -> out1, the rest -> out2; n0 = N for one),
// added to what is there when accumulate != 0: bias gradients straight into the gradient arena (train_amd.py:489 accumulates
// into .grad the same way)
extern "C" int hv_colsum_acc(const void* x, int64_t M, int N, int64_t ld, int dtype, float* workspace, float* out0, int n0, float* out1,
                             int n1, float* out2, int accumulate, void* stream) {
    return colsum_impl(x, M, N, ld, dtype, workspace, out0, n0, out1, n1, out2, accumulate, stream);
}

// dx = dy * gelu_tanh'(pre)   (FeedForward's GELU(approximate="tanh"), modules.py:712-719 through diffusers)
extern "C" int hv_gelu_tanh_bwd(const void* pre, const void* dy, void* dx, int64_t n, int dtype_in, int dtype_out, void* stream) {
    HV_REQUIRE(n >= 0 && pre && dy && dx, "hv_gelu_tanh_bwd: bad arguments");
    if (n == 0) return HV_OK;
    long long grid = (n + 255) / 256;
    if (grid > (long long)hv_num_sms() * 16) grid = (long long)hv_num_sms() * 16;
    cudaStream_t st = hv_stream(stream);
    const bool vec = (n % 8 == 0) && ((reinterpret_cast<uintptr_t>(pre) | reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(dx)) & 15) == 0;
    if (dtype_in == HV_BF16 && dtype_out == HV_BF16 && vec) {
        long long g8 = (n / 8 + 255) / 256;
        if (g8 > (long long)hv_num_sms() * 16) g8 = (long long)hv_num_sms() * 16;
        gelu_bwd_bf16x8_kernel<<<(unsigned)g8, 256, 0, st>>>((const uint4*)pre, (const uint4*)dy, (uint4*)dx, n / 8);
    } else if (dtype_in == HV_BF16 && dtype_out == HV_BF16) gelu_bwd_kernel<__nv_bfloat16, __nv_bfloat16><<<(unsigned)grid, 256, 0, st>>>((const __nv_bfloat16*)pre, (const __nv_bfloat16*)dy, (__nv_bfloat16*)dx, n);
    else if (dtype_in == HV_F32 && dtype_out == HV_F32) gelu_bwd_kernel<float, float><<<(unsigned)grid, 256, 0, st>>>((const float*)pre, (const float*)dy, (float*)dx, n);
    else if (dtype_in == HV_BF16 && dtype_out == HV_F32) gelu_bwd_kernel<__nv_bfloat16, float><<<(unsigned)grid, 256, 0, st>>>((const __nv_bfloat16*)pre, (const __nv_bfloat16*)dy, (float*)dx, n);
    else HV_REQUIRE(false, "hv_gelu_tanh_bwd: unsupported dtype combination %d -> %d", dtype_in, dtype_out);
    HV_LAUNCH_CHECK();
    return HV_OK;
}

// number of CTAs (= rows of the [cta][2][D] fp32 workspace) hv_layernorm_bwd uses
extern "C" int hv_layernorm_bwd_ctas() { return hv_num_sms() * 2; }

// LayerNorm backward over dense fp32 rows: dx (written, or accumulated when accumulate != 0), dw, db.
// workspace: hv_layernorm_bwd_ctas() * 2 * D floats.  dw / db may be NULL (elementwise_affine=False).
extern "C" int hv_layernorm_bwd(const float* x, const void* dy, int dy_dtype, const float* w, float eps, int64_t rows, int D,
                                float* dx, int accumulate, float* workspace, float* dw, float* db, void* stream) {
    HV_REQUIRE(rows >= 0 && D > 0 && D % 128 == 0 && D <= 1024, "hv_layernorm_bwd: D=%d must be a multiple of 128 and <= 1024", D);
    HV_REQUIRE(x && dy && dx && workspace, "hv_layernorm_bwd: null argument");
    if (rows == 0) return HV_OK;
    int ctas = hv_layernorm_bwd_ctas();
    if (ctas > (rows + 7) / 8) ctas = (int)((rows + 7) / 8);
    cudaStream_t st = hv_stream(stream);
    const size_t smem = (size_t)16 * D * sizeof(float);
#define LNB(NV, T) do { HV_CUDA(cudaFuncSetAttribute(ln_bwd_kernel<NV, T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
                        ln_bwd_kernel<NV, T><<<ctas, 256, smem, st>>>(x, (const T*)dy, w, eps, rows, D, dx, accumulate, workspace); } while (0)
#define LNB_T(NV) do { if (dy_dtype == HV_F32) LNB(NV, float); else LNB(NV, __nv_bfloat16); } while (0)
    switch (D / 128) {
        case 1: LNB_T(1); break; case 2: LNB_T(2); break; case 3: LNB_T(3); break; case 4: LNB_T(4); break;
        case 5: LNB_T(5); break; case 6: LNB_T(6); break; case 7: LNB_T(7); break; default: LNB_T(8); break;
    }
#undef LNB_T
#undef LNB
    HV_LAUNCH_CHECK();
    if (dw || db) {  // [ctas][2*D] -> column sums; dw = first D columns, db = last D
        colsum_final2_kernel<<<(2 * D + 31) / 32, 256, 0, st>>>(workspace, ctas, 2 * D, D, dw, db);
        HV_LAUNCH_CHECK();
    }
    return HV_OK;
}
