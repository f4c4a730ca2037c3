// fp32 SIMT GEMM with the fused epilogues — the fp32 parity path (1e-4 vs the reference needs
// true fp32 products; tensor-core TF32/bf16 cannot give that) and the tiny-K / skinny-N linears
// (motion_embed K=16, camera_motion_patch_embed K=8, proj_out N=8/16).
#include "epilogue.cuh"

template <typename AT>
__global__ void __launch_bounds__(256) gemm_simt_kernel(const AT* __restrict__ A, long long lda, const float* __restrict__ W,
                                                        long long ldw, long long M, int N, int K, HvEpiDev epi) {
    constexpr int BM = 64, BN = 64, BK = 16;
    __shared__ float As[BK][BM + 4];
    __shared__ float Ws[BK][BN + 4];
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const long long m0 = blockIdx.y * (long long)BM;
    const int n0 = blockIdx.x * BN;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    for (int k0 = 0; k0 < K; k0 += BK) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int idx = tid + i * 256;
            const int row = idx >> 4, kk = idx & 15;
            const long long gm = m0 + row;
            const int gk = k0 + kk;
            As[kk][row] = (gm < M && gk < K) ? hv_ld(A + gm * lda + gk) : 0.f;
            const int gn = n0 + row;
            Ws[kk][row] = (gn < N && gk < K) ? W[(long long)gn * ldw + gk] : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < BK; ++kk) {
            const float4 a = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
            const float4 b = *reinterpret_cast<const float4*>(&Ws[kk][tx * 4]);
            const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        __syncthreads();
    }
    const int col0 = n0 + tx * 4;
    if (col0 >= N) return;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const long long r = m0 + ty * 4 + i;
        if (r < M) hv_epilogue_row<4>(epi, r, col0, N, acc[i]);
    }
}

// per-head LayerNorm (64 wide) applied in place on a dense [M, N] matrix: one warp per (row, head)
template <typename T>
__global__ void __launch_bounds__(256) qknorm_rows_kernel(T* __restrict__ x, long long ldo, long long M, int heads_total,
                                                          int q_heads, const float* __restrict__ qw, const float* __restrict__ qb,
                                                          const float* __restrict__ kw, const float* __restrict__ kb, float eps) {
    const int lane = threadIdx.x & 31;
    const long long wid = blockIdx.x * (long long)(blockDim.x >> 5) + (threadIdx.x >> 5);
    if (wid >= M * heads_total) return;
    const long long r = wid / heads_total;
    const int h = (int)(wid - r * heads_total);
    T* p = x + r * ldo + (long long)h * 64;
    const float* w = h < q_heads ? qw : kw;
    const float* b = h < q_heads ? qb : kb;
    float a0 = hv_ld(p + lane), a1 = hv_ld(p + 32 + lane);
    const float mean = hv_warp_sum(a0 + a1) * (1.f / 64.f);
    const float d0 = a0 - mean, d1 = a1 - mean;
    const float rstd = rsqrtf(hv_warp_sum(d0 * d0 + d1 * d1) * (1.f / 64.f) + eps);
    hv_st(p + lane, d0 * rstd * w[lane] + b[lane]);
    hv_st(p + 32 + lane, d1 * rstd * w[32 + lane] + b[32 + lane]);
}

int hv_qknorm_rows(const HvEpilogue& e, int64_t M, cudaStream_t st) {
    const int heads_total = (e.q_cols + e.k_cols) / 64;
    if (heads_total == 0 || M == 0) return HV_OK;
    const long long warps = (long long)M * heads_total;
    dim3 grid((unsigned)((warps + 7) / 8));
    if (e.out_dtype == HV_F32)
        qknorm_rows_kernel<float><<<grid, 256, 0, st>>>((float*)e.out, e.ldo, M, heads_total, e.q_cols / 64, e.qn_w, e.qn_b, e.kn_w, e.kn_b, e.qk_eps);
    else
        qknorm_rows_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((__nv_bfloat16*)e.out, e.ldo, M, heads_total, e.q_cols / 64, e.qn_w, e.qn_b, e.kn_w, e.kn_b, e.qk_eps);
    HV_LAUNCH_CHECK();
    return HV_OK;
}

extern "C" int hv_linear_f32(const void* A, int a_dtype, int64_t lda, const float* W, int64_t ldw,
                             int64_t M, int N, int K, const HvEpilogue* epi, void* stream) {
    HV_REQUIRE(M >= 0 && N > 0 && K > 0 && lda >= K && ldw >= K, "hv_linear_f32: bad sizes");
    HV_REQUIRE(a_dtype == HV_F32 || a_dtype == HV_BF16, "hv_linear_f32: bad A dtype");
    HvEpiDev d;
    int rc = hv_epilogue_prepare(epi, M, N, d, "hv_linear_f32");
    if (rc) return rc;
    if (M == 0) return HV_OK;
    const bool qk = d.e.mode == HV_EPI_QKNORM;
    if (qk) d.e.mode = HV_EPI_STORE;  // store (+bias) first, normalise heads in a second pass
    dim3 grid((unsigned)((N + 63) / 64), (unsigned)((M + 63) / 64));
    HV_REQUIRE(grid.y <= 65535, "hv_linear_f32: M too large");
    cudaStream_t st = hv_stream(stream);
    if (a_dtype == HV_F32) gemm_simt_kernel<float><<<grid, 256, 0, st>>>((const float*)A, lda, W, ldw, M, N, K, d);
    else gemm_simt_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)A, lda, W, ldw, M, N, K, d);
    HV_LAUNCH_CHECK();
    if (qk) return hv_qknorm_rows(d.e, M, st);
    return HV_OK;
}
