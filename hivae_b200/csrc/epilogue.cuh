// Fused GEMM epilogues shared by the SIMT fp32 GEMM and the tcgen05 bf16 GEMM.
// A thread calls hv_epilogue_row() with NC consecutive accumulator columns of one output row.
#pragma once
#include "hv_common.cuh"

struct HvEpiDev {
    HvEpilogue e;
    int vec_ok;   // out / res / bias / gate bases and strides allow 16-byte accesses
    int vec_ok2;  // the same for the optional second output
};

// host-side validation + alignment analysis
static inline int hv_epilogue_prepare(const HvEpilogue* epi, int64_t M, int N, HvEpiDev& d, const char* who) {
    HV_REQUIRE(epi != nullptr, "%s: epilogue is NULL", who);
    d.e = *epi;
    const HvEpilogue& e = d.e;
    HV_REQUIRE(e.mode >= HV_EPI_STORE && e.mode <= HV_EPI_QKNORM, "%s: bad epilogue mode %d", who, e.mode);
    auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
    bool ok = (N % 4 == 0) && (!e.bias || al16(e.bias));
    if (e.mode == HV_EPI_RESIDUAL) {
        HV_REQUIRE(e.res != nullptr && e.L > 0 && e.row_off >= 0 && e.row_off + e.L <= e.frame_rows, "%s: bad residual row map", who);
        HV_REQUIRE(M % e.L == 0, "%s: M=%lld is not a multiple of L=%d", who, (long long)M, e.L);
        HV_REQUIRE(!e.gate || e.cond_div > 0, "%s: cond_div must be > 0", who);
        ok = ok && al16(e.res) && (!e.gate || (al16(e.gate) && e.gate_ld % 4 == 0 && e.gate_col_mot % 4 == 0 && e.gate_col_img % 4 == 0));
    } else {
        HV_REQUIRE(e.out != nullptr && e.ldo >= N, "%s: bad output", who);
        HV_REQUIRE(e.out_dtype == HV_F32 || e.out_dtype == HV_BF16, "%s: bad out dtype", who);
        const int per16 = e.out_dtype == HV_F32 ? 4 : 8;
        ok = ok && al16(e.out) && (e.ldo % per16 == 0) && (N % per16 == 0);
        if (e.mode == HV_EPI_QKNORM) {
            HV_REQUIRE(e.q_cols % 64 == 0 && e.k_cols % 64 == 0 && e.q_cols + e.k_cols <= N && N % 64 == 0,
                       "%s: qk-norm needs 64-wide heads (q_cols=%d k_cols=%d N=%d)", who, e.q_cols, e.k_cols, N);
            HV_REQUIRE((e.q_cols == 0 || (e.qn_w && e.qn_b)) && (e.k_cols == 0 || (e.kn_w && e.kn_b)), "%s: qk-norm params missing", who);
        }
    }
    d.vec_ok = ok ? 1 : 0;
    d.vec_ok2 = 0;
    if (e.out2) {
        HV_REQUIRE(e.mode == HV_EPI_GELU || e.mode == HV_EPI_QKNORM, "%s: a second output needs the GELU or QKNORM epilogue", who);
        HV_REQUIRE(e.ldo2 >= N, "%s: bad second output", who);
        const int per16 = e.out_dtype == HV_F32 ? 4 : 8;
        d.vec_ok2 = (ok && al16(e.out2) && e.ldo2 % per16 == 0) ? 1 : 0;
    }
    return HV_OK;
}

template <int NC>
__device__ __forceinline__ void hv_epi_store_dense(const HvEpilogue& e, int vec_ok, long long r, int col0, int N, const float (&v)[NC],
                                                   void* out_ptr = nullptr, long long out_ld = 0) {
    void* const base = out_ptr ? out_ptr : e.out;
    const long long ld = out_ptr ? out_ld : e.ldo;
    if (e.out_dtype == HV_F32) {
        float* o = reinterpret_cast<float*>(base) + r * ld + col0;
        if (vec_ok && col0 + NC <= N) {
#pragma unroll
            for (int j = 0; j < NC; j += 4) *reinterpret_cast<float4*>(o + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
        } else {
#pragma unroll
            for (int j = 0; j < NC; ++j) if (col0 + j < N) o[j] = v[j];
        }
    } else {
        __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(base) + r * ld + col0;
        if (vec_ok && (NC % 8 == 0) && col0 + NC <= N) {
#pragma unroll
            for (int j = 0; j < NC; j += 8)
                *reinterpret_cast<uint4*>(o + j) = make_uint4(hv_pack_bf16(v[j], v[j + 1]), hv_pack_bf16(v[j + 2], v[j + 3]),
                                                              hv_pack_bf16(v[j + 4], v[j + 5]), hv_pack_bf16(v[j + 6], v[j + 7]));
        } else if (vec_ok && col0 + NC <= N) {
#pragma unroll
            for (int j = 0; j < NC; j += 4)
                *reinterpret_cast<uint2*>(o + j) = make_uint2(hv_pack_bf16(v[j], v[j + 1]), hv_pack_bf16(v[j + 2], v[j + 3]));
        } else {
#pragma unroll
            for (int j = 0; j < NC; ++j) if (col0 + j < N) o[j] = __float2bfloat16(v[j]);
        }
    }
}

// acc: NC consecutive columns [col0, col0+NC) of output row r (r < M guaranteed by the caller).
template <int NC>
__device__ __forceinline__ void hv_epilogue_row(const HvEpiDev& d, long long r, int col0, int N, float (&acc)[NC]) {
    const HvEpilogue& e = d.e;
    const bool full = d.vec_ok && (col0 + NC <= N);
    if (e.bias) {
        if (full) {
#pragma unroll
            for (int j = 0; j < NC; j += 4) {
                float4 b = *reinterpret_cast<const float4*>(e.bias + col0 + j);
                acc[j] += b.x; acc[j + 1] += b.y; acc[j + 2] += b.z; acc[j + 3] += b.w;
            }
        } else {
#pragma unroll
            for (int j = 0; j < NC; ++j) if (col0 + j < N) acc[j] += e.bias[col0 + j];
        }
    }
    if (e.out2 && e.mode != HV_EPI_RESIDUAL) hv_epi_store_dense<NC>(e, d.vec_ok2, r, col0, N, acc, e.out2, e.ldo2);
    if (e.mode == HV_EPI_STORE) {
        hv_epi_store_dense<NC>(e, d.vec_ok, r, col0, N, acc);
    } else if (e.mode == HV_EPI_GELU) {
#pragma unroll
        for (int j = 0; j < NC; ++j) acc[j] = hv_gelu_tanh(acc[j]);
        hv_epi_store_dense<NC>(e, d.vec_ok, r, col0, N, acc);
    } else if (e.mode == HV_EPI_RESIDUAL) {
        HvRowMap map{e.L, (long long)e.frame_rows, e.row_off};
        int frame, abs_i;
        const long long rr = map.res_row(r, frame, abs_i);
        float* o = e.res + rr * N + col0;
        const float* g = nullptr;
        if (e.gate) {
            const bool img = (abs_i >= e.img_start) && (abs_i < e.img_start + e.img_len);
            g = e.gate + (long long)((frame / e.cond_div) * e.cond_mul) * e.gate_ld + (img ? e.gate_col_img : e.gate_col_mot) + col0;
        }
        if (full) {
#pragma unroll
            for (int j = 0; j < NC; j += 4) {
                float4 x = *reinterpret_cast<float4*>(o + j);
                float4 gg = g ? *reinterpret_cast<const float4*>(g + j) : make_float4(1.f, 1.f, 1.f, 1.f);
                x.x = fmaf(gg.x, acc[j], x.x); x.y = fmaf(gg.y, acc[j + 1], x.y);
                x.z = fmaf(gg.z, acc[j + 2], x.z); x.w = fmaf(gg.w, acc[j + 3], x.w);
                *reinterpret_cast<float4*>(o + j) = x;
            }
        } else {
#pragma unroll
            for (int j = 0; j < NC; ++j) if (col0 + j < N) o[j] = fmaf(g ? g[j] : 1.f, acc[j], o[j]);
        }
    } else {  // HV_EPI_QKNORM — only reached with NC == 64 and col0 % 64 == 0 (tcgen05 path)
        if constexpr (NC == 64) {
            const float *w = nullptr, *b = nullptr;
            if (col0 < e.q_cols) { w = e.qn_w; b = e.qn_b; }
            else if (col0 < e.q_cols + e.k_cols) { w = e.kn_w; b = e.kn_b; }
            if (w) {
                float s = 0.f;
#pragma unroll
                for (int j = 0; j < NC; ++j) s += acc[j];
                const float mean = s * (1.f / 64.f);
                float q = 0.f;
#pragma unroll
                for (int j = 0; j < NC; ++j) { float t = acc[j] - mean; q = fmaf(t, t, q); }
                const float rstd = rsqrtf(q * (1.f / 64.f) + e.qk_eps);
#pragma unroll
                for (int j = 0; j < NC; ++j) acc[j] = (acc[j] - mean) * rstd * __ldg(w + j) + __ldg(b + j);
            }
        }
        hv_epi_store_dense<NC>(e, d.vec_ok, r, col0, N, acc);
    }
}
