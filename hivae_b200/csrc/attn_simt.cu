// Generic fp32-math attention (any Lq/Lkv, head_dim 32/64, optional 0/1 mask): the fp32
// parity path, the tiny temporal (L=16) and camera cross-attention shapes, and model/attention.py.
// One CTA = 16 query rows of one (sequence, head); K/V streamed through shared memory in chunks of
// 64 keys; online softmax per query row kept in registers of the owning warp.
#include "attn.cuh"

template <typename T, int HD>
__global__ void __launch_bounds__(128) attn_simt_kernel(AttnArgs a) {
    constexpr int QT = 16, KT = 64, RPW = 4, DPL = HD / 32;  // rows per warp, dims per lane
    __shared__ float Ks[KT][HD + 1];
    __shared__ float Vs[KT][HD];
    __shared__ float Qs[QT][HD];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int qtiles = (a.Lq + QT - 1) / QT;
    const long long seq = blockIdx.x / qtiles;
    const int head = blockIdx.y, q0 = (int)(blockIdx.x % qtiles) * QT;
    const T* qp = reinterpret_cast<const T*>(a.q) + (long long)head * HD;
    const T* kp = reinterpret_cast<const T*>(a.k) + (long long)head * HD;
    const T* vp = reinterpret_cast<const T*>(a.v) + (long long)head * HD;
    for (int idx = threadIdx.x; idx < QT * HD; idx += blockDim.x) {
        const int r = idx / HD, d = idx - r * HD;
        Qs[r][d] = (q0 + r < a.Lq) ? hv_ld(qp + seq_row(a.qmap, seq, q0 + r) * a.ldq + d) * a.scale : 0.f;
    }
    float m[RPW], l[RPW], o[RPW][DPL];
#pragma unroll
    for (int i = 0; i < RPW; ++i) {
        m[i] = -INFINITY; l[i] = 0.f;
#pragma unroll
        for (int d = 0; d < DPL; ++d) o[i][d] = 0.f;
    }
    for (int k0 = 0; k0 < a.Lkv; k0 += KT) {
        __syncthreads();
        for (int idx = threadIdx.x; idx < KT * HD; idx += blockDim.x) {
            const int r = idx / HD, d = idx - r * HD;
            float kv = 0.f, vv = 0.f;
            if (k0 + r < a.Lkv) {
                const long long row = seq_row(a.kvmap, seq, k0 + r);
                kv = hv_ld(kp + row * a.ldk + d);
                vv = hv_ld(vp + row * a.ldv + d);
            }
            Ks[r][d] = kv; Vs[r][d] = vv;
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < RPW; ++i) {
            const int qr = warp * RPW + i;
            if (q0 + qr >= a.Lq) continue;  // warp-uniform
            float s0 = 0.f, s1 = 0.f;       // keys k0+lane and k0+32+lane
#pragma unroll 8
            for (int d = 0; d < HD; ++d) {
                const float qd = Qs[qr][d];
                s0 = fmaf(qd, Ks[lane][d], s0);
                s1 = fmaf(qd, Ks[lane + 32][d], s1);
            }
            if (a.mask) {
                const float* mr = a.mask + (long long)(q0 + qr) * a.Lkv + k0;
                if (k0 + lane < a.Lkv && mr[lane] == 0.f) s0 = -1e9f;
                if (k0 + 32 + lane < a.Lkv && mr[32 + lane] == 0.f) s1 = -1e9f;
            }
            if (k0 + lane >= a.Lkv) s0 = -INFINITY;
            if (k0 + 32 + lane >= a.Lkv) s1 = -INFINITY;
            const float mn = fmaxf(m[i], hv_warp_max(fmaxf(s0, s1)));
            const float corr = __expf(m[i] - mn);  // exp(-inf) = 0 on the first chunk
            const float p0 = __expf(s0 - mn), p1 = __expf(s1 - mn);
            l[i] = l[i] * corr + hv_warp_sum(p0 + p1);
            m[i] = mn;
#pragma unroll
            for (int d = 0; d < DPL; ++d) o[i][d] *= corr;
#pragma unroll 8
            for (int j = 0; j < 32; ++j) {
                const float pj0 = __shfl_sync(0xffffffffu, p0, j), pj1 = __shfl_sync(0xffffffffu, p1, j);
#pragma unroll
                for (int d = 0; d < DPL; ++d) {
                    o[i][d] = fmaf(pj0, Vs[j][d * 32 + lane], o[i][d]);
                    o[i][d] = fmaf(pj1, Vs[j + 32][d * 32 + lane], o[i][d]);
                }
            }
        }
    }
    T* op = reinterpret_cast<T*>(a.out) + (long long)head * HD;
#pragma unroll
    for (int i = 0; i < RPW; ++i) {
        const int qr = q0 + warp * RPW + i;
        if (qr >= a.Lq) continue;
        const float inv = 1.f / l[i];
        const long long row = seq_row(a.qmap, seq, qr);
#pragma unroll
        for (int d = 0; d < DPL; ++d) hv_st(op + row * a.ldo + d * 32 + lane, o[i][d] * inv);
    }
}

template <typename T>
static int launch_simt(const AttnArgs& a, int64_t seqs, cudaStream_t st) {
    dim3 grid((unsigned)(((a.Lq + 15) / 16) * seqs), (unsigned)a.heads);
    switch (a.hd) {
        case 32: attn_simt_kernel<T, 32><<<grid, 128, 0, st>>>(a); break;
        case 64: attn_simt_kernel<T, 64><<<grid, 128, 0, st>>>(a); break;
        default: hv_set_error("hv_attention: head_dim %d unsupported (32 or 64)", a.hd); return HV_EINVAL;
    }
    HV_LAUNCH_CHECK();
    return HV_OK;
}

int hv_attention_flash_bf16(const AttnArgs& a, int64_t seqs, cudaStream_t st);  // attn_flash.cu
int hv_attention_tc_bf16(const AttnArgs& a, int64_t seqs, cudaStream_t st);     // attn_tc.cu (1 = launched, 0 = not applicable)
int hv_attention_tc2_bf16(const AttnArgs& a, int64_t seqs, cudaStream_t st);    // attn_tc2.cu (same convention; tried first)
int hv_attention_small_bf16(const AttnArgs& a, int64_t seqs, cudaStream_t st);  // attn_small.cu (same convention)

// shared body of hv_attention / hv_attention_lse; *wrote_lse <- 1 when the kernel that ran filled a.lse_stats
static int attention_dispatch(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                              void* out, int64_t ldo, int dtype, int64_t seqs, int heads, int hd, int Lq, int Lkv,
                              const HvSeqMap* qmap, const HvSeqMap* kvmap, float scale, const float* mask, float* stats,
                              int* wrote_lse, void* stream) {
    HV_REQUIRE(seqs >= 0 && heads > 0 && Lq >= 0 && Lkv > 0, "hv_attention: bad sizes");
    HV_REQUIRE(qmap && kvmap && qmap->sdiv > 0 && kvmap->sdiv > 0, "hv_attention: bad sequence maps");
    HV_REQUIRE(dtype == HV_F32 || dtype == HV_BF16, "hv_attention: bad dtype");
    HV_REQUIRE(seqs * ((Lq + 15) / 16) < (1ll << 31) && heads <= 65535, "hv_attention: grid too large");
    *wrote_lse = 0;
    if (seqs == 0 || Lq == 0) return HV_OK;
    AttnArgs a{q, k, v, out, ldq, ldk, ldv, ldo, heads, hd, Lq, Lkv, *qmap, *kvmap, scale, mask};
    cudaStream_t st = hv_stream(stream);
    if (dtype == HV_BF16) {
        int rc = 0;
        if (Lq <= 16 && Lkv <= 16) rc = hv_attention_small_bf16(a, seqs, st);
        else if (Lkv >= 48) {
            a.lse_stats = stats;
            rc = hv_attention_tc2_bf16(a, seqs, st);
            if (rc == 1 && stats) *wrote_lse = 1;
            a.lse_stats = nullptr;
        }
        if (rc == 0 && Lkv >= 48) rc = hv_attention_tc_bf16(a, seqs, st);
        if (rc != 0) return rc < 0 ? rc : HV_OK;
    }
    if (dtype == HV_BF16 && hd == 64 && mask == nullptr && Lkv >= 48 &&
        (ldq % 8 == 0) && (ldk % 8 == 0) && (ldv % 8 == 0) && (ldo % 2 == 0))
        return hv_attention_flash_bf16(a, seqs, st);
    return dtype == HV_F32 ? launch_simt<float>(a, seqs, st) : launch_simt<__nv_bfloat16>(a, seqs, st);
}

extern "C" int hv_attention(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                            void* out, int64_t ldo, int dtype, int64_t seqs, int heads, int hd, int Lq, int Lkv,
                            const HvSeqMap* qmap, const HvSeqMap* kvmap, float scale, const float* mask, void* stream) {
    int wrote = 0;
    return attention_dispatch(q, ldq, k, ldk, v, ldv, out, ldo, dtype, seqs, heads, hd, Lq, Lkv, qmap, kvmap, scale, mask, nullptr, &wrote, stream);
}

// hv_attention that also hands the per-row log-sum-exp to the backward: `stats` is the workspace hv_attention_bwd takes
// (seqs * heads * Lq * 2 floats); returns 1 when element 0 of every row was written (then call hv_attention_bwd_lse with
// have_lse = 1 and the backward skips its own maximum / sum passes), 0 when the kernel that ran does not produce it, < 0 on error
extern "C" int hv_attention_lse(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                                void* out, int64_t ldo, int dtype, int64_t seqs, int heads, int hd, int Lq, int Lkv,
                                const HvSeqMap* qmap, const HvSeqMap* kvmap, float scale, const float* mask, float* stats, void* stream) {
    int wrote = 0;
    const int rc = attention_dispatch(q, ldq, k, ldk, v, ldv, out, ldo, dtype, seqs, heads, hd, Lq, Lkv, qmap, kvmap, scale, mask, stats, &wrote, stream);
    return rc != HV_OK ? rc : wrote;
}
