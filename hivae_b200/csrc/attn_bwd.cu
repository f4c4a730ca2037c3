// Attention backward, first (SIMT fp32-math) version: gradients of softmax(Q K^T * scale) V per (sequence, head)
// for head_dim 64, no mask — the autograd of F.scaled_dot_product_attention inside diffusers AttnProcessor2_0 at the
// call sites model/modules.py:699-707, 1043-1051.  Two kernels, each recomputing the
// probabilities it needs (flash-attention style, nothing of size Lq x Lkv ever goes to HBM):
//   dq kernel : one CTA per (16 query rows, head, sequence); K and V of the sequence in shared memory.
//               Per row: s = q K^T, lse, p = exp(s*scale - lse), dp = dO V^T, delta = dO . O,
//               ds = p (dp - delta) scale, dQ = ds K.  Also writes (lse, delta) per row for the second kernel.
//   dkv kernel: one CTA per (16 keys, head, sequence); Q and dO of the sequence in shared memory.
//               Per key: p, ds as above (from lse / delta), dV = p^T dO, dK = ds^T Q.
// Lanes own keys (or queries) for the 64-long dot products — rows are padded to 65 floats, so lanes reading different
// rows at the same column hit different banks — and then own columns for the rank-1 accumulations.
// This is the correctness baseline of the backward path (about 450 FMA per (query, key) pair on the FMA pipes); the
// tcgen05 version is next round's work.
#include "attn.cuh"

namespace {

constexpr int HDB = 64, PITCH = 65, ROWS_PER_CTA = 16;

template <typename T> __device__ __forceinline__ float ld1(const T* p, long long i);
template <> __device__ __forceinline__ float ld1<float>(const float* p, long long i) { return p[i]; }
template <> __device__ __forceinline__ float ld1<__nv_bfloat16>(const __nv_bfloat16* p, long long i) { return __bfloat162float(p[i]); }
template <typename T> __device__ __forceinline__ void st1(T* p, long long i, float v);
template <> __device__ __forceinline__ void st1<float>(float* p, long long i, float v) { p[i] = v; }
template <> __device__ __forceinline__ void st1<__nv_bfloat16>(__nv_bfloat16* p, long long i, float v) { p[i] = __float2bfloat16_rn(v); }

struct BwdArgs {
    const void *q, *k, *v, *o, *dout;
    void *dq, *dk, *dv;
    long long ldq, ldk, ldv, ldo, lddo, lddq, lddk, lddv;
    int heads, Lq, Lkv;
    HvSeqMap qmap, kvmap;
    float scale;
    float* stats;  // [seqs][heads][Lq][2] = (lse in the scaled-logit domain, delta)
};

// copies rows [0, L) x 64 columns of one head into shared memory (fp32, pitch 65)
template <typename T>
__device__ __forceinline__ void load_head(float* dst, const T* src, long long ld, const HvSeqMap& m, long long seq, int L, int col0) {
    for (int i = threadIdx.x; i < L * HDB; i += blockDim.x) {
        const int r = i >> 6, c = i & 63;
        dst[r * PITCH + c] = ld1<T>(src, seq_row(m, seq, r) * ld + col0 + c);
    }
}

template <typename T>
__global__ void __launch_bounds__(256) attn_bwd_dq_kernel(const BwdArgs a) {
    extern __shared__ float sm[];
    float* Ks = sm;                          // Lkv x 65
    float* Vs = Ks + a.Lkv * PITCH;          // Lkv x 65
    float* wq = Vs + a.Lkv * PITCH;          // 8 warps x 64   (query row)
    float* wdo = wq + 8 * HDB;               // 8 warps x 64   (dO row)
    float* wds = wdo + 8 * HDB;              // 8 warps x Lkv  (ds row)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long seq = blockIdx.z;
    const int head = blockIdx.y, col0 = head * HDB;
    load_head<T>(Ks, (const T*)a.k, a.ldk, a.kvmap, seq, a.Lkv, col0);
    load_head<T>(Vs, (const T*)a.v, a.ldv, a.kvmap, seq, a.Lkv, col0);
    __syncthreads();
    float* myq = wq + warp * HDB;
    float* mydo = wdo + warp * HDB;
    float* myds = wds + warp * a.Lkv;
    for (int rr = 0; rr < ROWS_PER_CTA / 8; ++rr) {
        const int i = blockIdx.x * ROWS_PER_CTA + rr * 8 + warp;
        if (i >= a.Lq) break;  // warp-uniform
        const long long row = seq_row(a.qmap, seq, i);
        float dpart = 0.f;
        for (int c = lane; c < HDB; c += 32) {
            myq[c] = ld1<T>((const T*)a.q, row * a.ldq + col0 + c);
            const float d = ld1<T>((const T*)a.dout, row * a.lddo + col0 + c);
            mydo[c] = d;
            dpart += d * ld1<T>((const T*)a.o, row * a.ldo + col0 + c);
        }
        const float delta = hv_warp_sum(dpart);
        __syncwarp();
        // lanes own keys j = lane, lane + 32, ...
        float mx = -INFINITY;
        for (int j = lane; j < a.Lkv; j += 32) {
            const float* kr = Ks + j * PITCH;
            float s = 0.f;
#pragma unroll 16
            for (int c = 0; c < HDB; ++c) s = fmaf(myq[c], kr[c], s);
            s *= a.scale;
            myds[j] = s;
            mx = fmaxf(mx, s);
        }
        mx = hv_warp_max(mx);
        float sum = 0.f;
        for (int j = lane; j < a.Lkv; j += 32) sum += __expf(myds[j] - mx);
        sum = hv_warp_sum(sum);
        const float lse = mx + __logf(sum);
        for (int j = lane; j < a.Lkv; j += 32) {
            const float* vr = Vs + j * PITCH;
            float dp = 0.f;
#pragma unroll 16
            for (int c = 0; c < HDB; ++c) dp = fmaf(mydo[c], vr[c], dp);
            const float p = __expf(myds[j] - lse);
            myds[j] = p * (dp - delta) * a.scale;
        }
        __syncwarp();
        // lanes own columns c = lane, lane + 32
        float g0 = 0.f, g1 = 0.f;
        for (int j = 0; j < a.Lkv; ++j) {
            const float ds = myds[j];
            g0 = fmaf(ds, Ks[j * PITCH + lane], g0);
            g1 = fmaf(ds, Ks[j * PITCH + lane + 32], g1);
        }
        st1<T>((T*)a.dq, row * a.lddq + col0 + lane, g0);
        st1<T>((T*)a.dq, row * a.lddq + col0 + lane + 32, g1);
        if (lane == 0) {
            float* st = a.stats + ((seq * a.heads + head) * a.Lq + i) * 2;
            st[0] = lse; st[1] = delta;
        }
        __syncwarp();
    }
}

template <typename T>
__global__ void __launch_bounds__(256) attn_bwd_dkv_kernel(const BwdArgs a) {
    extern __shared__ float sm[];
    float* Qs = sm;                          // Lq x 65
    float* Ds = Qs + a.Lq * PITCH;           // Lq x 65   (dO)
    float* slse = Ds + a.Lq * PITCH;         // Lq
    float* sdel = slse + a.Lq;               // Lq
    float* wk = sdel + a.Lq;                 // 8 warps x 64  (key row)
    float* wv = wk + 8 * HDB;                // 8 warps x 64  (value row)
    float* wp = wv + 8 * HDB;                // 8 warps x Lq  (p column)
    float* wds = wp + 8 * a.Lq;              // 8 warps x Lq  (ds column)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long seq = blockIdx.z;
    const int head = blockIdx.y, col0 = head * HDB;
    load_head<T>(Qs, (const T*)a.q, a.ldq, a.qmap, seq, a.Lq, col0);
    load_head<T>(Ds, (const T*)a.dout, a.lddo, a.qmap, seq, a.Lq, col0);
    const float* st = a.stats + (seq * a.heads + head) * a.Lq * 2;
    for (int i = threadIdx.x; i < a.Lq; i += blockDim.x) { slse[i] = st[2 * i]; sdel[i] = st[2 * i + 1]; }
    __syncthreads();
    float* myk = wk + warp * HDB;
    float* myv = wv + warp * HDB;
    float* myp = wp + warp * a.Lq;
    float* myds = wds + warp * a.Lq;
    for (int rr = 0; rr < ROWS_PER_CTA / 8; ++rr) {
        const int j = blockIdx.x * ROWS_PER_CTA + rr * 8 + warp;
        if (j >= a.Lkv) break;  // warp-uniform
        const long long row = seq_row(a.kvmap, seq, j);
        for (int c = lane; c < HDB; c += 32) {
            myk[c] = ld1<T>((const T*)a.k, row * a.ldk + col0 + c);
            myv[c] = ld1<T>((const T*)a.v, row * a.ldv + col0 + c);
        }
        __syncwarp();
        // lanes own queries i = lane, lane + 32, ...
        for (int i = lane; i < a.Lq; i += 32) {
            const float* qr = Qs + i * PITCH;
            const float* dr = Ds + i * PITCH;
            float s = 0.f, dp = 0.f;
#pragma unroll 16
            for (int c = 0; c < HDB; ++c) { s = fmaf(qr[c], myk[c], s); dp = fmaf(dr[c], myv[c], dp); }
            const float p = __expf(s * a.scale - slse[i]);
            myp[i] = p;
            myds[i] = p * (dp - sdel[i]) * a.scale;
        }
        __syncwarp();
        // lanes own columns c = lane, lane + 32
        float v0 = 0.f, v1 = 0.f, k0 = 0.f, k1 = 0.f;
        for (int i = 0; i < a.Lq; ++i) {
            const float p = myp[i], ds = myds[i];
            v0 = fmaf(p, Ds[i * PITCH + lane], v0); v1 = fmaf(p, Ds[i * PITCH + lane + 32], v1);
            k0 = fmaf(ds, Qs[i * PITCH + lane], k0); k1 = fmaf(ds, Qs[i * PITCH + lane + 32], k1);
        }
        st1<T>((T*)a.dv, row * a.lddv + col0 + lane, v0); st1<T>((T*)a.dv, row * a.lddv + col0 + lane + 32, v1);
        st1<T>((T*)a.dk, row * a.lddk + col0 + lane, k0); st1<T>((T*)a.dk, row * a.lddk + col0 + lane + 32, k1);
        __syncwarp();
    }
}

template <typename T>
int launch_bwd(const BwdArgs& a, int64_t seqs, cudaStream_t st) {
    const size_t sm1 = (size_t)(2 * a.Lkv * PITCH + 16 * HDB + 8 * a.Lkv) * sizeof(float);
    const size_t sm2 = (size_t)(2 * a.Lq * PITCH + 2 * a.Lq + 16 * HDB + 16 * a.Lq) * sizeof(float);
    HV_REQUIRE(sm1 <= 227 * 1024 && sm2 <= 227 * 1024, "hv_attention_bwd: Lq=%d / Lkv=%d need more than 227 KB of shared memory", a.Lq, a.Lkv);
    HV_CUDA(cudaFuncSetAttribute(attn_bwd_dq_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm1));
    HV_CUDA(cudaFuncSetAttribute(attn_bwd_dkv_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm2));
    dim3 g1((a.Lq + ROWS_PER_CTA - 1) / ROWS_PER_CTA, a.heads, (unsigned)seqs);
    attn_bwd_dq_kernel<T><<<g1, 256, sm1, st>>>(a);
    HV_LAUNCH_CHECK();
    dim3 g2((a.Lkv + ROWS_PER_CTA - 1) / ROWS_PER_CTA, a.heads, (unsigned)seqs);
    attn_bwd_dkv_kernel<T><<<g2, 256, sm2, st>>>(a);
    HV_LAUNCH_CHECK();
    return HV_OK;
}

}  // namespace

// Gradients of out = softmax(q k^T * scale) v for head_dim 64 (no mask).  All matrices are (rows, heads*64) with their own
// leading dimension, one dtype (HV_F32 / HV_BF16); `stats` = seqs * heads * Lq * 2 floats of workspace.
extern "C" int hv_attention_bwd(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                                const void* out, int64_t ldo, const void* dout, int64_t lddo,
                                void* dq, int64_t lddq, void* dk, int64_t lddk, void* dv, int64_t lddv, int dtype,
                                int64_t seqs, int heads, int hd, int Lq, int Lkv, const HvSeqMap* qmap, const HvSeqMap* kvmap,
                                float scale, float* stats, void* stream) {
    HV_REQUIRE(hd == HDB, "hv_attention_bwd: head_dim %d != 64", hd);
    HV_REQUIRE(seqs >= 0 && heads > 0 && heads <= 65535 && seqs <= 65535 && Lq > 0 && Lkv > 0, "hv_attention_bwd: bad sizes");
    HV_REQUIRE(qmap && kvmap && qmap->sdiv > 0 && kvmap->sdiv > 0 && stats, "hv_attention_bwd: bad sequence maps / workspace");
    HV_REQUIRE(dtype == HV_F32 || dtype == HV_BF16, "hv_attention_bwd: bad dtype");
    if (seqs == 0) return HV_OK;
    BwdArgs a{q, k, v, out, dout, dq, dk, dv, ldq, ldk, ldv, ldo, lddo, lddq, lddk, lddv, heads, Lq, Lkv, *qmap, *kvmap, scale, stats};
    return dtype == HV_F32 ? launch_bwd<float>(a, seqs, hv_stream(stream)) : launch_bwd<__nv_bfloat16>(a, seqs, hv_stream(stream));
}
