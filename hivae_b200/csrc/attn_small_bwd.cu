// Attention backward for tiny sequences (Lq, Lkv <= 16, head_dim 64, bf16, no mask): the per-pixel temporal attention of
// the DiT's BasicDiTBlock ('(n t) s d -> (n s) t d', model/transformer.py:2022-2028) and the camera encoder's 16 x 16
// cross-attention (transformer.py:538-543) — autograd of F.scaled_dot_product_attention at model/modules.py:1138-1146, 794-803.
//
// HBM-bound like its forward (attn_small.cu): q, k, v, out, dout are read once, dq, dk, dv written once (16 B per token and
// head column chunk).  One warp owns one (sequence, head); the four 16 x 64 tiles are brought in with 16-byte cp.async, all
// five small products (S, dP, dQ, dK, dV: 16 x 16 x 64 each) run in fp32 on the FMA pipe — 80 K FMA per problem, far below
// what the memory traffic allows — with P and dS exchanged through 2 KB of shared memory.  No atomics, fixed order.
#include "attn.cuh"

namespace {

__device__ __forceinline__ uint32_t smem_addr_b(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void cp_async16_b(void* dst, const void* src, bool valid) {
    const int sz = valid ? 16 : 0;  // src-size 0 -> 16 bytes of zeros
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_addr_b(dst)), "l"(src), "r"(sz) : "memory");
}
// 16 rows x 64 bf16; 16-byte chunk c of row r at chunk c ^ (r & 7)
__device__ __forceinline__ const uint4* chunk(const __nv_bfloat16* tile, int r, int c) {
    return reinterpret_cast<const uint4*>(tile + r * 64 + ((c ^ (r & 7)) << 3));
}
__device__ __forceinline__ void unpack8(const uint4 v, float (&f)[8]) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) { f[2 * j] = __uint_as_float(w[j] << 16); f[2 * j + 1] = __uint_as_float(w[j] & 0xffff0000u); }
}
__device__ __forceinline__ float dot8(const float (&a)[8], const float (&b)[8]) {
    return ((a[0] * b[0] + a[1] * b[1]) + (a[2] * b[2] + a[3] * b[3])) + ((a[4] * b[4] + a[5] * b[5]) + (a[6] * b[6] + a[7] * b[7]));
}

struct SmallBwd {
    const __nv_bfloat16 *q, *k, *v, *o, *dout;
    __nv_bfloat16 *dq, *dk, *dv;
    long long ldq, ldk, ldv, ldo, lddo, lddq, lddk, lddv;
    int heads, Lq, Lkv;
    HvSeqMap qmap, kvmap;
    float scale;
};

constexpr int WARPS_B = 8;
constexpr int WARP_SMEM = 4 * 16 * 64 * 2 + 2 * 16 * 17 * 4;  // four bf16 tiles + P, dS in fp32

__global__ void __launch_bounds__(WARPS_B * 32) attn_small_bwd_kernel(SmallBwd a) {
    extern __shared__ __align__(128) uint8_t smem_b[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long seq = blockIdx.x;
    const int head = blockIdx.y * WARPS_B + warp;
    if (head >= a.heads) return;  // whole warp; only warp-level synchronisation below
    __nv_bfloat16* Qs = reinterpret_cast<__nv_bfloat16*>(smem_b + warp * WARP_SMEM);
    __nv_bfloat16 *Ks = Qs + 1024, *Vs = Qs + 2048, *dOs = Qs + 3072;
    float* Pm = reinterpret_cast<float*>(Qs + 4096);
    float* dSm = Pm + 16 * 17;
    const int hc = head * 64;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int idx = lane + 32 * i, r = idx >> 3, c = idx & 7;
        const bool okq = r < a.Lq, okk = r < a.Lkv;
        const long long rq = okq ? seq_row(a.qmap, seq, r) : 0, rk = okk ? seq_row(a.kvmap, seq, r) : 0;
        cp_async16_b(Qs + r * 64 + ((c ^ (r & 7)) << 3), a.q + rq * a.ldq + hc + c * 8, okq);
        cp_async16_b(dOs + r * 64 + ((c ^ (r & 7)) << 3), a.dout + rq * a.lddo + hc + c * 8, okq);
        cp_async16_b(Ks + r * 64 + ((c ^ (r & 7)) << 3), a.k + rk * a.ldk + hc + c * 8, okk);
        cp_async16_b(Vs + r * 64 + ((c ^ (r & 7)) << 3), a.v + rk * a.ldv + hc + c * 8, okk);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    const int i = lane >> 1, hf = lane & 1, jb = hf * 8;  // this lane: query row i, keys [jb, jb + 8) / head columns [32 hf, 32 hf + 32)
    // D_i = sum_d dO O: the O half-row straight from global while the tiles are in flight
    float ov[4][8];
    {
        const bool ok = i < a.Lq;
        const long long rq = ok ? seq_row(a.qmap, seq, i) : 0;
        const uint4* po = reinterpret_cast<const uint4*>(a.o + rq * a.ldo + hc + hf * 32);
#pragma unroll
        for (int c = 0; c < 4; ++c) unpack8(ok ? __ldg(po + c) : make_uint4(0u, 0u, 0u, 0u), ov[c]);
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncwarp();
    float D = 0.f;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        float d[8];
        unpack8(*chunk(dOs, i, hf * 4 + c), d);
        D += dot8(d, ov[c]);
    }
    D += __shfl_xor_sync(0xffffffffu, D, 1);
    // ---- S = Q K^T and dP = dO V^T for (row i, keys jb .. jb+7)
    float s[8], dp[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) s[j] = dp[j] = 0.f;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        float qv[8], dv_[8];
        unpack8(*chunk(Qs, i, c), qv);
        unpack8(*chunk(dOs, i, c), dv_);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float kv[8], vv[8];
            unpack8(*chunk(Ks, jb + j, c), kv);
            unpack8(*chunk(Vs, jb + j, c), vv);
            s[j] += dot8(qv, kv);
            dp[j] += dot8(dv_, vv);
        }
    }
    // ---- softmax over the 16 keys of row i (two lanes), dS = P (dP - D) scale
    const float sl2 = a.scale * 1.4426950408889634f;
    float mx = -INFINITY;
#pragma unroll
    for (int j = 0; j < 8; ++j) { s[j] = (jb + j < a.Lkv) ? s[j] * sl2 : -INFINITY; mx = fmaxf(mx, s[j]); }
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
    float sum = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) { s[j] = exp2f(s[j] - mx); sum += s[j]; }
    sum += __shfl_xor_sync(0xffffffffu, sum, 1);
    const float inv = 1.f / sum;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const float p = s[j] * inv;
        Pm[i * 17 + jb + j] = p;
        dSm[i * 17 + jb + j] = p * (dp[j] - D) * a.scale;
    }
    __syncwarp();
    // ---- dQ[i][32 hf ..] = sum_j dS[i][j] K[j][..]
    {
        float acc[4][8];
#pragma unroll
        for (int c = 0; c < 4; ++c)
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[c][e] = 0.f;
#pragma unroll 4
        for (int j = 0; j < 16; ++j) {
            const float ds = dSm[i * 17 + j];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                float kv[8];
                unpack8(*chunk(Ks, j, hf * 4 + c), kv);
#pragma unroll
                for (int e = 0; e < 8; ++e) acc[c][e] = fmaf(ds, kv[e], acc[c][e]);
            }
        }
        if (i < a.Lq) {
            uint4* dst = reinterpret_cast<uint4*>(a.dq + seq_row(a.qmap, seq, i) * a.lddq + hc + hf * 32);
#pragma unroll
            for (int c = 0; c < 4; ++c)
                dst[c] = make_uint4(hv_pack_bf16(acc[c][0], acc[c][1]), hv_pack_bf16(acc[c][2], acc[c][3]),
                                    hv_pack_bf16(acc[c][4], acc[c][5]), hv_pack_bf16(acc[c][6], acc[c][7]));
        }
    }
    // ---- dV[j][32 hf ..] = sum_i P[i][j] dO[i][..],  dK[j][32 hf ..] = sum_i dS[i][j] Q[i][..]   (this lane: key row j = lane / 2)
    {
        const int j = i;
        float av[4][8], ak[4][8];
#pragma unroll
        for (int c = 0; c < 4; ++c)
#pragma unroll
            for (int e = 0; e < 8; ++e) av[c][e] = ak[c][e] = 0.f;
#pragma unroll 4
        for (int r = 0; r < 16; ++r) {
            const float p = Pm[r * 17 + j], ds = dSm[r * 17 + j];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                float dv_[8], qv[8];
                unpack8(*chunk(dOs, r, hf * 4 + c), dv_);
                unpack8(*chunk(Qs, r, hf * 4 + c), qv);
#pragma unroll
                for (int e = 0; e < 8; ++e) { av[c][e] = fmaf(p, dv_[e], av[c][e]); ak[c][e] = fmaf(ds, qv[e], ak[c][e]); }
            }
        }
        if (j < a.Lkv) {
            const long long rk = seq_row(a.kvmap, seq, j);
            uint4* pv = reinterpret_cast<uint4*>(a.dv + rk * a.lddv + hc + hf * 32);
            uint4* pk = reinterpret_cast<uint4*>(a.dk + rk * a.lddk + hc + hf * 32);
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                pv[c] = make_uint4(hv_pack_bf16(av[c][0], av[c][1]), hv_pack_bf16(av[c][2], av[c][3]),
                                   hv_pack_bf16(av[c][4], av[c][5]), hv_pack_bf16(av[c][6], av[c][7]));
                pk[c] = make_uint4(hv_pack_bf16(ak[c][0], ak[c][1]), hv_pack_bf16(ak[c][2], ak[c][3]),
                                   hv_pack_bf16(ak[c][4], ak[c][5]), hv_pack_bf16(ak[c][6], ak[c][7]));
            }
        }
    }
}

}  // namespace

// 1 = launched, 0 = shape not handled here, < 0 = error
int hv_attention_small_bwd_bf16(const void* q, long long ldq, const void* k, long long ldk, const void* v, long long ldv,
                                const void* o, long long ldo, const void* dout, long long lddo, void* dq, long long lddq,
                                void* dk, long long lddk, void* dv, long long lddv, long long seqs, int heads, int Lq, int Lkv,
                                const HvSeqMap& qm, const HvSeqMap& km, float scale, cudaStream_t st) {
    if (Lq > 16 || Lkv > 16 || Lq < 1 || Lkv < 1 || seqs > 2147483647ll) return 0;
    if ((ldq % 8) || (ldk % 8) || (ldv % 8) || (ldo % 8) || (lddo % 8) || (lddq % 8) || (lddk % 8) || (lddv % 8)) return 0;
    auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
    if (!al16(q) || !al16(k) || !al16(v) || !al16(o) || !al16(dout) || !al16(dq) || !al16(dk) || !al16(dv)) return 0;
    SmallBwd a;
    a.q = (const __nv_bfloat16*)q; a.k = (const __nv_bfloat16*)k; a.v = (const __nv_bfloat16*)v; a.o = (const __nv_bfloat16*)o;
    a.dout = (const __nv_bfloat16*)dout; a.dq = (__nv_bfloat16*)dq; a.dk = (__nv_bfloat16*)dk; a.dv = (__nv_bfloat16*)dv;
    a.ldq = ldq; a.ldk = ldk; a.ldv = ldv; a.ldo = ldo; a.lddo = lddo; a.lddq = lddq; a.lddk = lddk; a.lddv = lddv;
    a.heads = heads; a.Lq = Lq; a.Lkv = Lkv; a.qmap = qm; a.kvmap = km; a.scale = scale;
    const int smem = WARPS_B * WARP_SMEM;
    static bool configured_of[HV_MAX_DEVICES] = {};  // function attributes are per device
    bool& configured = configured_of[hv_device_slot()];
    if (!configured) { HV_CUDA(cudaFuncSetAttribute(attn_small_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); configured = true; }
    dim3 grid((unsigned)seqs, (unsigned)((heads + WARPS_B - 1) / WARPS_B));
    attn_small_bwd_kernel<<<grid, WARPS_B * 32, smem, st>>>(a);
    HV_LAUNCH_CHECK();
    return 1;
}
