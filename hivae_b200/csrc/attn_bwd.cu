// Attention backward: gradients of softmax(Q K^T * scale) V per (sequence, head)
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
// The SIMT fp32-math kernels below serve fp32 tensors (1e-4-level parity) and odd alignments (about 450 FMA per
// (query, key) pair on the FMA pipes); bf16 tensors take the warp-level tensor-core kernels further down.
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


// ------------------------------------------------------------------------------------------------------------------
// bf16 tensor-core version (warp-level mma.sync.m16n8k16, fp32 accumulation), same two-kernel structure and the same
// (lse, delta) hand-over.  A CTA is four warps; each warp owns 16 query rows (dq kernel) or 16 keys (dkv kernel), holds
// their Q / dO (K / V) rows as A fragments in registers and walks the other dimension in chunks of 32, whose rows sit
// in shared memory as bf16 with a pitch of 72 elements (fragment loads and ldmatrix rows hit 32 distinct banks):
//   dq : S = Q K^T and dP = dO V^T (B fragments = plain 32-bit loads of K / V rows), P = exp(S scale - lse),
//        dS = P (dP - delta) scale re-packed from the accumulator layout into A fragments, dQ += dS K with the B
//        fragments of K taken through ldmatrix.trans.  lse comes from a first pass of S with an online max / sum.
//   dkv: S^T = K Q^T, dP^T = V dO^T, P^T / dS^T as above (lse / delta indexed by column), dV += P^T dO, dK += dS^T Q.
// Padded rows are zero in shared memory; padded keys are masked, padded queries carry lse = +inf (p = 0).
constexpr int MP = 72;

__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ldsm_x4_trans(uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3, const void* p) {
    const uint32_t addr = (uint32_t)__cvta_generic_to_shared(p);
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];\n"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float2 unpack_bf16(uint32_t u) {
    return __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u));
}

__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" :: "n"(N) : "memory"); }

// rows [0, L) x 64 columns of one head -> shared memory [Lp][MP] bf16, rows [L, Lp) zero
__device__ __forceinline__ void stage_head(__nv_bfloat16* dst, const __nv_bfloat16* src, long long ld, const HvSeqMap& m,
                                           long long seq, int L, int Lp, int col0, int vec) {
    if (vec) {  // 16-byte cp.async: every copy of the head is in flight at once (the caller commits / waits)
        for (int i = threadIdx.x; i < Lp * 8; i += blockDim.x) {
            const int r = i >> 3, c = (i & 7) * 8;
            if (r < L) {
                const uint32_t d = (uint32_t)__cvta_generic_to_shared(dst + r * MP + c);
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" :: "r"(d), "l"(src + seq_row(m, seq, r) * ld + col0 + c) : "memory");
            } else {
                *reinterpret_cast<uint4*>(dst + r * MP + c) = make_uint4(0u, 0u, 0u, 0u);
            }
        }
    } else {
        for (int i = threadIdx.x; i < Lp * 32; i += blockDim.x) {
            const int r = i >> 5, c = (i & 31) * 2;
            uint32_t v = 0u;
            if (r < L) v = *reinterpret_cast<const uint32_t*>(src + seq_row(m, seq, r) * ld + col0 + c);
            *reinterpret_cast<uint32_t*>(dst + r * MP + c) = v;
        }
    }
}

// A fragments (four 16-wide k tiles) of rows r0 .. r0+15 of one head, straight from global memory; rows >= L are zero
__device__ __forceinline__ void load_a_frags(uint32_t (&f)[4][4], const __nv_bfloat16* src, long long ld, const HvSeqMap& m,
                                             long long seq, int r0, int L, int col0, int g, int tig) {
    const __nv_bfloat16* p0 = (r0 + g < L) ? src + seq_row(m, seq, r0 + g) * ld + col0 + 2 * tig : nullptr;
    const __nv_bfloat16* p1 = (r0 + g + 8 < L) ? src + seq_row(m, seq, r0 + g + 8) * ld + col0 + 2 * tig : nullptr;
#pragma unroll
    for (int kt = 0; kt < 4; ++kt) {
        f[kt][0] = p0 ? *reinterpret_cast<const uint32_t*>(p0 + kt * 16) : 0u;
        f[kt][1] = p1 ? *reinterpret_cast<const uint32_t*>(p1 + kt * 16) : 0u;
        f[kt][2] = p0 ? *reinterpret_cast<const uint32_t*>(p0 + kt * 16 + 8) : 0u;
        f[kt][3] = p1 ? *reinterpret_cast<const uint32_t*>(p1 + kt * 16 + 8) : 0u;
    }
}

__device__ __forceinline__ void ldsm_x4(uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3, const void* p) {
    const uint32_t addr = (uint32_t)__cvta_generic_to_shared(p);
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];\n"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}

// B fragments of 32 rows of a [.][MP] shared array: one ldmatrix.x4 fetches two k tiles of an 8-row block; lane l addresses
// row (l & 7), columns 8 (l >> 3) of the 8 x 32 block.
// two independent products over the same 32 rows (S = A1 X1^T and dP = A2 X2^T) with their MMAs interleaved, so that
// consecutive tensor instructions never touch the same accumulator (dependent MMAs are eight instructions apart)
__device__ __forceinline__ void mma_rows32x2(float (&acc1)[4][4], const uint32_t (&af1)[4][4], const __nv_bfloat16* X1,
                                             float (&acc2)[4][4], const uint32_t (&af2)[4][4], const __nv_bfloat16* X2, int base, int lane) {
    const int off = (base + (lane & 7)) * MP + (lane >> 3) * 8;
#pragma unroll
    for (int kt = 0; kt < 4; kt += 2) {
        uint32_t b1[4][4], b2[4][4];
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) {
            ldsm_x4(b1[nt][0], b1[nt][1], b1[nt][2], b1[nt][3], X1 + off + nt * 8 * MP + kt * 16);
            ldsm_x4(b2[nt][0], b2[nt][1], b2[nt][2], b2[nt][3], X2 + off + nt * 8 * MP + kt * 16);
        }
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) mma16816(acc1[nt], af1[kt], b1[nt][0], b1[nt][1]);
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) mma16816(acc2[nt], af2[kt], b2[nt][0], b2[nt][1]);
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) mma16816(acc1[nt], af1[kt + 1], b1[nt][2], b1[nt][3]);
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) mma16816(acc2[nt], af2[kt + 1], b2[nt][2], b2[nt][3]);
    }
}

// one product with the K dimension split over two accumulator sets (same reason), summed at the end
__device__ __forceinline__ void mma_rows32_split(float (&acc)[4][4], const uint32_t (&af)[4][4], const __nv_bfloat16* X, int base, int lane) {
    const int off = (base + (lane & 7)) * MP + (lane >> 3) * 8;
    float hi[4][4] = {};
    uint32_t b1[4][4], b2[4][4];
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
        ldsm_x4(b1[nt][0], b1[nt][1], b1[nt][2], b1[nt][3], X + off + nt * 8 * MP);
        ldsm_x4(b2[nt][0], b2[nt][1], b2[nt][2], b2[nt][3], X + off + nt * 8 * MP + 32);
    }
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) mma16816(acc[nt], af[0], b1[nt][0], b1[nt][1]);
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) mma16816(hi[nt], af[2], b2[nt][0], b2[nt][1]);
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) mma16816(acc[nt], af[1], b1[nt][2], b1[nt][3]);
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) mma16816(hi[nt], af[3], b2[nt][2], b2[nt][3]);
#pragma unroll
    for (int nt = 0; nt < 4; ++nt)
#pragma unroll
        for (int e = 0; e < 4; ++e) acc[nt][e] += hi[nt][e];
}

// acc (16 x 64 as eight 8-wide column tiles) += A(16 x 32, two k tiles of fragments) . X[base .. base+32)[0..64)
__device__ __forceinline__ void mma_cols64(float (&acc)[8][4], const uint32_t (&af)[2][4], const __nv_bfloat16* X, int base, int lane) {
#pragma unroll
    for (int kt = 0; kt < 2; ++kt) {
        const __nv_bfloat16* xr = X + (base + kt * 16 + (lane & 15)) * MP + (lane >> 4) * 8;
#pragma unroll
        for (int nc = 0; nc < 8; nc += 2) {
            uint32_t b0, b1, b2, b3;
            ldsm_x4_trans(b0, b1, b2, b3, xr + nc * 8);
            mma16816(acc[nc], af[kt], b0, b1);
            mma16816(acc[nc + 1], af[kt], b2, b3);
        }
    }
}

// accumulator tiles of a 16 x 32 block -> A fragments of its two 16-wide k tiles
__device__ __forceinline__ void acc_to_a(uint32_t (&af)[2][4], const float (&x)[4][4]) {
#pragma unroll
    for (int kt = 0; kt < 2; ++kt) {
        af[kt][0] = pack_bf16(x[2 * kt][0], x[2 * kt][1]);
        af[kt][1] = pack_bf16(x[2 * kt][2], x[2 * kt][3]);
        af[kt][2] = pack_bf16(x[2 * kt + 1][0], x[2 * kt + 1][1]);
        af[kt][3] = pack_bf16(x[2 * kt + 1][2], x[2 * kt + 1][3]);
    }
}

__device__ __forceinline__ void store_rows(__nv_bfloat16* dst, long long ld, const HvSeqMap& m, long long seq, int r0, int L,
                                           int col0, int g, int tig, const float (&acc)[8][4]) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int r = r0 + g + 8 * h;
        if (r >= L) continue;
        __nv_bfloat16* p = dst + seq_row(m, seq, r) * ld + col0 + 2 * tig;
#pragma unroll
        for (int nc = 0; nc < 8; ++nc) *reinterpret_cast<uint32_t*>(p + nc * 8) = pack_bf16(acc[nc][2 * h], acc[nc][2 * h + 1]);
    }
}

__global__ void __launch_bounds__(128, 2) attn_bwd_dq_mma_kernel(const BwdArgs a, int Lp, int vec) {
    extern __shared__ __align__(16) unsigned char smraw[];
    __nv_bfloat16* Ks = reinterpret_cast<__nv_bfloat16*>(smraw);
    __nv_bfloat16* Vs = Ks + Lp * MP;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, tig = lane & 3;
    const long long seq = blockIdx.z;
    const int head = blockIdx.y, col0 = head * HDB;
    stage_head(Ks, (const __nv_bfloat16*)a.k, a.ldk, a.kvmap, seq, a.Lkv, Lp, col0, vec);
    cp_async_commit();
    stage_head(Vs, (const __nv_bfloat16*)a.v, a.ldv, a.kvmap, seq, a.Lkv, Lp, col0, vec);
    cp_async_commit();
    const int r0 = blockIdx.x * 64 + warp * 16;
    uint32_t qa[4][4], da[4][4];
    load_a_frags(qa, (const __nv_bfloat16*)a.q, a.ldq, a.qmap, seq, r0, a.Lq, col0, g, tig);
    load_a_frags(da, (const __nv_bfloat16*)a.dout, a.lddo, a.qmap, seq, r0, a.Lq, col0, g, tig);
    float del[2] = {0.f, 0.f};
    {
        uint32_t oa[4][4];
        load_a_frags(oa, (const __nv_bfloat16*)a.o, a.ldo, a.qmap, seq, r0, a.Lq, col0, g, tig);
#pragma unroll
        for (int kt = 0; kt < 4; ++kt)
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float2 d = unpack_bf16(da[kt][e]), o = unpack_bf16(oa[kt][e]);
                del[e & 1] += d.x * o.x + d.y * o.y;
            }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            del[h] += __shfl_xor_sync(0xffffffffu, del[h], 1);
            del[h] += __shfl_xor_sync(0xffffffffu, del[h], 2);
        }
    }
    cp_async_wait<1>();  // K has landed (V still in flight: pass 1 does not read it)
    __syncthreads();
    // pass 1: log-sum-exp of the scaled scores (online max / sum over this thread's columns, then across the quad)
    const bool active = r0 < a.Lq;  // warp-uniform; idle warps of the last tile only keep the barriers
    float mx[2] = {-1e30f, -1e30f}, sm[2] = {0.f, 0.f};
    for (int jc = 0; active && jc < Lp; jc += 32) {
        float s[4][4] = {};
        mma_rows32_split(s, qa, Ks, jc, lane);
        // branch-free online update: chunk maximum first, then one rescale of the running sum per row half
        float cm[2] = {mx[0], mx[1]};
#pragma unroll
        for (int nt = 0; nt < 4; ++nt)
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const bool valid = jc + nt * 8 + 2 * tig + (e & 1) < a.Lkv;
                s[nt][e] = valid ? s[nt][e] * a.scale : -1e30f;
                cm[e >> 1] = fmaxf(cm[e >> 1], s[nt][e]);
            }
#pragma unroll
        for (int h = 0; h < 2; ++h) { sm[h] *= __expf(mx[h] - cm[h]); mx[h] = cm[h]; }
#pragma unroll
        for (int nt = 0; nt < 4; ++nt)
#pragma unroll
            for (int e = 0; e < 4; ++e) sm[e >> 1] += s[nt][e] > -1e29f ? __expf(s[nt][e] - mx[e >> 1]) : 0.f;
    }
    float lse[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        float M = fmaxf(mx[h], __shfl_xor_sync(0xffffffffu, mx[h], 1));
        M = fmaxf(M, __shfl_xor_sync(0xffffffffu, M, 2));
        float L = sm[h] * __expf(mx[h] - M);
        L += __shfl_xor_sync(0xffffffffu, L, 1);
        L += __shfl_xor_sync(0xffffffffu, L, 2);
        lse[h] = M + __logf(L);
        const int i = r0 + g + 8 * h;
        if (tig == 0 && i < a.Lq) {
            float* st = a.stats + ((seq * a.heads + head) * a.Lq + i) * 2;
            st[0] = lse[h]; st[1] = del[h];
        }
    }
    cp_async_wait<0>();
    __syncthreads();
    if (!active) return;
    // pass 2
    float dq[8][4] = {};
    for (int jc = 0; jc < Lp; jc += 32) {
        float s[4][4] = {}, dp[4][4] = {};
        mma_rows32x2(s, qa, Ks, dp, da, Vs, jc, lane);
#pragma unroll
        for (int nt = 0; nt < 4; ++nt)
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int h = e >> 1;
                const float p = (jc + nt * 8 + 2 * tig + (e & 1) < a.Lkv) ? __expf(s[nt][e] * a.scale - lse[h]) : 0.f;
                s[nt][e] = p * (dp[nt][e] - del[h]) * a.scale;
            }
        uint32_t dsa[2][4];
        acc_to_a(dsa, s);
        mma_cols64(dq, dsa, Ks, jc, lane);
    }
    store_rows((__nv_bfloat16*)a.dq, a.lddq, a.qmap, seq, r0, a.Lq, col0, g, tig, dq);
}

__global__ void __launch_bounds__(128, 2) attn_bwd_dkv_mma_kernel(const BwdArgs a, int Lp, int vec) {
    extern __shared__ __align__(16) unsigned char smraw[];
    __nv_bfloat16* Qs = reinterpret_cast<__nv_bfloat16*>(smraw);
    __nv_bfloat16* Ds = Qs + Lp * MP;
    float* slse = reinterpret_cast<float*>(Ds + Lp * MP);
    float* sdel = slse + Lp;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, tig = lane & 3;
    const long long seq = blockIdx.z;
    const int head = blockIdx.y, col0 = head * HDB;
    stage_head(Qs, (const __nv_bfloat16*)a.q, a.ldq, a.qmap, seq, a.Lq, Lp, col0, vec);
    stage_head(Ds, (const __nv_bfloat16*)a.dout, a.lddo, a.qmap, seq, a.Lq, Lp, col0, vec);
    const float* st = a.stats + (seq * a.heads + head) * a.Lq * 2;
    for (int i = threadIdx.x; i < Lp; i += blockDim.x) {
        slse[i] = i < a.Lq ? st[2 * i] : INFINITY;
        sdel[i] = i < a.Lq ? st[2 * i + 1] : 0.f;
    }
    cp_async_commit();
    const int j0 = blockIdx.x * 64 + warp * 16;
    uint32_t ka[4][4], va[4][4];
    load_a_frags(ka, (const __nv_bfloat16*)a.k, a.ldk, a.kvmap, seq, j0, a.Lkv, col0, g, tig);
    load_a_frags(va, (const __nv_bfloat16*)a.v, a.ldv, a.kvmap, seq, j0, a.Lkv, col0, g, tig);
    cp_async_wait<0>();
    __syncthreads();
    if (j0 >= a.Lkv) return;  // warp-uniform, no block-wide barrier below
    float dk[8][4] = {}, dv[8][4] = {};
    for (int ic = 0; ic < Lp; ic += 32) {
        float s[4][4] = {}, dp[4][4] = {};
        mma_rows32x2(s, ka, Qs, dp, va, Ds, ic, lane);
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) {
            const int i = ic + nt * 8 + 2 * tig;
            const float l0 = slse[i], l1 = slse[i + 1], d0 = sdel[i], d1 = sdel[i + 1];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float p = __expf(s[nt][e] * a.scale - ((e & 1) ? l1 : l0));
                dp[nt][e] = p * (dp[nt][e] - ((e & 1) ? d1 : d0)) * a.scale;
                s[nt][e] = p;
            }
        }
        uint32_t pa[2][4], dsa[2][4];
        acc_to_a(pa, s);
        acc_to_a(dsa, dp);
        mma_cols64(dv, pa, Ds, ic, lane);
        mma_cols64(dk, dsa, Qs, ic, lane);
    }
    store_rows((__nv_bfloat16*)a.dv, a.lddv, a.kvmap, seq, j0, a.Lkv, col0, g, tig, dv);
    store_rows((__nv_bfloat16*)a.dk, a.lddk, a.kvmap, seq, j0, a.Lkv, col0, g, tig, dk);
}

inline bool al(const void* p, long long ld, int bytes) {
    return (reinterpret_cast<uintptr_t>(p) % bytes) == 0 && (ld * 2) % bytes == 0;
}

// returns HV_OK after launching, or 1 when the shapes / alignments need the SIMT kernels
int launch_bwd_mma(const BwdArgs& a, int64_t seqs, cudaStream_t st) {
    static const bool off = [] { const char* e = getenv("HV_ATTN_BWD_SIMT"); return e && *e == '1'; }();
    if (off) return 1;
    const bool ok4 = al(a.q, a.ldq, 4) && al(a.k, a.ldk, 4) && al(a.v, a.ldv, 4) && al(a.o, a.ldo, 4) && al(a.dout, a.lddo, 4) &&
                     al(a.dq, a.lddq, 4) && al(a.dk, a.lddk, 4) && al(a.dv, a.lddv, 4);
    const int Lpk = (a.Lkv + 31) / 32 * 32, Lpq = (a.Lq + 31) / 32 * 32;
    const size_t sm1 = (size_t)2 * Lpk * MP * 2, sm2 = (size_t)2 * Lpq * MP * 2 + (size_t)2 * Lpq * 4;
    if (!ok4 || sm1 > 227 * 1024 || sm2 > 227 * 1024) return 1;
    const int vec1 = al(a.k, a.ldk, 16) && al(a.v, a.ldv, 16), vec2 = al(a.q, a.ldq, 16) && al(a.dout, a.lddo, 16);
    HV_CUDA(cudaFuncSetAttribute(attn_bwd_dq_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm1));
    HV_CUDA(cudaFuncSetAttribute(attn_bwd_dkv_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm2));
    dim3 g1((a.Lq + 63) / 64, a.heads, (unsigned)seqs);
    attn_bwd_dq_mma_kernel<<<g1, 128, sm1, st>>>(a, Lpk, vec1);
    HV_LAUNCH_CHECK();
    dim3 g2((a.Lkv + 63) / 64, a.heads, (unsigned)seqs);
    attn_bwd_dkv_mma_kernel<<<g2, 128, sm2, st>>>(a, Lpq, vec2);
    HV_LAUNCH_CHECK();
    return HV_OK;
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

int hv_attention_bwd_tc_bf16(const void* q, long long ldq, const void* k, long long ldk, const void* v, long long ldv,
                             const void* o, long long ldo, const void* dout, long long lddo, void* dq, long long lddq,
                             void* dk, long long lddk, void* dv, long long lddv, long long seqs, int heads, int Lq, int Lkv,
                             const HvSeqMap& qm, const HvSeqMap& km, float scale, float* stats, int have_lse, cudaStream_t st);  // attn_bwd_tc.cu

int hv_attention_small_bwd_bf16(const void* q, long long ldq, const void* k, long long ldk, const void* v, long long ldv,
                                const void* o, long long ldo, const void* dout, long long lddo, void* dq, long long lddq,
                                void* dk, long long lddk, void* dv, long long lddv, long long seqs, int heads, int Lq, int Lkv,
                                const HvSeqMap& qm, const HvSeqMap& km, float scale, cudaStream_t st);  // attn_small_bwd.cu

// Gradients of out = softmax(q k^T * scale) v for head_dim 64 (no mask).  All matrices are (rows, heads*64) with their own
// leading dimension, one dtype (HV_F32 / HV_BF16); `stats` = seqs * heads * Lq * 2 floats of workspace.
static int attention_bwd_dispatch(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                                  const void* out, int64_t ldo, const void* dout, int64_t lddo,
                                  void* dq, int64_t lddq, void* dk, int64_t lddk, void* dv, int64_t lddv, int dtype,
                                  int64_t seqs, int heads, int hd, int Lq, int Lkv, const HvSeqMap* qmap, const HvSeqMap* kvmap,
                                  float scale, float* stats, int have_lse, void* stream) {
    HV_REQUIRE(hd == HDB, "hv_attention_bwd: head_dim %d != 64", hd);
    HV_REQUIRE(seqs >= 0 && heads > 0 && heads <= 65535 && seqs <= 65535 && Lq > 0 && Lkv > 0, "hv_attention_bwd: bad sizes");
    HV_REQUIRE(qmap && kvmap && qmap->sdiv > 0 && kvmap->sdiv > 0 && stats, "hv_attention_bwd: bad sequence maps / workspace");
    HV_REQUIRE(dtype == HV_F32 || dtype == HV_BF16, "hv_attention_bwd: bad dtype");
    if (seqs == 0) return HV_OK;
    BwdArgs a{q, k, v, out, dout, dq, dk, dv, ldq, ldk, ldv, ldo, lddo, lddq, lddk, lddv, heads, Lq, Lkv, *qmap, *kvmap, scale, stats};
    if (dtype == HV_F32) return launch_bwd<float>(a, seqs, hv_stream(stream));
    // bf16: tcgen05 / TMEM kernels (attn_bwd_tc.cu) for dense sequences of 16..288 tokens, then the warp-level tensor-core
    // kernels, then SIMT
    const int rs = hv_attention_small_bwd_bf16(q, ldq, k, ldk, v, ldv, out, ldo, dout, lddo, dq, lddq, dk, lddk, dv, lddv, seqs, heads, Lq, Lkv,
                                               *qmap, *kvmap, scale, hv_stream(stream));  // L <= 16: one warp per (sequence, head)
    if (rs != 0) return rs < 0 ? rs : HV_OK;
    const int rt = hv_attention_bwd_tc_bf16(q, ldq, k, ldk, v, ldv, out, ldo, dout, lddo, dq, lddq, dk, lddk, dv, lddv, seqs, heads, Lq, Lkv,
                                            *qmap, *kvmap, scale, stats, have_lse, hv_stream(stream));
    if (rt != 0) return rt < 0 ? rt : HV_OK;
    const int rc = launch_bwd_mma(a, seqs, hv_stream(stream));  // shape / alignment outside the tcgen05 kernels
    return rc == 1 ? launch_bwd<__nv_bfloat16>(a, seqs, hv_stream(stream)) : rc;
}

extern "C" int hv_attention_bwd(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                                const void* out, int64_t ldo, const void* dout, int64_t lddo,
                                void* dq, int64_t lddq, void* dk, int64_t lddk, void* dv, int64_t lddv, int dtype,
                                int64_t seqs, int heads, int hd, int Lq, int Lkv, const HvSeqMap* qmap, const HvSeqMap* kvmap,
                                float scale, float* stats, void* stream) {
    return attention_bwd_dispatch(q, ldq, k, ldk, v, ldv, out, ldo, dout, lddo, dq, lddq, dk, lddk, dv, lddv, dtype, seqs, heads, hd, Lq, Lkv,
                                  qmap, kvmap, scale, stats, 0, stream);
}

// hv_attention_bwd for a forward that ran through hv_attention_lse and returned 1: element 0 of every `stats` row already holds
// the log2-domain log-sum-exp, so the tcgen05 dq kernel skips its row-maximum pass and both cross-warp exchanges (kernels that
// do not use it recompute it).  have_lse = 0 is hv_attention_bwd.
extern "C" int hv_attention_bwd_lse(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv,
                                    const void* out, int64_t ldo, const void* dout, int64_t lddo,
                                    void* dq, int64_t lddq, void* dk, int64_t lddk, void* dv, int64_t lddv, int dtype,
                                    int64_t seqs, int heads, int hd, int Lq, int Lkv, const HvSeqMap* qmap, const HvSeqMap* kvmap,
                                    float scale, float* stats, int have_lse, void* stream) {
    return attention_bwd_dispatch(q, ldq, k, ldk, v, ldv, out, ldo, dout, lddo, dq, lddq, dk, lddk, dv, lddv, dtype, seqs, heads, hd, Lq, Lkv,
                                  qmap, kvmap, scale, stats, have_lse != 0, stream);
}
