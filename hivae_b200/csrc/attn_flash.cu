// bf16 flash attention (head_dim 64, no mask) for the spatial sequences (L = 264..274).
// CTA = 4 warps x 16 query rows; K/V streamed in 64-key chunks through a cp.async double buffer
// (XOR-swizzled 128-byte rows, ldmatrix fragments); QK^T and PV on mma.sync m16n8k16 bf16 with
// fp32 accumulation; online softmax in registers (exp2 with the scale folded in).
// First-generation tensor path for attention (3 % of the step's FLOPs); the tcgen05/TMEM variant
// replaces it once the GEMMs are at roofline ("next").
#include "attn.cuh"

namespace {

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void cp_async16(void* dst, const void* src, bool valid) {
    const int sz = valid ? 16 : 0;  // src-size 0 -> 16 bytes of zeros
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_addr(dst)), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], const void* p) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(smem_addr(p)));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t (&r)[4], const void* p) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(smem_addr(p)));
}
__device__ __forceinline__ void mma_bf16(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// tile of 64 rows x 64 bf16 (128 B per row); 16-byte chunk c of row r lives at chunk c ^ (r & 7)
__device__ __forceinline__ __nv_bfloat16* tile_ptr(__nv_bfloat16* tile, int r, int c) {
    return tile + r * 64 + ((c ^ (r & 7)) << 3);
}

__device__ __forceinline__ void load_tile(__nv_bfloat16* tile, const __nv_bfloat16* base, long long ld, const HvSeqMap& map,
                                          long long seq, int tok0, int L, int tid) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int idx = tid + i * 128, r = idx >> 3, c = idx & 7;
        const bool ok = tok0 + r < L;
        const long long row = ok ? seq_row(map, seq, tok0 + r) : 0;
        cp_async16(tile_ptr(tile, r, c), base + row * ld + c * 8, ok);
    }
}

__global__ void __launch_bounds__(128) attn_flash_kernel(AttnArgs a) {
    __shared__ __align__(128) __nv_bfloat16 Qs[64 * 64];
    __shared__ __align__(128) __nv_bfloat16 Ks[2][64 * 64];
    __shared__ __align__(128) __nv_bfloat16 Vs[2][64 * 64];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int qtiles = (a.Lq + 63) / 64;
    const long long seq = blockIdx.x / qtiles;
    const int q0 = (int)(blockIdx.x % qtiles) * 64, head = blockIdx.y;
    const __nv_bfloat16* qp = reinterpret_cast<const __nv_bfloat16*>(a.q) + head * 64;
    const __nv_bfloat16* kp = reinterpret_cast<const __nv_bfloat16*>(a.k) + head * 64;
    const __nv_bfloat16* vp = reinterpret_cast<const __nv_bfloat16*>(a.v) + head * 64;
    const int nchunks = (a.Lkv + 63) / 64;

    load_tile(Qs, qp, a.ldq, a.qmap, seq, q0, a.Lq, tid);
    load_tile(Ks[0], kp, a.ldk, a.kvmap, seq, 0, a.Lkv, tid);
    load_tile(Vs[0], vp, a.ldv, a.kvmap, seq, 0, a.Lkv, tid);
    cp_commit();

    const float sl2 = a.scale * 1.4426950408889634f;
    float o[8][4];
#pragma unroll
    for (int j = 0; j < 8; ++j) { o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f; }
    float mrow[2] = {-INFINITY, -INFINITY}, lrow[2] = {0.f, 0.f};
    uint32_t qf[4][4];

    for (int c = 0; c < nchunks; ++c) {
        if (c + 1 < nchunks) {
            load_tile(Ks[(c + 1) & 1], kp, a.ldk, a.kvmap, seq, (c + 1) * 64, a.Lkv, tid);
            load_tile(Vs[(c + 1) & 1], vp, a.ldv, a.kvmap, seq, (c + 1) * 64, a.Lkv, tid);
            cp_commit();
            cp_wait<1>();
        } else {
            cp_wait<0>();
        }
        __syncthreads();
        if (c == 0) {
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) ldsm_x4(qf[ks], tile_ptr(Qs, warp * 16 + (lane & 15), ks * 2 + (lane >> 4)));
        }
        __nv_bfloat16* Kt = Ks[c & 1];
        __nv_bfloat16* Vt = Vs[c & 1];
        // ---- S = Q K^T (16 x 64 per warp)
        float s[8][4];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
#pragma unroll
            for (int kp2 = 0; kp2 < 2; ++kp2) {
                uint32_t b[4];
                ldsm_x4(b, tile_ptr(Kt, j * 8 + (lane & 7), kp2 * 4 + (lane >> 3)));
                mma_bf16(s[j], qf[2 * kp2], b[0], b[1]);
                mma_bf16(s[j], qf[2 * kp2 + 1], b[2], b[3]);
            }
        }
        // ---- online softmax (rows g = lane/4 and g + 8)
        const int kbase = c * 64 + 2 * (lane & 3);
        float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
        for (int j = 0; j < 8; ++j) {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int key = kbase + j * 8 + (e & 1);
                s[j][e] = key < a.Lkv ? s[j][e] * sl2 : -INFINITY;
                mx[e >> 1] = fmaxf(mx[e >> 1], s[j][e]);
            }
        }
        float corr[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            mx[h] = fmaxf(mx[h], __shfl_xor_sync(0xffffffffu, mx[h], 1));
            mx[h] = fmaxf(mx[h], __shfl_xor_sync(0xffffffffu, mx[h], 2));
            const float mn = fmaxf(mrow[h], mx[h]);
            corr[h] = exp2f(mrow[h] - mn);
            mrow[h] = mn;
        }
        float ps[2] = {0.f, 0.f};
        uint32_t pf[4][4];  // P as A fragments for the 4 key steps of 16
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float p0 = exp2f(s[j][0] - mrow[0]), p1 = exp2f(s[j][1] - mrow[0]);
            const float p2 = exp2f(s[j][2] - mrow[1]), p3 = exp2f(s[j][3] - mrow[1]);
            ps[0] += p0 + p1; ps[1] += p2 + p3;
            pf[j >> 1][(j & 1) * 2 + 0] = hv_pack_bf16(p0, p1);
            pf[j >> 1][(j & 1) * 2 + 1] = hv_pack_bf16(p2, p3);
        }
#pragma unroll
        for (int h = 0; h < 2; ++h) lrow[h] = lrow[h] * corr[h] + ps[h];  // quad-partial sums, reduced at the end
#pragma unroll
        for (int j = 0; j < 8; ++j) { o[j][0] *= corr[0]; o[j][1] *= corr[0]; o[j][2] *= corr[1]; o[j][3] *= corr[1]; }
        // ---- O += P V
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
            for (int dp = 0; dp < 4; ++dp) {
                uint32_t b[4];
                ldsm_x4_t(b, tile_ptr(Vt, kk * 16 + (lane & 7) + ((lane >> 3) & 1) * 8, dp * 2 + (lane >> 4)));
                mma_bf16(o[2 * dp], pf[kk], b[0], b[1]);
                mma_bf16(o[2 * dp + 1], pf[kk], b[2], b[3]);
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        lrow[h] += __shfl_xor_sync(0xffffffffu, lrow[h], 1);
        lrow[h] += __shfl_xor_sync(0xffffffffu, lrow[h], 2);
    }
    __nv_bfloat16* op = reinterpret_cast<__nv_bfloat16*>(a.out) + head * 64 + 2 * (lane & 3);
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int qr = q0 + warp * 16 + (lane >> 2) + h * 8;
        if (qr < a.Lq) {
            const float inv = 1.f / lrow[h];
            __nv_bfloat16* orow = op + seq_row(a.qmap, seq, qr) * a.ldo;
#pragma unroll
            for (int j = 0; j < 8; ++j)
                *reinterpret_cast<uint32_t*>(orow + j * 8) = hv_pack_bf16(o[j][2 * h] * inv, o[j][2 * h + 1] * inv);
        }
    }
}

}  // namespace

int hv_attention_flash_bf16(const AttnArgs& a, int64_t seqs, cudaStream_t st) {
    dim3 grid((unsigned)(((a.Lq + 63) / 64) * seqs), (unsigned)a.heads);
    attn_flash_kernel<<<grid, 128, 0, st>>>(a);
    HV_LAUNCH_CHECK();
    return HV_OK;
}
