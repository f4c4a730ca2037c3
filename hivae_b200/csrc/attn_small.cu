// bf16 attention for tiny sequences (Lq, Lkv <= 16, head_dim 64, no mask): the per-pixel temporal
// attention of the DiT's BasicDiTBlock ('(n t) s d -> (n s) t d', 16 frames; model/transformer.py:2022-2028)
// and the camera encoder's 16 x 16 cross-attention (transformer.py:538-543).
//
// HBM-bound: 2048 x 16 (sequence, head) problems of 16 x 16 x 64 each per call at batch 8; every byte of
// q/k/v is read once and every output byte written once.  One warp owns one (sequence, head): the three
// 16 x 64 tiles are brought in with 16-byte cp.async (each token's head slice is one full 128-byte line),
// S = Q K^T and O = P V run on mma.sync m16n8k16 (a 16-row problem cannot fill a 128-row tcgen05 tile),
// the softmax lives in the accumulator registers, and the output goes back through shared memory as
// 128-byte row segments.  A CTA = 8 warps = 8 heads of one sequence, so each token row is touched as one
// contiguous 1 KB run.
#include "attn.cuh"

namespace {

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void cp_async16(void* dst, const void* src, bool valid) {
    const int sz = valid ? 16 : 0;  // src-size 0 -> 16 bytes of zeros
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_addr(dst)), "l"(src), "r"(sz) : "memory");
}
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
// 16 rows x 64 bf16; 16-byte chunk c of row r at chunk c ^ (r & 7)
__device__ __forceinline__ __nv_bfloat16* tile_ptr(__nv_bfloat16* tile, int r, int c) { return tile + r * 64 + ((c ^ (r & 7)) << 3); }

constexpr int WARPS = 8;

__global__ void __launch_bounds__(WARPS * 32) attn_small_kernel(AttnArgs a) {
    __shared__ __align__(128) __nv_bfloat16 tiles[WARPS][3][16 * 64];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long seq = blockIdx.x;
    const int head = blockIdx.y * WARPS + warp;
    if (head >= a.heads) return;  // whole warp; only warp-level synchronisation below
    __nv_bfloat16 *Qs = tiles[warp][0], *Ks = tiles[warp][1], *Vs = tiles[warp][2];
    const __nv_bfloat16* qp = reinterpret_cast<const __nv_bfloat16*>(a.q) + head * 64;
    const __nv_bfloat16* kp = reinterpret_cast<const __nv_bfloat16*>(a.k) + head * 64;
    const __nv_bfloat16* vp = reinterpret_cast<const __nv_bfloat16*>(a.v) + head * 64;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int idx = lane + 32 * i, r = idx >> 3, c = idx & 7;
        const bool okq = r < a.Lq, okk = r < a.Lkv;
        const long long rq = okq ? seq_row(a.qmap, seq, r) : 0, rk = okk ? seq_row(a.kvmap, seq, r) : 0;
        cp_async16(tile_ptr(Qs, r, c), qp + rq * a.ldq + c * 8, okq);
        cp_async16(tile_ptr(Ks, r, c), kp + rk * a.ldk + c * 8, okk);
        cp_async16(tile_ptr(Vs, r, c), vp + rk * a.ldv + c * 8, okk);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncwarp();

    // ---- S = Q K^T : 16 x 16, two 8-key column tiles
    float s[2][4];
#pragma unroll
    for (int j = 0; j < 2; ++j) { s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f; }
#pragma unroll
    for (int kp2 = 0; kp2 < 2; ++kp2) {
        uint32_t qa[4], qb[4];
        ldsm_x4(qa, tile_ptr(Qs, lane & 15, (2 * kp2) * 2 + (lane >> 4)));
        ldsm_x4(qb, tile_ptr(Qs, lane & 15, (2 * kp2 + 1) * 2 + (lane >> 4)));
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            uint32_t b[4];
            ldsm_x4(b, tile_ptr(Ks, j * 8 + (lane & 7), kp2 * 4 + (lane >> 3)));
            mma_bf16(s[j], qa, b[0], b[1]);
            mma_bf16(s[j], qb, b[2], b[3]);
        }
    }
    // ---- softmax over the 16 keys of rows g = lane/4 (regs 0,1) and g + 8 (regs 2,3)
    const float sl2 = a.scale * 1.4426950408889634f;
    const int kb = 2 * (lane & 3);
    float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
    for (int j = 0; j < 2; ++j)
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const int key = j * 8 + kb + (e & 1);
            s[j][e] = key < a.Lkv ? s[j][e] * sl2 : -INFINITY;
            mx[e >> 1] = fmaxf(mx[e >> 1], s[j][e]);
        }
    float sum[2] = {0.f, 0.f};
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        mx[h] = fmaxf(mx[h], __shfl_xor_sync(0xffffffffu, mx[h], 1));
        mx[h] = fmaxf(mx[h], __shfl_xor_sync(0xffffffffu, mx[h], 2));
    }
    uint32_t pa[4];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
        const float p0 = exp2f(s[j][0] - mx[0]), p1 = exp2f(s[j][1] - mx[0]);
        const float p2 = exp2f(s[j][2] - mx[1]), p3 = exp2f(s[j][3] - mx[1]);
        sum[0] += p0 + p1; sum[1] += p2 + p3;
        pa[2 * j] = hv_pack_bf16(p0, p1);
        pa[2 * j + 1] = hv_pack_bf16(p2, p3);
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        sum[h] += __shfl_xor_sync(0xffffffffu, sum[h], 1);
        sum[h] += __shfl_xor_sync(0xffffffffu, sum[h], 2);
    }
    // ---- O = P V : 16 x 64
    float o[8][4];
#pragma unroll
    for (int j = 0; j < 8; ++j) { o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f; }
#pragma unroll
    for (int dp = 0; dp < 4; ++dp) {
        uint32_t b[4];
        ldsm_x4_t(b, tile_ptr(Vs, (lane & 7) + ((lane >> 3) & 1) * 8, dp * 2 + (lane >> 4)));
        mma_bf16(o[2 * dp], pa, b[0], b[1]);
        mma_bf16(o[2 * dp + 1], pa, b[2], b[3]);
    }
    // ---- out = O / rowsum, staged through the Q tile so that global stores are 128-byte row segments
    __syncwarp();
    const float inv0 = 1.f / sum[0], inv1 = 1.f / sum[1];
    const int g = lane >> 2;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        *reinterpret_cast<uint32_t*>(tile_ptr(Qs, g, j) + kb) = hv_pack_bf16(o[j][0] * inv0, o[j][1] * inv0);
        *reinterpret_cast<uint32_t*>(tile_ptr(Qs, g + 8, j) + kb) = hv_pack_bf16(o[j][2] * inv1, o[j][3] * inv1);
    }
    __syncwarp();
    __nv_bfloat16* op = reinterpret_cast<__nv_bfloat16*>(a.out) + head * 64;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int idx = lane + 32 * i, r = idx >> 3, c = idx & 7;
        if (r < a.Lq)
            *reinterpret_cast<uint4*>(op + seq_row(a.qmap, seq, r) * a.ldo + c * 8) = *reinterpret_cast<const uint4*>(tile_ptr(Qs, r, c));
    }
}

}  // namespace

// 1 = launched, 0 = shape not handled here, < 0 = error
int hv_attention_small_bf16(const AttnArgs& a, int64_t seqs, cudaStream_t st) {
    if (a.hd != 64 || a.mask || a.Lq > 16 || a.Lkv > 16 || a.Lkv < 1) return 0;
    if ((a.ldq % 8) || (a.ldk % 8) || (a.ldv % 8) || (a.ldo % 8)) return 0;
    auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
    if (!al16(a.q) || !al16(a.k) || !al16(a.v) || !al16(a.out)) return 0;
    dim3 grid((unsigned)seqs, (unsigned)((a.heads + WARPS - 1) / WARPS));
    attn_small_kernel<<<grid, WARPS * 32, 0, st>>>(a);
    HV_LAUNCH_CHECK();
    return 1;
}
