// tcgen05 / TMEM attention for the spatial sequences of the DiT and the object encoder
// (head_dim 64, bf16, no mask, 16 <= Lkv <= 320, rows of a sequence contiguous):
//   softmax(Q K^T * scale) V   per (sequence, head)
// (F.scaled_dot_product_attention inside diffusers AttnProcessor2_0; call sites
//  model/modules.py:699-707, 1043-1051 of the reference).
//
// The whole key range of a sequence (<= 320 keys) fits in TMEM, so there is no online-softmax
// rescaling: per 128-query tile
//   S[128 x KP]  = Q K^T           tcgen05.mma, A = Q tile, B = K (both K-major, SWIZZLE_128B via TMA)
//   P            = exp2(S*c - max) 8 softmax warps (2 threads per query row, half the keys each),
//                                  bf16 P written to shared memory in the K-major SWIZZLE_128B layout
//   O[128 x 64]  = P V             tcgen05.mma, A = P (smem), B = V as an MN-major operand (V is stored
//                                  [key][dim], exactly what TMA delivers; no transpose pass)
//   out          = O / rowsum      read back with tcgen05.ld, scaled, stored as bf16
// CTAs are persistent over (sequence, head) pairs; K is reloaded while the previous pair's last
// tile is still in its softmax / PV phase, V while its epilogue runs.
//
// Warp roles (288 threads): warps 0..7 softmax + epilogue (TMEM lane quarter = warp % 4, key half =
// warp / 4), warp 8 = TMA producer + MMA issuer + TMEM owner (one elected lane).
#include <cuda.h>
#include "attn.cuh"
#include "tc_ptx.cuh"

int hv_make_tmap_bf16(const void* ptr, long long rows, int K, long long ld, int box_rows, CUtensorMap* out);  // gemm_tc.cu

namespace {

constexpr int QT = 128, HD = 64, KMAX = 320, THREADS = 288;
constexpr int O_COL = 0, S_COL = 64;  // TMEM columns: O accumulator [0,64), S [64, 64+KP)
constexpr uint32_t TMEM_COLS = 512;

struct TcArgs {
    __nv_bfloat16* out;
    long long ldo, q_sa, kv_sa;
    int heads, Lq, Lkv, KP;  // KP = Lkv rounded up to 16
    long long items;         // seqs * heads
    float sl2;               // scale * log2(e)
};

__global__ void __launch_bounds__(THREADS, 1)
attn_tc_kernel(const __grid_constant__ CUtensorMap tq, const __grid_constant__ CUtensorMap tk,
               const __grid_constant__ CUtensorMap tv, const TcArgs a) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t q_bar[2], k_bar, v_bar, s_full, p_full, o_full;
    __shared__ uint32_t tmem_slot;
    __shared__ float smax[2][QT], ssum[2][QT];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int KP = a.KP, KP32 = (KP + 31) & ~31;
    const int nq = (a.Lq + QT - 1) / QT;
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* Qs = smem;                       // 2 x [128 x 64] bf16
    uint8_t* Ks = Qs + 2 * QT * 128;          // [KP32 x 64]
    uint8_t* Vs = Ks + KP32 * 128;            // [KP32 x 64]
    uint8_t* Ps = Vs + KP32 * 128;            // nchunks x [128 x 64]

    if (threadIdx.x == 0) {
        tc::prefetch_tmap(&tq); tc::prefetch_tmap(&tk); tc::prefetch_tmap(&tv);
        tc::mbar_init(&q_bar[0], 1); tc::mbar_init(&q_bar[1], 1);
        tc::mbar_init(&k_bar, 1); tc::mbar_init(&v_bar, 1);
        tc::mbar_init(&s_full, 1); tc::mbar_init(&p_full, 256); tc::mbar_init(&o_full, 1);
        tc::fence_barrier_init();
    }
    if (warp == 8) tc::tmem_alloc(&tmem_slot, TMEM_COLS);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem_base = tmem_slot;

    if (warp == 8) {
        if (lane == 0) {
            // ================================================= TMA producer + MMA issuer
            auto issue_s = [&](long long T) {
                tc::fence_after_sync();
                const uint64_t da = tc::make_smem_desc_k128(tc::smem_u32(Qs + (T & 1) * (QT * 128)));
                for (int n0 = 0; n0 < KP; n0 += 256) {
                    const int nn = KP - n0 < 256 ? KP - n0 : 256;
                    const uint32_t idesc = tc::make_idesc_bf16(QT, nn);
                    const uint64_t db = tc::make_smem_desc_k128(tc::smem_u32(Ks + n0 * 128));
#pragma unroll
                    for (int k = 0; k < HD / 16; ++k)
                        tc::umma_bf16(tmem_base + S_COL + n0, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, k != 0);
                }
                tc::umma_commit(&s_full);
            };
            auto issue_pv = [&]() {
                tc::fence_after_sync();
                constexpr uint32_t idesc = tc::make_idesc_bf16_bmn(QT, HD);
                for (int j = 0; j < KP / 16; ++j) {
                    const uint64_t da = tc::make_smem_desc_k128(tc::smem_u32(Ps + (j >> 2) * (QT * 128))) + (uint64_t)(2 * (j & 3));
                    const uint64_t db = tc::make_smem_desc_k128(tc::smem_u32(Vs + j * 2048));  // 16 keys x 128 B, MN-major
                    tc::umma_bf16(tmem_base + O_COL, da, db, idesc, j != 0);
                }
                tc::umma_commit(&o_full);
            };
            long long T0 = 0, it = 0;
            for (long long w = blockIdx.x; w < a.items; w += gridDim.x, ++it, T0 += nq) {
                const long long seq = w / a.heads;
                const int head = (int)(w - seq * a.heads), col = head * HD;
                const long long qrow0 = seq * a.q_sa, kvrow0 = seq * a.kv_sa;
                // K (free: every S MMA of the previous pair completed before its p_full)
                tc::mbar_expect_tx(&k_bar, KP32 * 128);
                for (int b = 0; b < KP32 / 32; ++b) tc::tma_load_2d(Ks + b * 4096, &tk, &k_bar, col, (int)(kvrow0 + b * 32));
                for (int i = 0; i < (nq < 2 ? nq : 2); ++i) {
                    const long long T = T0 + i;
                    uint8_t* dst = Qs + (T & 1) * (QT * 128);
                    tc::mbar_expect_tx(&q_bar[T & 1], QT * 128);
#pragma unroll
                    for (int b = 0; b < QT / 32; ++b)
                        tc::tma_load_2d(dst + b * 4096, &tq, &q_bar[T & 1], col, (int)(qrow0 + i * QT + b * 32));
                }
                if (it > 0) tc::mbar_wait(&o_full, (uint32_t)((T0 - 1) & 1));  // V free once the previous PV retired
                tc::mbar_expect_tx(&v_bar, KP32 * 128);
                for (int b = 0; b < KP32 / 32; ++b) tc::tma_load_2d(Vs + b * 4096, &tv, &v_bar, col, (int)(kvrow0 + b * 32));
                tc::mbar_wait(&k_bar, (uint32_t)(it & 1));
                tc::mbar_wait(&q_bar[T0 & 1], (uint32_t)((T0 >> 1) & 1));
                issue_s(T0);
                for (int i = 0; i < nq; ++i) {
                    const long long T = T0 + i;
                    tc::mbar_wait(&p_full, (uint32_t)(T & 1));  // P(T) written, S free
                    if (i == 0) tc::mbar_wait(&v_bar, (uint32_t)(it & 1));
                    issue_pv();
                    if (i + 1 < nq) {
                        tc::mbar_wait(&q_bar[(T + 1) & 1], (uint32_t)(((T + 1) >> 1) & 1));
                        issue_s(T + 1);
                    }
                    if (i + 2 < nq) {  // Q buffer (T & 1) is free: S(T) retired
                        uint8_t* dst = Qs + (T & 1) * (QT * 128);
                        tc::mbar_expect_tx(&q_bar[T & 1], QT * 128);
#pragma unroll
                        for (int b = 0; b < QT / 32; ++b)
                            tc::tma_load_2d(dst + b * 4096, &tq, &q_bar[T & 1], col, (int)(qrow0 + (i + 2) * QT + b * 32));
                    }
                }
            }
        }
    } else {
        // ===================================================== softmax + epilogue warps
        const int q = warp & 3, h = warp >> 2, row = q * 32 + lane;
        const int nb = KP >> 4, hb = (nb + 1) >> 1;
        const int c_begin = h ? hb * 16 : 0, c_end = h ? KP : hb * 16;
        const uint32_t t_lane = tmem_base + ((uint32_t)(q * 32) << 16);
        uint8_t* prow = Ps + (row >> 3) * 1024 + (row & 7) * 128;  // this row inside every 64-key P chunk
        long long T = 0;
        for (long long w = blockIdx.x; w < a.items; w += gridDim.x) {
            const long long seq = w / a.heads;
            const int head = (int)(w - seq * a.heads);
            for (int i = 0; i < nq; ++i, ++T) {
                tc::mbar_wait(&s_full, (uint32_t)(T & 1));
                tc::fence_after_sync();
                // ---- pass 1: row maximum over this thread's half of the keys
                float mx = -INFINITY;
                for (int c = c_begin; c < c_end; c += 16) {
                    float v[16];
                    tc::tmem_ld16(t_lane + S_COL + c, v);
#pragma unroll
                    for (int j = 0; j < 16; ++j) if (c + j < a.Lkv) mx = fmaxf(mx, v[j]);
                }
                smax[h][row] = mx;
                tc::named_bar_sync(1, 256);
                mx = fmaxf(mx, smax[h ^ 1][row]);
                const float msc = mx * a.sl2;
                // ---- pass 2: P = exp2(S*c - max*c) -> bf16 -> shared memory (K-major, 128 B swizzle)
                float sum = 0.f;
                for (int c = c_begin; c < c_end; c += 16) {
                    float v[16];
                    tc::tmem_ld16(t_lane + S_COL + c, v);
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        v[j] = (c + j < a.Lkv) ? tc::ex2_approx(fmaf(v[j], a.sl2, -msc)) : 0.f;
                        sum += v[j];
                    }
                    uint8_t* pc = prow + (c >> 6) * (QT * 128);
                    const int jc = (c & 63) >> 3;
                    *reinterpret_cast<uint4*>(pc + (((jc) ^ (row & 7)) << 4)) =
                        make_uint4(hv_pack_bf16(v[0], v[1]), hv_pack_bf16(v[2], v[3]), hv_pack_bf16(v[4], v[5]), hv_pack_bf16(v[6], v[7]));
                    *reinterpret_cast<uint4*>(pc + (((jc + 1) ^ (row & 7)) << 4)) =
                        make_uint4(hv_pack_bf16(v[8], v[9]), hv_pack_bf16(v[10], v[11]), hv_pack_bf16(v[12], v[13]), hv_pack_bf16(v[14], v[15]));
                }
                ssum[h][row] = sum;
                tc::fence_before_sync();
                tc::fence_proxy_async();
                tc::mbar_arrive(&p_full);
                tc::named_bar_sync(1, 256);
                const float inv = 1.f / (sum + ssum[h ^ 1][row]);
                // ---- epilogue: O / rowsum -> bf16 (this thread: 32 of the 64 output dims of its row)
                tc::mbar_wait(&o_full, (uint32_t)(T & 1));
                tc::fence_after_sync();
                float o[32];
                tc::tmem_ld32(t_lane + O_COL + h * 32, o);
                tc::fence_before_sync();
                const int qr = i * QT + row;
                if (qr < a.Lq) {
                    __nv_bfloat16* op = a.out + (seq * a.q_sa + qr) * a.ldo + head * HD + h * 32;
#pragma unroll
                    for (int j = 0; j < 32; j += 8)
                        *reinterpret_cast<uint4*>(op + j) =
                            make_uint4(hv_pack_bf16(o[j] * inv, o[j + 1] * inv), hv_pack_bf16(o[j + 2] * inv, o[j + 3] * inv),
                                       hv_pack_bf16(o[j + 4] * inv, o[j + 5] * inv), hv_pack_bf16(o[j + 6] * inv, o[j + 7] * inv));
                }
            }
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 8) {
        tc::fence_after_sync();
        tc::tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

}  // namespace

// returns 1 if the shape is handled here (and launched), 0 if the caller should use another kernel, < 0 on error
int hv_attention_tc_bf16(const AttnArgs& a, int64_t seqs, cudaStream_t st) {
    const HvSeqMap &qm = a.qmap, &km = a.kvmap;
    const bool dense = qm.sdiv == 1 && qm.sc == 1 && km.sdiv == 1 && km.sc == 1;
    if (!dense || a.hd != HD || a.mask || a.Lkv < 16 || a.Lkv > KMAX) return 0;
    if ((a.ldq % 8) || (a.ldk % 8) || (a.ldv % 8) || (a.ldo % 8)) return 0;
    auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
    if (!al16(a.q) || !al16(a.k) || !al16(a.v) || !al16(a.out)) return 0;
    const long long qrows = (seqs - 1) * qm.sa + a.Lq, kvrows = (seqs - 1) * km.sa + a.Lkv;
    if (qrows >= (1ll << 31) || kvrows >= (1ll << 31)) return 0;
    CUtensorMap tq, tk, tv;
    int rc;
    const int cols = a.heads * HD;
    if ((rc = hv_make_tmap_bf16(a.q, qrows, cols, a.ldq, 32, &tq)) != HV_OK) return rc;
    if ((rc = hv_make_tmap_bf16(a.k, kvrows, cols, a.ldk, 32, &tk)) != HV_OK) return rc;
    if ((rc = hv_make_tmap_bf16(a.v, kvrows, cols, a.ldv, 32, &tv)) != HV_OK) return rc;
    TcArgs t;
    t.out = reinterpret_cast<__nv_bfloat16*>(a.out);
    t.ldo = a.ldo; t.q_sa = qm.sa; t.kv_sa = km.sa;
    t.heads = a.heads; t.Lq = a.Lq; t.Lkv = a.Lkv; t.KP = (a.Lkv + 15) & ~15;
    t.items = seqs * a.heads;
    t.sl2 = a.scale * 1.4426950408889634f;
    const int KP32 = (t.KP + 31) & ~31, nchunks = (t.KP + 63) >> 6;
    const int smem = 2 * QT * 128 + 2 * KP32 * 128 + nchunks * QT * 128 + 1024;
    static int configured = 0;
    if (configured < smem) {
        HV_CUDA(cudaFuncSetAttribute(attn_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        configured = smem;
    }
    const int sms = hv_num_sms();
    const int grid = (int)(t.items < sms ? t.items : sms);
    attn_tc_kernel<<<grid, THREADS, smem, st>>>(tq, tk, tv, t);
    HV_LAUNCH_CHECK();
    return 1;
}
