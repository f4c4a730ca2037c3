// tcgen05 / TMEM attention BACKWARD for the spatial sequences of the DiT and the object encoder (head_dim 64, bf16, no
// mask, 16 <= L <= 288, rows of a sequence contiguous): gradients of out = softmax(Q K^T * scale) V, i.e. autograd of
// F.scaled_dot_product_attention at model/modules.py:699-707, 1043-1051 of the reference.
//
//   P = softmax(S), S = Q K^T * scale;  dP = dO V^T;  D_i = sum_d dO_id O_id;  dS = P o (dP - D) * scale
//   dQ = dS K        dK = dS^T Q        dV = P^T dO
//
// Two kernels, both "S-like MMA -> elementwise on TMEM -> PV-like MMA" like the forward (attn_tc.cu):
//   dq kernel,  one 128-QUERY tile at a time:  S = Q K^T into TMEM (<= 288 columns); the 256 compute threads (thread = (row,
//       key half)) take the row max, then P~ = exp2(S c - max c) which they KEEP IN REGISTERS as packed bf16 (72 registers) —
//       so that dP = dO V^T can overwrite the very same TMEM columns; dS~ = P~ (dP - D) -> bf16 -> shared memory (K-major,
//       128 B swizzle) -> dQ~ = dS~ K with K as an MN-major operand (no transpose) -> dQ = dQ~ * scale / rowsum.
//       Also writes (log2-sum-exp, D) per query row for the second kernel.
//   dkv kernel, one 128-KEY tile at a time:  S^T = K_t Q^T (keys on the TMEM lanes, <= 288 query columns),
//       P^T = exp2(S^T c - lse_q) exact (lse from the dq kernel), kept in registers AND staged to shared memory,
//       dV_t = P^T dO (dO MN-major);  dP^T = V_t dO^T overwrites S^T;  dS^T = P^T (dP^T - D_q) scale -> the same staging
//       buffer once the dV MMA has retired -> dK_t = dS^T Q (Q MN-major).
// No atomics anywhere: every output element is produced by exactly one CTA (bit-reproducible).
//
// Warp roles (320 threads): warps 0..7 compute (TMEM lane quarter = warp % 4, column half = warp / 4), warp 8 = MMA issuer +
// TMEM owner, warp 9 = TMA producer.  CTAs are persistent over (sequence, head) pairs.
#include <cuda.h>
#include <cstdlib>
#include "attn.cuh"
#include "tc_ptx.cuh"

int hv_make_tmap_bf16(const void* ptr, long long rows, int K, long long ld, int box_rows, CUtensorMap* out);  // gemm_tc.cu

namespace {

constexpr int QT = 128, HD = 64, LMAX = 288, THREADS = 320;
constexpr int ACC_COL = 0, S_COL = 128;  // TMEM: accumulators (dQ | dV, dK) in [0, 128), S / dP in [128, 416)
constexpr uint32_t TMEM_COLS = 512;
constexpr int TBYTES = QT * 128;         // one 128-row x 64-column bf16 tile

struct BwdTc {
    const __nv_bfloat16 *o, *dout;
    __nv_bfloat16 *dq, *dk, *dv;
    long long ldo, lddo, lddq, lddk, lddv, q_sa, kv_sa;
    int heads, Lq, Lkv, QP, KP;  // QP / KP = Lq / Lkv rounded up to 16
    long long items;             // seqs * heads
    float sl2, scale;            // scale * log2(e), scale
    float* stats;                // [item][Lq][2] = (log2-domain log-sum-exp, D)
    unsigned long long* dbg;     // phase cycle counters (HV_ATTN_BWD_PROF=1) or nullptr: [kernel][cta][16]
};

// cycles since the previous tick of (warp 0, lane 0) -> counter `phase` of this CTA (only when profiling is on)
#define BWD_TICK(phase) do { if (prof) { const long long t_ = clock64(); a.dbg[(KERNEL_ID * 148 + (blockIdx.x % 148)) * 16 + (phase)] += (unsigned long long)(t_ - tk0); tk0 = t_; } } while (0)

__device__ __forceinline__ void ld16_issue(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void ld16_wait(uint32_t (&r)[16]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                   "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15])
                 :
                 : "memory");
}

// 16 consecutive K-dimension elements (columns c .. c+15 of `row`) of a bf16 A operand staged as 64-column chunks of
// [128 rows x 128 B] in the K-major SWIZZLE_128B layout
__device__ __forceinline__ void stage16(uint8_t* base, int row, int c, const float (&v)[16]) {
    uint8_t* pc = base + (c >> 6) * TBYTES + (row >> 3) * 1024 + (row & 7) * 128;
    const int jc = (c & 63) >> 3;
    *reinterpret_cast<uint4*>(pc + ((jc ^ (row & 7)) << 4)) =
        make_uint4(hv_pack_bf16(v[0], v[1]), hv_pack_bf16(v[2], v[3]), hv_pack_bf16(v[4], v[5]), hv_pack_bf16(v[6], v[7]));
    *reinterpret_cast<uint4*>(pc + (((jc + 1) ^ (row & 7)) << 4)) =
        make_uint4(hv_pack_bf16(v[8], v[9]), hv_pack_bf16(v[10], v[11]), hv_pack_bf16(v[12], v[13]), hv_pack_bf16(v[14], v[15]));
}
__device__ __forceinline__ float bf_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf_hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }

// S-like product into TMEM: D[128 x n] = A[128 x 64] B[n x 64]^T, both K-major tiles; n <= 288 is split into two MMAs
__device__ __forceinline__ void issue_s_like(uint32_t d_tmem, const uint8_t* a_tile, const uint8_t* b_rows, int n, int half_rows) {
    const uint64_t da = tc::make_smem_desc_k128(tc::smem_u32(a_tile));
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int r0 = h ? half_rows : 0, cols = h ? n - half_rows : half_rows;
        if (cols <= 0) continue;
        const uint64_t db = tc::make_smem_desc_k128(tc::smem_u32(b_rows + r0 * 128));
        const uint32_t idesc = tc::make_idesc_bf16(QT, cols);
#pragma unroll
        for (int k = 0; k < HD / 16; ++k) tc::umma_bf16(d_tmem + (uint32_t)r0, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, k != 0);
    }
}
// PV-like product: D[128 x 64] = A[128 x kdim] B[kdim x 64], A staged K-major in 64-column chunks, B = [kdim rows][64] rows of
// 128 B as TMA delivers them (MN-major operand)
__device__ __forceinline__ void issue_pv_like(uint32_t d_tmem, const uint8_t* a_stage, const uint8_t* b_rows, int kdim) {
    constexpr uint32_t idesc = tc::make_idesc_bf16_bmn(QT, HD);
    for (int j = 0; j < kdim / 16; ++j) {
        const uint64_t da = tc::make_smem_desc_k128(tc::smem_u32(a_stage + (j >> 2) * TBYTES)) + (uint64_t)(2 * (j & 3));
        const uint64_t db = tc::make_smem_desc_k128(tc::smem_u32(b_rows + j * 2048));
        tc::umma_bf16(d_tmem, da, db, idesc, j != 0);
    }
}

// ------------------------------------------------------------------------------------------------ dQ (+ stats)
// PIPE: keep the next 16-column TMEM load in flight while the current one is processed (32 more live registers)
template <bool PIPE>
__global__ void __launch_bounds__(THREADS, 1)
attn_bwd_dq_tc_kernel(const __grid_constant__ CUtensorMap tq, const __grid_constant__ CUtensorMap tk,
                      const __grid_constant__ CUtensorMap tv, const __grid_constant__ CUtensorMap tdo, const BwdTc a) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t kv_full, kv_free, q_full, q_free, do_full, do_free, s_full, s_free, dp_full, ds_full, dq_full;
    __shared__ uint32_t tmem_slot;
    __shared__ float smax[2][QT], ssum[2][QT];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int KP = a.KP, KP32 = (KP + 31) & ~31;
    const int nq = (a.Lq + QT - 1) / QT;
    const int nb = KP >> 4, HB = ((nb + 1) >> 1) * 16;  // key half 0 = [0, HB), half 1 = [HB, KP)
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* Qs = smem;                    // [128 x 64]
    uint8_t* dOs = Qs + TBYTES;            // [128 x 64]
    uint8_t* Ks = dOs + TBYTES;            // [KP32 x 64]
    uint8_t* Vs = Ks + KP32 * 128;         // [KP32 x 64]
    uint8_t* dSs = Vs + KP32 * 128;        // nchunks x [128 x 64]

    if (threadIdx.x == 0) {
        tc::prefetch_tmap(&tq); tc::prefetch_tmap(&tk); tc::prefetch_tmap(&tv); tc::prefetch_tmap(&tdo);
        tc::mbar_init(&kv_full, 1); tc::mbar_init(&kv_free, 1); tc::mbar_init(&q_full, 1); tc::mbar_init(&q_free, 1);
        tc::mbar_init(&do_full, 1); tc::mbar_init(&do_free, 1); tc::mbar_init(&s_full, 1); tc::mbar_init(&s_free, 8);
        tc::mbar_init(&dp_full, 1); tc::mbar_init(&ds_full, 8); tc::mbar_init(&dq_full, 1);   // compute warps arrive once per warp
        tc::fence_barrier_init();
    }
    if (warp == 8) tc::tmem_alloc(&tmem_slot, TMEM_COLS);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem_base = tmem_slot;

    if (warp == 9) {
        // ===================================================== TMA producer
        if (lane == 0) {
            long long T = 0, it = 0;
            for (long long w = blockIdx.x; w < a.items; w += gridDim.x, ++it) {
                const long long seq = w / a.heads;
                const int col = (int)(w - seq * a.heads) * HD;
                const long long qrow0 = seq * a.q_sa, kvrow0 = seq * a.kv_sa;
                tc::mbar_wait(&kv_free, (uint32_t)((it & 1) ^ 1));
                tc::mbar_expect_tx(&kv_full, 2 * KP32 * 128);
                for (int j = 0; j < KP32 / 32; ++j) {
                    tc::tma_load_2d(Ks + j * 4096, &tk, &kv_full, col, (int)(kvrow0 + j * 32));
                    tc::tma_load_2d(Vs + j * 4096, &tv, &kv_full, col, (int)(kvrow0 + j * 32));
                }
                for (int i = 0; i < nq; ++i, ++T) {
                    tc::mbar_wait(&q_free, (uint32_t)((T & 1) ^ 1));
                    tc::mbar_expect_tx(&q_full, TBYTES);
                    for (int j = 0; j < 4; ++j) tc::tma_load_2d(Qs + j * 4096, &tq, &q_full, col, (int)(qrow0 + i * QT + j * 32));
                    tc::mbar_wait(&do_free, (uint32_t)((T & 1) ^ 1));
                    tc::mbar_expect_tx(&do_full, TBYTES);
                    for (int j = 0; j < 4; ++j) tc::tma_load_2d(dOs + j * 4096, &tdo, &do_full, col, (int)(qrow0 + i * QT + j * 32));
                }
            }
        }
    } else if (warp == 8) {
        // ===================================================== MMA issuer
        if (lane == 0) {
            long long T = 0, it = 0;
            for (long long w = blockIdx.x; w < a.items; w += gridDim.x, ++it) {
                tc::mbar_wait(&kv_full, (uint32_t)(it & 1));
                for (int i = 0; i < nq; ++i, ++T) {
                    const uint32_t ph = (uint32_t)(T & 1);
                    tc::mbar_wait(&q_full, ph);
                    tc::fence_after_sync();
                    issue_s_like(tmem_base + S_COL, Qs, Ks, KP, HB);       // S = Q K^T  (the previous tile's dP was consumed: ds_full below)
                    tc::umma_commit(&s_full);
                    tc::umma_commit(&q_free);
                    tc::mbar_wait(&do_full, ph);
                    tc::mbar_wait(&s_free, ph);                            // P~ is in registers, S may be overwritten
                    tc::fence_after_sync();
                    issue_s_like(tmem_base + S_COL, dOs, Vs, KP, HB);      // dP = dO V^T
                    tc::umma_commit(&dp_full);
                    tc::umma_commit(&do_free);
                    tc::mbar_wait(&ds_full, ph);                           // dS~ staged; the previous tile's dQ was read back
                    tc::fence_after_sync();
                    issue_pv_like(tmem_base + ACC_COL, dSs, Ks, KP);       // dQ~ = dS~ K
                    tc::umma_commit(&dq_full);
                    if (i == nq - 1) tc::umma_commit(&kv_free);
                }
            }
        }
    } else {
        // ===================================================== compute warps
        const int q = warp & 3, h = warp >> 2, row = q * 32 + lane;
        const int c_begin = h ? HB : 0, c_end = h ? KP : HB;
        const uint32_t t_lane = tmem_base + ((uint32_t)(q * 32) << 16);
        const uint32_t s_addr = t_lane + S_COL;
        long long T = 0;
        for (long long w = blockIdx.x; w < a.items; w += gridDim.x) {
            const long long seq = w / a.heads;
            const int head = (int)(w - seq * a.heads);
            for (int i = 0; i < nq; ++i, ++T) {
                const uint32_t ph = (uint32_t)(T & 1);
                const int qr = i * QT + row;
                const bool valid = qr < a.Lq;
                constexpr int KERNEL_ID = 0;
                const bool prof = a.dbg != nullptr && threadIdx.x == 0;
                long long tk0 = prof ? clock64() : 0;
                // D = sum_d dO O of this thread's query row (both key halves compute it: 2 x 128 B per row)
                float Dr = 0.f;
                if (valid) {
                    const uint4* po = reinterpret_cast<const uint4*>(a.o + (seq * a.q_sa + qr) * a.ldo + head * HD);
                    const uint4* pd = reinterpret_cast<const uint4*>(a.dout + (seq * a.q_sa + qr) * a.lddo + head * HD);
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const uint4 x = __ldg(po + j), y = __ldg(pd + j);
                        Dr += bf_lo(x.x) * bf_lo(y.x) + bf_hi(x.x) * bf_hi(y.x) + bf_lo(x.y) * bf_lo(y.y) + bf_hi(x.y) * bf_hi(y.y)
                            + bf_lo(x.z) * bf_lo(y.z) + bf_hi(x.z) * bf_hi(y.z) + bf_lo(x.w) * bf_lo(y.w) + bf_hi(x.w) * bf_hi(y.w);
                    }
                }
                BWD_TICK(0);   // D loads
                tc::mbar_wait(&s_full, ph);
                tc::fence_after_sync();
                BWD_TICK(1);   // wait S
                // ---- pass 1: row maximum over this thread's key half
                float mx = -INFINITY;
                {
                    uint32_t buf[2][16];
                    if (PIPE) ld16_issue(s_addr + c_begin, buf[0]);
#pragma unroll
                    for (int k = 0; k < 9; ++k) {
                        const int c = c_begin + 16 * k;
                        if (c < c_end) {  // warp-uniform; the next chunk is in flight while this one is reduced
                            if (!PIPE) ld16_issue(s_addr + c, buf[k & 1]);
                            ld16_wait(buf[k & 1]);
                            if (PIPE && c + 16 < c_end) ld16_issue(s_addr + c + 16, buf[(k + 1) & 1]);
                            const uint32_t (&va)[16] = buf[k & 1];
                            if (c + 16 <= a.Lkv) {
                                float m4[4];
#pragma unroll
                                for (int j = 0; j < 4; ++j)
                                    m4[j] = fmaxf(fmaxf(__uint_as_float(va[j]), __uint_as_float(va[j + 4])), fmaxf(__uint_as_float(va[j + 8]), __uint_as_float(va[j + 12])));
                                mx = fmaxf(mx, fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3])));
                            } else {
#pragma unroll
                                for (int j = 0; j < 16; ++j) if (c + j < a.Lkv) mx = fmaxf(mx, __uint_as_float(va[j]));
                            }
                        }
                    }
                }
                smax[h][row] = mx;
                BWD_TICK(2);   // pass 1
                tc::named_bar_sync(1, 256);
                mx = fmaxf(mx, smax[h ^ 1][row]);
                const float msc = mx * a.sl2;
                BWD_TICK(3);   // max exchange
                // ---- pass 2: P~ = exp2(S c - max c) -> registers (packed bf16), row sum
                uint32_t preg[72];
                float sum = 0.f;
                {
                    uint32_t buf[2][16];
                    if (PIPE) ld16_issue(s_addr + c_begin, buf[0]);
#pragma unroll
                    for (int k = 0; k < 9; ++k) {
                        const int c = c_begin + 16 * k;
                        if (c < c_end) {  // warp-uniform
                            if (!PIPE) ld16_issue(s_addr + c, buf[k & 1]);
                            ld16_wait(buf[k & 1]);
                            if (PIPE && c + 16 < c_end) ld16_issue(s_addr + c + 16, buf[(k + 1) & 1]);
                            const uint32_t (&va)[16] = buf[k & 1];
                            float p[16];
#pragma unroll
                            for (int j = 0; j < 16; ++j)
                                p[j] = (c + j < a.Lkv) ? tc::ex2_approx(fmaf(__uint_as_float(va[j]), a.sl2, -msc)) : 0.f;
#pragma unroll
                            for (int j = 0; j < 8; ++j) {
                                preg[8 * k + j] = hv_pack_bf16(p[2 * j], p[2 * j + 1]);
                                sum += bf_lo(preg[8 * k + j]) + bf_hi(preg[8 * k + j]);  // the sum of what the MMAs will see
                            }
                        }
                    }
                }
                BWD_TICK(4);   // pass 2
                tc::fence_before_sync();
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(&s_free);
                ssum[h][row] = sum;
                tc::named_bar_sync(1, 256);
                const float l = sum + ssum[h ^ 1][row];
                if (h == 0 && valid) {
                    float2* st = reinterpret_cast<float2*>(a.stats) + (w * a.Lq + qr);
                    *st = make_float2(msc + __log2f(l), Dr);
                }
                // ---- pass 3: dS~ = P~ (dP - D) -> bf16 -> shared memory
                BWD_TICK(5);   // sum exchange + stats
                tc::mbar_wait(&dp_full, ph);
                tc::fence_after_sync();
                BWD_TICK(6);   // wait dP
                {
                    uint32_t buf[2][16];
                    if (PIPE) ld16_issue(s_addr + c_begin, buf[0]);
#pragma unroll
                    for (int k = 0; k < 9; ++k) {
                        const int c = c_begin + 16 * k;
                        if (c < c_end) {
                            if (!PIPE) ld16_issue(s_addr + c, buf[k & 1]);
                            ld16_wait(buf[k & 1]);
                            if (PIPE && c + 16 < c_end) ld16_issue(s_addr + c + 16, buf[(k + 1) & 1]);
                            const uint32_t (&va)[16] = buf[k & 1];
                            float ds[16];
#pragma unroll
                            for (int j = 0; j < 8; ++j) {
                                ds[2 * j] = bf_lo(preg[8 * k + j]) * (__uint_as_float(va[2 * j]) - Dr);
                                ds[2 * j + 1] = bf_hi(preg[8 * k + j]) * (__uint_as_float(va[2 * j + 1]) - Dr);
                            }
                            stage16(dSs, row, c, ds);
                        }
                    }
                }
                tc::fence_before_sync();
                tc::fence_proxy_async();
                __syncwarp();
                BWD_TICK(7);   // pass 3
                if (lane == 0) tc::mbar_arrive(&ds_full);
                // ---- dQ = dQ~ * scale / rowsum
                tc::mbar_wait(&dq_full, ph);
                tc::fence_after_sync();
                BWD_TICK(8);   // wait dQ
                {
                    float o[32];
                    tc::tmem_ld32(t_lane + ACC_COL + h * 32, o);
                    if (valid) {
                        const float f = a.scale / l;
                        __nv_bfloat16* dst = a.dq + (seq * a.q_sa + qr) * a.lddq + head * HD + h * 32;
#pragma unroll
                        for (int j = 0; j < 32; j += 8)
                            *reinterpret_cast<uint4*>(dst + j) = make_uint4(hv_pack_bf16(o[j] * f, o[j + 1] * f), hv_pack_bf16(o[j + 2] * f, o[j + 3] * f),
                                                                            hv_pack_bf16(o[j + 4] * f, o[j + 5] * f), hv_pack_bf16(o[j + 6] * f, o[j + 7] * f));
                    }
                }
                tc::fence_before_sync();
                BWD_TICK(9);   // epilogue
                if (prof) a.dbg[(KERNEL_ID * 148 + (blockIdx.x % 148)) * 16 + 15] += 1;
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

// ------------------------------------------------------------------------------------------------ dK, dV
template <bool PIPE>
__global__ void __launch_bounds__(THREADS, 1)
attn_bwd_dkv_tc_kernel(const __grid_constant__ CUtensorMap tq, const __grid_constant__ CUtensorMap tk,
                       const __grid_constant__ CUtensorMap tv, const __grid_constant__ CUtensorMap tdo, const BwdTc a) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t qdo_full, qdo_free, kt_full, kt_free, st_full, p_full, dpt_full, dv_done, ds_full, acc_full;
    __shared__ uint32_t tmem_slot;
    __shared__ float sLse[LMAX], sD[LMAX];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int QP = a.QP, QP32 = (QP + 31) & ~31;
    const int nk = (a.Lkv + QT - 1) / QT;
    const int nb = QP >> 4, HB = ((nb + 1) >> 1) * 16;  // query half 0 = [0, HB), half 1 = [HB, QP)
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* Qs = smem;                    // [QP32 x 64]
    uint8_t* dOs = Qs + QP32 * 128;        // [QP32 x 64]
    uint8_t* Kt = dOs + QP32 * 128;        // [128 x 64]
    uint8_t* Vt = Kt + TBYTES;             // [128 x 64]
    uint8_t* Ps = Vt + TBYTES;             // nchunks x [128 x 64]: P^T, then dS^T

    if (threadIdx.x == 0) {
        tc::prefetch_tmap(&tq); tc::prefetch_tmap(&tk); tc::prefetch_tmap(&tv); tc::prefetch_tmap(&tdo);
        tc::mbar_init(&qdo_full, 1); tc::mbar_init(&qdo_free, 1); tc::mbar_init(&kt_full, 1); tc::mbar_init(&kt_free, 1);
        tc::mbar_init(&st_full, 1); tc::mbar_init(&p_full, 8); tc::mbar_init(&dpt_full, 1); tc::mbar_init(&dv_done, 1);
        tc::mbar_init(&ds_full, 8); tc::mbar_init(&acc_full, 1);   // compute warps arrive once per warp
        tc::fence_barrier_init();
    }
    if (warp == 8) tc::tmem_alloc(&tmem_slot, TMEM_COLS);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem_base = tmem_slot;

    if (warp == 9) {
        // ===================================================== TMA producer
        if (lane == 0) {
            long long T = 0, it = 0;
            for (long long w = blockIdx.x; w < a.items; w += gridDim.x, ++it) {
                const long long seq = w / a.heads;
                const int col = (int)(w - seq * a.heads) * HD;
                const long long qrow0 = seq * a.q_sa, kvrow0 = seq * a.kv_sa;
                tc::mbar_wait(&qdo_free, (uint32_t)((it & 1) ^ 1));
                tc::mbar_expect_tx(&qdo_full, 2 * QP32 * 128);
                for (int j = 0; j < QP32 / 32; ++j) {
                    tc::tma_load_2d(Qs + j * 4096, &tq, &qdo_full, col, (int)(qrow0 + j * 32));
                    tc::tma_load_2d(dOs + j * 4096, &tdo, &qdo_full, col, (int)(qrow0 + j * 32));
                }
                for (int t = 0; t < nk; ++t, ++T) {
                    tc::mbar_wait(&kt_free, (uint32_t)((T & 1) ^ 1));
                    tc::mbar_expect_tx(&kt_full, 2 * TBYTES);
                    for (int j = 0; j < 4; ++j) {
                        tc::tma_load_2d(Kt + j * 4096, &tk, &kt_full, col, (int)(kvrow0 + t * QT + j * 32));
                        tc::tma_load_2d(Vt + j * 4096, &tv, &kt_full, col, (int)(kvrow0 + t * QT + j * 32));
                    }
                }
            }
        }
    } else if (warp == 8) {
        // ===================================================== MMA issuer
        if (lane == 0) {
            long long T = 0, it = 0;
            for (long long w = blockIdx.x; w < a.items; w += gridDim.x, ++it) {
                tc::mbar_wait(&qdo_full, (uint32_t)(it & 1));
                for (int t = 0; t < nk; ++t, ++T) {
                    const uint32_t ph = (uint32_t)(T & 1);
                    tc::mbar_wait(&kt_full, ph);
                    tc::fence_after_sync();
                    issue_s_like(tmem_base + S_COL, Kt, Qs, QP, HB);        // S^T = K_t Q^T
                    tc::umma_commit(&st_full);
                    tc::mbar_wait(&p_full, ph);                             // P^T in registers + staged; accumulators of the previous tile read back
                    tc::fence_after_sync();
                    issue_s_like(tmem_base + S_COL, Vt, dOs, QP, HB);       // dP^T = V_t dO^T over the S^T columns
                    tc::umma_commit(&dpt_full);
                    tc::umma_commit(&kt_free);
                    issue_pv_like(tmem_base + ACC_COL, Ps, dOs, QP);        // dV_t = P^T dO
                    tc::umma_commit(&dv_done);
                    tc::mbar_wait(&ds_full, ph);                            // dS^T staged (after dv_done)
                    tc::fence_after_sync();
                    issue_pv_like(tmem_base + ACC_COL + HD, Ps, Qs, QP);    // dK_t = dS^T Q
                    tc::umma_commit(&acc_full);
                    if (t == nk - 1) tc::umma_commit(&qdo_free);
                }
            }
        }
    } else {
        // ===================================================== compute warps: thread = (key row, query half)
        const int q = warp & 3, h = warp >> 2, row = q * 32 + lane;
        const int c_begin = h ? HB : 0, c_end = h ? QP : HB;
        const uint32_t t_lane = tmem_base + ((uint32_t)(q * 32) << 16);
        const uint32_t s_addr = t_lane + S_COL;
        long long T = 0;
        for (long long w = blockIdx.x; w < a.items; w += gridDim.x) {
            const long long seq = w / a.heads;
            const int head = (int)(w - seq * a.heads);
            // per-query statistics of the dq kernel -> shared memory (the previous item's readers are past their last use:
            // every thread went through the acc_full wait of its last tile after reading them)
            tc::named_bar_sync(1, 256);
            for (int i = threadIdx.x; i < QP; i += 256) {
                const float2 s = i < a.Lq ? reinterpret_cast<const float2*>(a.stats)[w * a.Lq + i] : make_float2(0.f, 0.f);
                sLse[i] = s.x; sD[i] = s.y;
            }
            tc::named_bar_sync(1, 256);
            for (int t = 0; t < nk; ++t, ++T) {
                const uint32_t ph = (uint32_t)(T & 1);
                const int kr = t * QT + row;
                const bool valid = kr < a.Lkv;
                constexpr int KERNEL_ID = 1;
                const bool prof = a.dbg != nullptr && threadIdx.x == 0;
                long long tk0 = prof ? clock64() : 0;
                tc::mbar_wait(&st_full, ph);
                tc::fence_after_sync();
                BWD_TICK(1);   // wait S^T
                // ---- P^T = exp2(S^T c - lse_q) -> registers (packed bf16) and the staging buffer
                uint32_t preg[72];
                {
                    uint32_t buf[2][16];
                    if (PIPE) ld16_issue(s_addr + c_begin, buf[0]);
#pragma unroll
                    for (int k = 0; k < 9; ++k) {
                        const int c = c_begin + 16 * k;
                        if (c < c_end) {  // warp-uniform
                            if (!PIPE) ld16_issue(s_addr + c, buf[k & 1]);
                            ld16_wait(buf[k & 1]);
                            if (PIPE && c + 16 < c_end) ld16_issue(s_addr + c + 16, buf[(k + 1) & 1]);
                            const uint32_t (&va)[16] = buf[k & 1];
                            float p[16];
#pragma unroll
                            for (int j = 0; j < 16; ++j)
                                p[j] = (c + j < a.Lq) ? tc::ex2_approx(fmaf(__uint_as_float(va[j]), a.sl2, -sLse[c + j])) : 0.f;
#pragma unroll
                            for (int j = 0; j < 8; ++j) preg[8 * k + j] = hv_pack_bf16(p[2 * j], p[2 * j + 1]);
                            stage16(Ps, row, c, p);
                        }
                    }
                }
                BWD_TICK(4);   // pass 2
                tc::fence_before_sync();
                tc::fence_proxy_async();
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(&p_full);
                // ---- dS^T = P^T (dP^T - D_q) scale -> the same staging buffer, once the dV MMA has read P^T
                tc::mbar_wait(&dpt_full, ph);
                tc::mbar_wait(&dv_done, ph);
                tc::fence_after_sync();
                BWD_TICK(6);   // wait dP^T + dV
                {
                    uint32_t buf[2][16];
                    if (PIPE) ld16_issue(s_addr + c_begin, buf[0]);
#pragma unroll
                    for (int k = 0; k < 9; ++k) {
                        const int c = c_begin + 16 * k;
                        if (c < c_end) {
                            if (!PIPE) ld16_issue(s_addr + c, buf[k & 1]);
                            ld16_wait(buf[k & 1]);
                            if (PIPE && c + 16 < c_end) ld16_issue(s_addr + c + 16, buf[(k + 1) & 1]);
                            const uint32_t (&va)[16] = buf[k & 1];
                            float ds[16];
#pragma unroll
                            for (int j = 0; j < 8; ++j) {
                                ds[2 * j] = bf_lo(preg[8 * k + j]) * (__uint_as_float(va[2 * j]) - sD[c + 2 * j]) * a.scale;
                                ds[2 * j + 1] = bf_hi(preg[8 * k + j]) * (__uint_as_float(va[2 * j + 1]) - sD[c + 2 * j + 1]) * a.scale;
                            }
                            stage16(Ps, row, c, ds);
                        }
                    }
                }
                tc::fence_before_sync();
                tc::fence_proxy_async();
                __syncwarp();
                BWD_TICK(7);   // pass 3
                if (lane == 0) tc::mbar_arrive(&ds_full);
                // ---- dV_t, dK_t -> global
                tc::mbar_wait(&acc_full, ph);
                tc::fence_after_sync();
                BWD_TICK(8);   // wait dK
                {
                    float ov[32], ok[32];
                    tc::tmem_ld32(t_lane + ACC_COL + h * 32, ov);
                    tc::tmem_ld32(t_lane + ACC_COL + HD + h * 32, ok);
                    if (valid) {
                        __nv_bfloat16* pv = a.dv + (seq * a.kv_sa + kr) * a.lddv + head * HD + h * 32;
                        __nv_bfloat16* pk = a.dk + (seq * a.kv_sa + kr) * a.lddk + head * HD + h * 32;
#pragma unroll
                        for (int j = 0; j < 32; j += 8) {
                            *reinterpret_cast<uint4*>(pv + j) = make_uint4(hv_pack_bf16(ov[j], ov[j + 1]), hv_pack_bf16(ov[j + 2], ov[j + 3]),
                                                                           hv_pack_bf16(ov[j + 4], ov[j + 5]), hv_pack_bf16(ov[j + 6], ov[j + 7]));
                            *reinterpret_cast<uint4*>(pk + j) = make_uint4(hv_pack_bf16(ok[j], ok[j + 1]), hv_pack_bf16(ok[j + 2], ok[j + 3]),
                                                                           hv_pack_bf16(ok[j + 4], ok[j + 5]), hv_pack_bf16(ok[j + 6], ok[j + 7]));
                        }
                    }
                }
                tc::fence_before_sync();
                BWD_TICK(9);   // epilogue
                if (prof) a.dbg[(KERNEL_ID * 148 + (blockIdx.x % 148)) * 16 + 15] += 1;
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

unsigned long long* g_bwd_prof = nullptr;
unsigned long long* bwd_prof_buffer() {
    static const bool on = [] { const char* e = getenv("HV_ATTN_BWD_PROF"); return e && atoi(e) != 0; }();
    if (on && !g_bwd_prof && cudaMalloc(&g_bwd_prof, 2 * 148 * 16 * sizeof(unsigned long long)) == cudaSuccess)
        cudaMemset(g_bwd_prof, 0, 2 * 148 * 16 * sizeof(unsigned long long));
    return on ? g_bwd_prof : nullptr;
}

}  // namespace

// Diagnostics (HV_ATTN_BWD_PROF=1): copies the phase cycle counters of the tcgen05 attention-backward kernels to `host`
// (2 kernels x 148 CTAs x 16 u64: phases 0..9 as tagged in the kernels, [15] = tiles) and clears them; 0 when profiling is off
extern "C" int hv_attn_bwd_prof_read(unsigned long long* host) {
    if (!g_bwd_prof) return 0;
    cudaDeviceSynchronize();
    cudaMemcpy(host, g_bwd_prof, 2 * 148 * 16 * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
    cudaMemset(g_bwd_prof, 0, 2 * 148 * 16 * sizeof(unsigned long long));
    return 1;
}

// returns 1 when the shape is handled here (both kernels launched), 0 when the caller should use the warp-level kernels,
// < 0 on error.  Arguments as hv_attention_bwd (attn_bwd.cu); HV_ATTN_BWD_TC=0 in the environment disables this path.
int hv_attention_bwd_tc_bf16(const void* q, long long ldq, const void* k, long long ldk, const void* v, long long ldv,
                             const void* o, long long ldo, const void* dout, long long lddo, void* dq, long long lddq,
                             void* dk, long long lddk, void* dv, long long lddv, long long seqs, int heads, int Lq, int Lkv,
                             const HvSeqMap& qm, const HvSeqMap& km, float scale, float* stats, cudaStream_t st) {
    static const bool enabled = [] { const char* e = getenv("HV_ATTN_BWD_TC"); return !(e && atoi(e) == 0); }();
    if (!enabled) return 0;
    const bool dense = qm.sdiv == 1 && qm.sc == 1 && km.sdiv == 1 && km.sc == 1;
    if (!dense || Lq < 16 || Lkv < 16 || Lq > LMAX || Lkv > LMAX) return 0;
    if ((ldq % 8) || (ldk % 8) || (ldv % 8) || (ldo % 8) || (lddo % 8) || (lddq % 8) || (lddk % 8) || (lddv % 8)) return 0;
    auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
    if (!al16(q) || !al16(k) || !al16(v) || !al16(o) || !al16(dout) || !al16(dq) || !al16(dk) || !al16(dv)) return 0;
    const long long qrows = (seqs - 1) * qm.sa + Lq, kvrows = (seqs - 1) * km.sa + Lkv;
    if (qrows >= (1ll << 31) || kvrows >= (1ll << 31)) return 0;
    CUtensorMap tq, tk, tv, tdo;
    int rc;
    const int cols = heads * HD;
    if ((rc = hv_make_tmap_bf16(q, qrows, cols, ldq, 32, &tq)) != HV_OK) return rc;
    if ((rc = hv_make_tmap_bf16(k, kvrows, cols, ldk, 32, &tk)) != HV_OK) return rc;
    if ((rc = hv_make_tmap_bf16(v, kvrows, cols, ldv, 32, &tv)) != HV_OK) return rc;
    if ((rc = hv_make_tmap_bf16(dout, qrows, cols, lddo, 32, &tdo)) != HV_OK) return rc;
    BwdTc a;
    a.o = reinterpret_cast<const __nv_bfloat16*>(o); a.dout = reinterpret_cast<const __nv_bfloat16*>(dout);
    a.dq = reinterpret_cast<__nv_bfloat16*>(dq); a.dk = reinterpret_cast<__nv_bfloat16*>(dk); a.dv = reinterpret_cast<__nv_bfloat16*>(dv);
    a.ldo = ldo; a.lddo = lddo; a.lddq = lddq; a.lddk = lddk; a.lddv = lddv; a.q_sa = qm.sa; a.kv_sa = km.sa;
    a.heads = heads; a.Lq = Lq; a.Lkv = Lkv; a.QP = (Lq + 15) & ~15; a.KP = (Lkv + 15) & ~15;
    a.items = seqs * heads;
    a.sl2 = scale * 1.4426950408889634f; a.scale = scale;
    a.stats = stats;
    a.dbg = bwd_prof_buffer();
    const int KP32 = (a.KP + 31) & ~31, QP32 = (a.QP + 31) & ~31;
    const int smem_dq = 2 * TBYTES + 2 * KP32 * 128 + ((a.KP + 63) >> 6) * TBYTES + 1024;
    const int smem_dkv = 2 * QP32 * 128 + 2 * TBYTES + ((a.QP + 63) >> 6) * TBYTES + 1024;
    static const int pipe = [] { const char* e = getenv("HV_ATTN_BWD_PIPE"); return e ? atoi(e) : 0; }();
    static int conf_dq = 0, conf_dkv = 0;
    if (conf_dq < smem_dq) {
        HV_CUDA(cudaFuncSetAttribute(attn_bwd_dq_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_dq));
        HV_CUDA(cudaFuncSetAttribute(attn_bwd_dq_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_dq));
        conf_dq = smem_dq;
    }
    if (conf_dkv < smem_dkv) {
        HV_CUDA(cudaFuncSetAttribute(attn_bwd_dkv_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_dkv));
        HV_CUDA(cudaFuncSetAttribute(attn_bwd_dkv_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_dkv));
        conf_dkv = smem_dkv;
    }
    const int sms = hv_num_sms();
    const int grid = (int)(a.items < sms ? a.items : sms);
    if (pipe & 1) attn_bwd_dq_tc_kernel<true><<<grid, THREADS, smem_dq, st>>>(tq, tk, tv, tdo, a);
    else attn_bwd_dq_tc_kernel<false><<<grid, THREADS, smem_dq, st>>>(tq, tk, tv, tdo, a);
    HV_LAUNCH_CHECK();
    if (pipe & 2) attn_bwd_dkv_tc_kernel<true><<<grid, THREADS, smem_dkv, st>>>(tq, tk, tv, tdo, a);
    else attn_bwd_dkv_tc_kernel<false><<<grid, THREADS, smem_dkv, st>>>(tq, tk, tv, tdo, a);
    HV_LAUNCH_CHECK();
    return 1;
}
