// tcgen05 / TMEM attention for the spatial sequences of the DiT and the object encoder
// (head_dim 64, bf16, no mask, 16 <= Lkv <= 288, rows of a sequence contiguous):
//   softmax(Q K^T * scale) V   per (sequence, head)
// (F.scaled_dot_product_attention inside diffusers AttnProcessor2_0; call sites
//  model/modules.py:699-707, 1043-1051 of the reference).
//
// The whole key range of a sequence (<= 288 keys) fits in TMEM, so there is no online-softmax
// rescaling.  Per 128-query tile T:
//   S[128 x KP]  = Q K^T           tcgen05.mma, A = Q tile, B = K (both K-major, SWIZZLE_128B via TMA),
//                                  issued as two key halves H0 / H1 into a ring of THREE half-buffers
//   P            = exp2(S*c - max) 8 softmax warps: thread (row, h) owns key half h of its query row,
//                                  bf16 P written to shared memory in the K-major SWIZZLE_128B layout
//   O[128 x 64]  = P V             tcgen05.mma, A = P (smem), B = V as an MN-major operand (V is stored
//                                  [key][dim], exactly what TMA delivers; no transpose pass)
//   out          = O / rowsum      read back with tcgen05.ld, scaled, stored as bf16
// Software pipeline (the point of the three half-buffers): while the softmax warps work on tile T the
// tensor core already holds S(T+1); the epilogue of tile T-1 runs between the max pass and the exp pass
// of tile T, when PV(T-1) has long retired.  The softmax warps therefore never wait for an MMA in steady
// state.  Measured bound (tools/probes, clock64 per phase): TMEM read bandwidth.  tcgen05.ld delivers
// ~64 B/clk/SM, the two passes (max, exp) each read the 128 x 288 fp32 tile (147 KB) -> ~4600 of the
// ~7000 clk a tile takes; MUFU.EX2 (2300 clk/tile) and the MMAs (1150 clk/tile) hide under it.  A variant
// that keeps the scores in registers between the passes (16 softmax warps, 72 scores per thread) read
// TMEM once but lost more to register spills and barrier cost than it gained (299 vs 188 us at the DiT
// shape) and was not kept.  Neither was a later one-pass variant with EIGHT softmax warps that holds all 144 scores
// of a thread in registers (three warpgroups, setmaxnreg 216 / 56 so that nothing spills): 189 us against 169 us
// for this two-pass kernel on the same box, i.e. halving the TMEM reads does not pay — the per-tile time
// (~7000 clk) is dominated by dependent-latency chains of the two softmax warps per SM sub-partition, which is
// what the next redesign has to attack (more independent rows in flight per warp, not fewer TMEM reads).
// Measured phase split (HV_ATTN_PROF=1, tools/attn_phases.py, L = 274, cycles per tile averaged over full and tail
// tiles): wait S 215-307, max pass ~1600, max exchange 600-750, O epilogue ~1250 (includes the o_full wait), exp pass
// 3230-3390, publish P + sums ~480; total ~7550.  The MUFU floor of the exp pass is 2304 clk per full tile, the
// max pass moves 147 KB of TMEM at ~92 B/clk/SM (8 warps x one 2 KB tcgen05.ld in flight / ~170 clk).
// Tail tiles (<= 32 query rows left, e.g. the third tile of a 274-token sequence):
// the producer replicates the 32-row Q box into all four lane quarters, so every softmax warp can reach the
// rows and the keys are split eight ways instead of two (a third of a full tile's softmax time).
//
// Warp roles (320 threads): warps 0..7 softmax + epilogue (TMEM lane quarter = warp % 4, key half =
// warp / 4), warp 8 = MMA issuer + TMEM owner, warp 9 = TMA producer (one elected lane each).
// CTAs are persistent over (sequence, head) pairs; K is double-buffered across pairs.
#include <cuda.h>
#include <cstdlib>
#include "attn.cuh"
#include "tc_ptx.cuh"

int hv_make_tmap_bf16(const void* ptr, long long rows, int K, long long ld, int box_rows, CUtensorMap* out);  // gemm_tc.cu

namespace {

constexpr int QT = 128, HD = 64, KMAX = 288, THREADS = 320;
constexpr int O_COL = 0, S_COL = 64, SBUF = 144;  // TMEM columns: O [0,64), three S half-buffers of 144 columns
constexpr uint32_t TMEM_COLS = 512;
constexpr int QBYTES = QT * 128;

struct TcArgs {
    __nv_bfloat16* out;
    long long ldo, q_sa, kv_sa;
    int heads, Lq, Lkv, KP;  // KP = Lkv rounded up to 16
    long long items;         // seqs * heads
    float sl2;               // scale * log2(e)
    unsigned long long* dbg; // phase cycle counters (HV_ATTN_PROF=1, tools/attn_phases.py) or nullptr
};

__device__ __forceinline__ void tmem_ld16_issue(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}
// wait for the outstanding tcgen05.ld; the registers are in/out operands so that no use of them can be
// scheduled above the wait
__device__ __forceinline__ void tmem_ld16_wait(uint32_t (&r)[16]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                   "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15])
                 :
                 : "memory");
}

template <bool PROF>  // PROF: phase cycle counters (a separate instantiation, the production kernel carries none of it)
__global__ void __launch_bounds__(THREADS, 1)
attn_tc_kernel(const __grid_constant__ CUtensorMap tq, const __grid_constant__ CUtensorMap tk,
               const __grid_constant__ CUtensorMap tv, const TcArgs a) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t q_full[2], q_free[2], k_full[2], k_free[2], v_full, v_free, s_full[2][2], p_full, o_full;
    __shared__ uint32_t tmem_slot;
    __shared__ float smax[2][QT], ssum[2][QT];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int KP = a.KP, KP32 = (KP + 31) & ~31;
    const int nq = (a.Lq + QT - 1) / QT;
    const int nb = KP >> 4, hb = (nb + 1) >> 1, HB = hb * 16;  // key half 0 = [0, HB), half 1 = [HB, KP)
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* Qs = smem;                       // 2 x [128 x 64] bf16
    uint8_t* Ks = Qs + 2 * QBYTES;            // 2 x [KP32 x 64]
    uint8_t* Vs = Ks + 2 * KP32 * 128;        // [KP32 x 64]
    uint8_t* Ps = Vs + KP32 * 128;            // nchunks x [128 x 64]
    const long long n_items = a.items > blockIdx.x ? (a.items - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const long long total = n_items * nq;     // query tiles of this CTA

    if (threadIdx.x == 0) {
        tc::prefetch_tmap(&tq); tc::prefetch_tmap(&tk); tc::prefetch_tmap(&tv);
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            tc::mbar_init(&q_full[i], 1); tc::mbar_init(&q_free[i], 1);
            tc::mbar_init(&k_full[i], 1); tc::mbar_init(&k_free[i], 1);
            tc::mbar_init(&s_full[i][0], 1); tc::mbar_init(&s_full[i][1], 1);
        }
        tc::mbar_init(&v_full, 1); tc::mbar_init(&v_free, 1);
        tc::mbar_init(&p_full, 256); tc::mbar_init(&o_full, 1);
        tc::fence_barrier_init();
    }
    if (warp == 8) tc::tmem_alloc(&tmem_slot, TMEM_COLS);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem_base = tc::warp_uniform(tmem_slot);  // uniform register: operand of warp-converged MMA issue

    if (warp == 9) {
        // ===================================================== TMA producer (all 32 lanes, one elected lane issues: tc_ptx.cuh *_w)
        {
            long long T = 0, it = 0;
            auto load_q = [&](long long qrow0, int col, int i) {
                const int b = (int)(T & 1);
                tc::mbar_wait(&q_free[b], (uint32_t)(((T >> 1) & 1) ^ 1));
                tc::mbar_expect_tx_w(&q_full[b], QBYTES);
                const bool tail = a.Lq - i * QT <= 32;  // tail tile: the same 32-row box goes to all four lane quarters
#pragma unroll
                for (int j = 0; j < QT / 32; ++j)
                    tc::tma_load_2d_w(tc::smem_u32(Qs + b * QBYTES + j * 4096), &tq, &q_full[b], col, (int)(qrow0 + i * QT + (tail ? 0 : j * 32)));
                ++T;
            };
            for (long long w = blockIdx.x; w < a.items; w += gridDim.x, ++it) {
                const long long seq = w / a.heads;
                const int col = (int)(w - seq * a.heads) * HD;
                const long long qrow0 = seq * a.q_sa, kvrow0 = seq * a.kv_sa;
                const int kb = (int)(it & 1);
                tc::mbar_wait(&k_free[kb], (uint32_t)(((it >> 1) & 1) ^ 1));
                tc::mbar_expect_tx_w(&k_full[kb], KP32 * 128);
                for (int j = 0; j < KP32 / 32; ++j)
                    tc::tma_load_2d_w(tc::smem_u32(Ks + kb * KP32 * 128 + j * 4096), &tk, &k_full[kb], col, (int)(kvrow0 + j * 32));
                load_q(qrow0, col, 0);
                if (nq > 1) load_q(qrow0, col, 1);
                tc::mbar_wait(&v_free, (uint32_t)((it & 1) ^ 1));
                tc::mbar_expect_tx_w(&v_full, KP32 * 128);
                for (int j = 0; j < KP32 / 32; ++j)
                    tc::tma_load_2d_w(tc::smem_u32(Vs + j * 4096), &tv, &v_full, col, (int)(kvrow0 + j * 32));
                for (int i = 2; i < nq; ++i) load_q(qrow0, col, i);
            }
        }
    } else if (warp == 8) {
        // ===================================================== MMA issuer
        if (total > 0) {  // all 32 lanes with warp-uniform values, one elected lane issues (tc_ptx.cuh *_w)
            auto issue_s = [&](long long T, int h) {  // S(T), key half h -> ring buffer (2T + h) % 3
                const long long it = T / nq;
                const int ncols = h ? KP - HB : HB;
                if (ncols > 0) {
                    const uint64_t da = tc::make_smem_desc_k128(tc::smem_u32(Qs + (T & 1) * QBYTES));
                    const uint64_t db = tc::make_smem_desc_k128(tc::smem_u32(Ks + (it & 1) * KP32 * 128 + h * HB * 128));
                    const uint32_t idesc = tc::make_idesc_bf16(QT, ncols);
                    const uint32_t d = tmem_base + S_COL + (uint32_t)((2 * T + h) % 3) * SBUF;
#pragma unroll
                    for (int k = 0; k < HD / 16; ++k) tc::umma_bf16_w(d, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, k != 0);
                }
                tc::umma_commit_w(&s_full[T & 1][h]);  // one barrier per (tile parity, half): never more than one phase ahead of a waiter
            };
            auto issue_pv = [&]() {
                constexpr uint32_t idesc = tc::make_idesc_bf16_bmn(QT, HD);
                for (int j = 0; j < KP / 16; ++j) {
                    const uint64_t da = tc::make_smem_desc_k128(tc::smem_u32(Ps + (j >> 2) * QBYTES)) + (uint64_t)(2 * (j & 3));
                    const uint64_t db = tc::make_smem_desc_k128(tc::smem_u32(Vs + j * 2048));  // 16 keys x 128 B, MN-major
                    tc::umma_bf16_w(tmem_base + O_COL, da, db, idesc, j != 0);
                }
                tc::umma_commit_w(&o_full);
            };
            auto after_s = [&](long long T) {  // both halves of S(T) issued: its Q buffer and (last tile) its K buffer retire with them
                tc::umma_commit_w(&q_free[T & 1]);
                if (T % nq == nq - 1) tc::umma_commit_w(&k_free[(T / nq) & 1]);
            };
            tc::mbar_wait(&k_full[0], 0);
            tc::mbar_wait(&q_full[0], 0);
            tc::fence_after_sync();
            issue_s(0, 0); issue_s(0, 1); after_s(0);
            for (long long T = 0; T < total; ++T) {
                const long long it = T / nq;
                const int i = (int)(T - it * nq);
                if (T + 1 < total) {
                    const long long it1 = (T + 1) / nq;
                    if (i + 1 == nq) tc::mbar_wait(&k_full[it1 & 1], (uint32_t)((it1 >> 1) & 1));
                    tc::mbar_wait(&q_full[(T + 1) & 1], (uint32_t)(((T + 1) >> 1) & 1));
                    tc::fence_after_sync();
                    issue_s(T + 1, 0);  // ring buffer of S(T-1).H1: its exp pass ended with p_full(T-1)
                }
                tc::mbar_wait(&p_full, (uint32_t)(T & 1));  // P(T) written; O(T-1) read back
                if (i == 0) tc::mbar_wait(&v_full, (uint32_t)(it & 1));
                tc::fence_after_sync();
                // S(T+1).H1 (ring buffer of S(T).H0) goes ahead of PV(T): the softmax warps need it first, PV(T)
                // is only read back after their next max pass
                if (T + 1 < total) { issue_s(T + 1, 1); after_s(T + 1); }
                issue_pv();
                if (i == nq - 1) tc::umma_commit_w(&v_free);
            }
        }
    } else {
        // ===================================================== softmax + epilogue warps
        const int q = warp & 3, h = warp >> 2, row = q * 32 + lane;
        const int c_begin = h ? HB : 0, c_end = h ? KP : HB;
        const uint32_t t_lane = tmem_base + ((uint32_t)(q * 32) << 16);
        uint8_t* prow_full = Ps + (row >> 3) * 1024 + (row & 7) * 128;  // this row inside every 64-key P chunk
        uint8_t* prow_tail = Ps + (lane >> 3) * 1024 + (lane & 7) * 128;  // tail tiles: row = lane in every quarter
        long long T = 0;
        // deferred epilogue of the previous tile
        bool pend = false;
        __nv_bfloat16* pend_ptr = nullptr;
        float pend_inv = 0.f;
        bool pend_store = false, pend_active = false;
        auto epilogue = [&]() {
            tc::mbar_wait(&o_full, (uint32_t)((T - 1) & 1));
            tc::fence_after_sync();
            if (pend_active) {  // warp-uniform (tcgen05.ld is .sync.aligned); only the store is per-row
                float o[32];
                tc::tmem_ld32(t_lane + O_COL + h * 32, o);
                const float inv = pend_inv;
                if (pend_store) {
#pragma unroll
                    for (int j = 0; j < 32; j += 8)
                        *reinterpret_cast<uint4*>(pend_ptr + j) =
                            make_uint4(hv_pack_bf16(o[j] * inv, o[j + 1] * inv), hv_pack_bf16(o[j + 2] * inv, o[j + 3] * inv),
                                       hv_pack_bf16(o[j + 4] * inv, o[j + 5] * inv), hv_pack_bf16(o[j + 6] * inv, o[j + 7] * inv));
                }
            }
            tc::fence_before_sync();
        };
        for (long long w = blockIdx.x; w < a.items; w += gridDim.x) {
            const long long seq = w / a.heads;
            const int head = (int)(w - seq * a.heads);
            for (int i = 0; i < nq; ++i, ++T) {
                const bool active = i * QT + q * 32 < a.Lq;  // warp-uniform: this warp has query rows in the tile
                // tail tile (<= 32 query rows left): the producer replicated the rows into all four lane quarters, so
                // the eight softmax warps split the KEYS eight ways (16-key chunks k = q, q+4, ... of key half h)
                // instead of six of them idling for a whole tile
                const bool tail = a.Lq - i * QT <= 32;
                const int nch = (c_end - c_begin) >> 4;
                uint8_t* prow = tail ? prow_tail : prow_full;
                const uint32_t s_addr = t_lane + S_COL + (uint32_t)((2 * T + h) % 3) * SBUF - (uint32_t)c_begin;
                const bool prof = PROF && a.dbg != nullptr && lane == 0 && q == 0;
                long long tk0 = 0, tk1 = 0;
                auto tick = [&](int phase) {  // cycles since the previous tick -> counter `phase` of (CTA, key half)
                    if (PROF && prof) { tk1 = clock64(); if (phase >= 0) a.dbg[(blockIdx.x * 2 + h) * 8 + phase] += (unsigned long long)(tk1 - tk0); tk0 = tk1; }
                };
                tick(-1);
                tc::mbar_wait(&s_full[T & 1][h], (uint32_t)((T >> 1) & 1));
                tc::fence_after_sync();
                tick(0);
                // ---- pass 1: row maximum over this thread's half of the keys
                float mx = -INFINITY;
                if (tail) {
                    uint32_t va[16];
                    for (int k = q; k < nch; k += 4) {
                        const int c = c_begin + (k << 4);
                        tmem_ld16_issue(s_addr + c, va);
                        tmem_ld16_wait(va);
#pragma unroll
                        for (int j = 0; j < 16; ++j) if (c + j < a.Lkv) mx = fmaxf(mx, __uint_as_float(va[j]));
                    }
                } else if (active && c_begin < c_end) {
                    uint32_t va[16], vb[16];
                    float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
                    int c = c_begin;
                    tmem_ld16_issue(s_addr + c, va);
                    while (true) {
                        tmem_ld16_wait(va);
                        bool more = c + 16 < c_end;
                        if (more) tmem_ld16_issue(s_addr + c + 16, vb);
                        if (c + 16 <= a.Lkv) {
#pragma unroll
                            for (int j = 0; j < 4; ++j)  // four independent chains instead of one 16-long dependent one
                                mx4[j] = fmaxf(fmaxf(mx4[j], __uint_as_float(va[j])), fmaxf(fmaxf(__uint_as_float(va[j + 4]), __uint_as_float(va[j + 8])), __uint_as_float(va[j + 12])));
                        } else {
#pragma unroll
                            for (int j = 0; j < 16; ++j) if (c + j < a.Lkv) mx = fmaxf(mx, __uint_as_float(va[j]));
                        }
                        if (!more) break;
                        c += 16;
                        tmem_ld16_wait(vb);
                        more = c + 16 < c_end;
                        if (more) tmem_ld16_issue(s_addr + c + 16, va);
                        if (c + 16 <= a.Lkv) {
#pragma unroll
                            for (int j = 0; j < 4; ++j)  // four independent chains instead of one 16-long dependent one
                                mx4[j] = fmaxf(fmaxf(mx4[j], __uint_as_float(vb[j])), fmaxf(fmaxf(__uint_as_float(vb[j + 4]), __uint_as_float(vb[j + 8])), __uint_as_float(vb[j + 12])));
                        } else {
#pragma unroll
                            for (int j = 0; j < 16; ++j) if (c + j < a.Lkv) mx = fmaxf(mx, __uint_as_float(vb[j]));
                        }
                        if (!more) break;
                        c += 16;
                    }
                    mx = fmaxf(fmaxf(mx, mx4[0]), fmaxf(fmaxf(mx4[1], mx4[2]), mx4[3]));
                }
                smax[h][row] = mx;
                tick(1);
                tc::named_bar_sync(1, 256);
                if (tail) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) mx = fmaxf(mx, fmaxf(smax[0][j * 32 + lane], smax[1][j * 32 + lane]));
                } else {
                    mx = fmaxf(mx, smax[h ^ 1][row]);
                }
                const float msc = mx * a.sl2;
                tick(2);
                // ---- epilogue of the previous tile.  Its wait on o_full is ALSO what protects the single P buffer in shared
                // memory: PV(T-1) must have finished reading P(T-1) before the exp pass below overwrites it, so this
                // cannot move behind the exp pass (tried: 166 us instead of 170, but P(T) then races with PV(T-1)).
                if (pend) epilogue();
                tick(3);
                // ---- pass 2: P = exp2(S*c - max*c) -> bf16 -> shared memory (K-major, 128 B swizzle)
                float sum4[4] = {0.f, 0.f, 0.f, 0.f};
                auto emit = [&](uint32_t (&r)[16], int c) {
                        float v[16];
                        if (c + 16 <= a.Lkv) {
#pragma unroll
                            for (int j = 0; j < 16; ++j) v[j] = tc::ex2_approx(fmaf(__uint_as_float(r[j]), a.sl2, -msc));
                        } else {
#pragma unroll
                            for (int j = 0; j < 16; ++j) v[j] = (c + j < a.Lkv) ? tc::ex2_approx(fmaf(__uint_as_float(r[j]), a.sl2, -msc)) : 0.f;
                        }
#pragma unroll
                        for (int j = 0; j < 4; ++j) sum4[j] += (v[j] + v[j + 4]) + (v[j + 8] + v[j + 12]);
                        uint8_t* pc = prow + (c >> 6) * QBYTES;
                        const int jc = (c & 63) >> 3;
                        *reinterpret_cast<uint4*>(pc + (((jc) ^ (row & 7)) << 4)) =
                            make_uint4(hv_pack_bf16(v[0], v[1]), hv_pack_bf16(v[2], v[3]), hv_pack_bf16(v[4], v[5]), hv_pack_bf16(v[6], v[7]));
                        *reinterpret_cast<uint4*>(pc + (((jc + 1) ^ (row & 7)) << 4)) =
                            make_uint4(hv_pack_bf16(v[8], v[9]), hv_pack_bf16(v[10], v[11]), hv_pack_bf16(v[12], v[13]), hv_pack_bf16(v[14], v[15]));
                };
                if (tail) {
                    uint32_t va[16];
                    for (int k = q; k < nch; k += 4) {
                        const int c = c_begin + (k << 4);
                        tmem_ld16_issue(s_addr + c, va);
                        tmem_ld16_wait(va);
                        emit(va, c);
                    }
                } else if (active && c_begin < c_end) {
                    uint32_t va[16], vb[16];
                    int c = c_begin;
                    tmem_ld16_issue(s_addr + c, va);
                    while (true) {
                        tmem_ld16_wait(va);
                        bool more = c + 16 < c_end;
                        if (more) tmem_ld16_issue(s_addr + c + 16, vb);
                        emit(va, c);
                        if (!more) break;
                        c += 16;
                        tmem_ld16_wait(vb);
                        more = c + 16 < c_end;
                        if (more) tmem_ld16_issue(s_addr + c + 16, va);
                        emit(vb, c);
                        if (!more) break;
                        c += 16;
                    }
                }
                float sum = (sum4[0] + sum4[1]) + (sum4[2] + sum4[3]);
                ssum[h][row] = sum;
                tick(4);
                tc::fence_before_sync();   // orders this thread's tcgen05.ld of S / O before the MMAs that overwrite them
                tc::fence_proxy_async();   // P (generic-proxy stores) -> visible to the tensor core
                tc::mbar_arrive(&p_full);
                tc::named_bar_sync(1, 256);
                tick(5);
                if (PROF && prof) a.dbg[(blockIdx.x * 2 + h) * 8 + (tail ? 7 : 6)] += 1;
                const int qr = i * QT + row;
                pend = true;
                pend_active = active;
                pend_store = qr < a.Lq;
                if (tail) {
                    sum = 0.f;
#pragma unroll
                    for (int j = 0; j < 4; ++j) sum += ssum[0][j * 32 + lane] + ssum[1][j * 32 + lane];
                } else {
                    sum += ssum[h ^ 1][row];
                }
                pend_inv = 1.f / sum;
                pend_ptr = a.out + (seq * a.q_sa + qr) * a.ldo + head * HD + h * 32;
            }
        }
        if (pend) epilogue();
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 8) {
        tc::fence_after_sync();
        tc::tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

}  // namespace

// HV_ATTN_PROF=1: per-(CTA, key half) cycle counters of the softmax phases, 8 x u64 each:
// 0 wait S, 1 max pass, 2 max exchange, 3 O epilogue, 4 exp pass, 5 publish P + sum exchange, 6 full tiles, 7 tail tiles
static unsigned long long* g_prof = nullptr;
static unsigned long long* hv_attn_prof_buffer() {
    static const bool on = [] { const char* e = getenv("HV_ATTN_PROF"); return e && atoi(e) != 0; }();
    if (on && !g_prof && cudaMalloc(&g_prof, 148 * 2 * 8 * sizeof(unsigned long long)) == cudaSuccess)
        cudaMemset(g_prof, 0, 148 * 2 * 8 * sizeof(unsigned long long));
    return on ? g_prof : nullptr;
}
// copies the counters to `host` (148 * 16 u64) and clears them; returns 0 when profiling is off
extern "C" int hv_attn_prof_read(unsigned long long* host) {
    if (!g_prof) return 0;
    cudaDeviceSynchronize();
    cudaMemcpy(host, g_prof, 148 * 2 * 8 * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
    cudaMemset(g_prof, 0, 148 * 2 * 8 * sizeof(unsigned long long));
    return 1;
}

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
    t.dbg = hv_attn_prof_buffer();
    const int KP32 = (t.KP + 31) & ~31, nchunks = (t.KP + 63) >> 6;
    const int smem = 2 * QBYTES + 3 * KP32 * 128 + nchunks * QBYTES + 1024;
    static int configured_of[HV_MAX_DEVICES] = {};  // function attributes are per device
    int& configured = configured_of[hv_device_slot()];
    if (configured < smem) {
        HV_CUDA(cudaFuncSetAttribute(attn_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        HV_CUDA(cudaFuncSetAttribute(attn_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        configured = smem;
    }
    const int sms = hv_num_sms();
    const int grid = (int)(t.items < sms ? t.items : sms);
    if (t.dbg) attn_tc_kernel<true><<<grid, THREADS, smem, st>>>(tq, tk, tv, t);
    else attn_tc_kernel<false><<<grid, THREADS, smem, st>>>(tq, tk, tv, t);
    HV_LAUNCH_CHECK();
    return 1;
}
