// tcgen05 / TMEM attention, second design: ONE pass over S, P kept in tensor memory, no barrier between softmax warps.
// Same contract as attn_tc.cu (head_dim 64, bf16, no mask, 16 <= Lkv <= 288, rows of a sequence contiguous):
//   softmax(Q K^T * scale) V   per (sequence, head)
// (F.scaled_dot_product_attention inside diffusers AttnProcessor2_0; call sites model/modules.py:699-707, 1043-1051).
//
// What the phase counters of the first design showed (attn_tc.cu header): of ~7550 clk per 128-query tile, 2300 went
// into a separate row-maximum pass over S plus the cross-warp exchange of the maxima, 480 into publishing P through
// shared memory, and everything ran in lockstep behind two named barriers per tile.  This design removes all three:
//   * no maximum pass.  softmax is invariant to the constant that is subtracted, and the whole key range is in TMEM
//     at once (no online rescaling), so ANY m_i >= max_j s_ij that does not underflow works.  An extra warp computes
//     the Cauchy-Schwarz bound m_i = |q_i| * max_j |k_j| * scale from the Q and K tiles in shared memory while the
//     tensor core computes S.  exp2 arguments then lie in [-2 m_i, 0]; as long as m_i * log2(e) <= 56 nothing
//     underflows (fp32 and bf16 share the exponent range), relative precision is that of the exact-maximum version.
//     Tiles whose bound is larger (never with LayerNorm'ed q / k) take the exact path: maximum pass + exchange.
//   * P never touches shared memory: every softmax thread overwrites its S row in place with packed bf16
//     (tcgen05.st) and P V is a tcgen05.mma with the A operand in TMEM.
//   * no named barrier on the fast path: a thread owns (query row, key half); the row sums of the two halves meet in
//     shared memory and are read in the epilogue, which sits behind the o_full mbarrier anyway.  The two warps of an
//     SM sub-partition therefore drift apart: the key-half-0 warp runs exp(T+1) while the key-half-1 warp waits for
//     P V(T) and does the epilogue of T first — MUFU and the TMEM port are busy from one warp or the other.
// TMEM: O [0,64) + a ring of three S half-buffers of 144 columns (P(T).h lives in the first 72 columns of its S).
// Tail tiles (<= 32 query rows left, e.g. the third tile of a 274-token sequence) keep the first design's scheme: the
// producer replicates the 32-row Q box into all four lane quarters, the eight softmax warps split the KEYS eight ways
// and P goes through shared memory (rows 0..31 only), P V as an SS MMA.
// Measured (tools/probes/mma_probe.cu, tmem_probe.cu): an M = 128 MMA issues every 65-88 clk whatever N <= 128 is, so the
// 18 N = 64 MMAs of P V cost as much as 18 N = 128 ones; MUFU.EX2 runs at 15.8 / clk / SM, which puts the floor of the
// exponentials of a 128 x 288 tile at 2330 clk.  Same-box A/B against attn_tc.cu (128 sequences x 16 heads,
// profiles/r02_q_attn_ab.txt): L = 256 0.077 against 0.107 ms (447 / 321 TFLOP/s), L = 274 0.117 against 0.150 ms (336 / 263).
// How it got there, every step read off the timeline of CTA 0 (HV_ATTN_TIMELINE=1, tools/attn_timeline.py, profiles/r02_[mpq]_*):
//   0.097 / 0.144  first version of this design (one pass, P in TMEM, O read back by the key-half-0 warps)
//   0.091 / 0.134  exp loop software-pipelined by hand: row sums / packing / P store of chunk i-1 inside the MUFU phase of
//                  chunk i (331 -> 275 clk per 16-key chunk; one warp per SM sub-partition runs it most of the time)
//   0.087 / 0.125  P V issued per key half as soon as that half of P is complete (two p_full barriers)
//   0.083 / 0.125  K of the next item requested BEFORE the wait for the single V buffer (it arrived ~3000 clk late)
//   0.077 / 0.117  two bounds warps + sqrt.approx: the bounds of an item's first tile are ready before its S
// What a tile costs now (L = 256, ~4300 clk): exp pass of the key-half-1 warps ~2350 (8 chunks at ~290 clk against the MUFU
// bound of 128), P V.H1 issue ~750, S(T+1).H1 issue ~400, hand-overs ~500.  Tried and rejected: forcing the program order
// MUFU group / finish part / MUFU group through value-neutral LOP3 dependencies (0.093 against 0.091 at that stage).
//
// Warp roles (384 threads): warps 0..7 softmax + epilogue (TMEM lane quarter = warp % 4, key half = warp / 4),
// warp 8 = MMA issuer + TMEM owner, warp 9 = TMA producer, warps 10 / 11 = norms / bounds (rows split between them).  Warps 8 and 9 run with all 32
// lanes converged and warp-uniform operands; one elected lane issues (tc_ptx.cuh *_w).
#include <cuda.h>
#include <cstdlib>
#include "attn.cuh"
#include "tc_ptx.cuh"

int hv_make_tmap_bf16(const void* ptr, long long rows, int K, long long ld, int box_rows, CUtensorMap* out);  // gemm_tc.cu

namespace {

constexpr int QT = 128, HD = 64, KMAX = 288, THREADS = 384;
constexpr int O_COL = 0, S_COL = 64, SBUF = 144;
constexpr uint32_t TMEM_COLS = 512;
constexpr int QBYTES = QT * 128;
constexpr float BOUND_MAX = 56.f;  // log2 units: exp2 arguments stay above -2 * 56 = -112 > -126

struct Tc2Args {
    __nv_bfloat16* out;
    long long ldo, q_sa, kv_sa;
    int heads, Lq, Lkv, KP;
    long long items;
    float sl2;
    int force_exact;  // HV_ATTN_EXACT=1: always take the maximum pass (tests)
    float* lse;       // optional [item][Lq][2] floats: element 0 <- msc + log2(sum) per query row (for the backward)
    unsigned long long* tl;  // HV_ATTN_TIMELINE=1: clock64 stamps of CTA 0, 64 tiles x 16 events (tools/attn_timeline.py)
};

__device__ __forceinline__ void ld16_issue(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}
// wait for the outstanding tcgen05.ld; the registers are in/out operands so that no use of them moves above the wait
__device__ __forceinline__ void ld16_wait(uint32_t (&r)[16]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                   "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15])
                 :
                 : "memory");
}
// UPPER BOUND of the sum of squares of row r of a [rows x 64] bf16 tile stored by TMA with SWIZZLE_128B (32-row boxes back to
// back), within +-1.6 %: packed bf16x2 fused multiply-adds (one instruction per two elements, 8 independent chains of 4),
// partial sums widened to fp32 at the end.  All terms are positive, so every rounding is a RELATIVE error of at most 2^-8 of
// the partial sum: (1 + 2^-8)^4 - 1 = 1.6 %; the caller folds that into its head room.  The chunk order is rotated per lane
// so that a warp reading 32 consecutive rows is bank-conflict free.
__device__ __forceinline__ float row_sumsq(const uint8_t* tile, int r, int lane) {
    const uint8_t* p = tile + (r >> 3) * 1024 + (r & 7) * 128;
    uint32_t acc[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const uint4 u = *reinterpret_cast<const uint4*>(p + (((j ^ lane) & 7) << 4));
        const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            uint32_t& c = acc[(j & 1) * 4 + e];
            asm("fma.rn.bf16x2 %0, %1, %1, %0;" : "+r"(c) : "r"(w[e]));
        }
    }
    float s0 = 0.f, s1 = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        s0 += __uint_as_float(acc[i] << 16);
        s1 += __uint_as_float(acc[i] & 0xffff0000u);
    }
    return s0 + s1;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

template <bool PROF>  // PROF: timeline stamps (a separate instantiation, the production kernel carries none of it)
__global__ void __launch_bounds__(THREADS, 1)
attn_tc2_kernel(const __grid_constant__ CUtensorMap tq, const __grid_constant__ CUtensorMap tk,
                const __grid_constant__ CUtensorMap tv, const Tc2Args a) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t q_full[2], q_free[2], k_full[2], k_free[2], v_full, v_free, s_full[2][2], p_full[2], o_full, n_full[4];
    __shared__ uint32_t tmem_slot;
    __shared__ float qn[4][QT];      // per tile (ring of 4): the bound m_i * log2(e) of every query row
    __shared__ int exact_f[4][2];    // per tile and bounds warp: bound too large -> exact maximum pass
    __shared__ float kpart[2][2];    // per item parity and bounds warp: max_j |k_j|^2 over that warp's keys
    __shared__ float smax[2][QT], ssum[2][2][QT];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int KP = a.KP, KP32 = (KP + 31) & ~31;
    const int nq = (a.Lq + QT - 1) / QT;
    const int nb = KP >> 4, hb = (nb + 1) >> 1, HB = hb * 16;  // key half 0 = [0, HB), half 1 = [HB, KP)
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* Qs = smem;                       // 2 x [128 x 64] bf16
    uint8_t* Ks = Qs + 2 * QBYTES;            // 2 x [KP32 x 64]
    uint8_t* Vs = Ks + 2 * KP32 * 128;        // [KP32 x 64]
    uint8_t* Ps = Vs + KP32 * 128;            // tail tiles only: nchunks x [128 x 64] (rows 0..31 written)
    const long long n_items = a.items > blockIdx.x ? (a.items - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const int ntiles = (int)(n_items * nq);   // query tiles of this CTA

    if (threadIdx.x == 0) {
        tc::prefetch_tmap(&tq); tc::prefetch_tmap(&tk); tc::prefetch_tmap(&tv);
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            tc::mbar_init(&q_full[i], 1); tc::mbar_init(&q_free[i], 3);   // free: MMA commit + the two bounds warps
            tc::mbar_init(&k_full[i], 1); tc::mbar_init(&k_free[i], 3);
            tc::mbar_init(&s_full[i][0], 1); tc::mbar_init(&s_full[i][1], 1);
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) tc::mbar_init(&n_full[i], 2);   // one arrival per bounds warp
        tc::mbar_init(&v_full, 1); tc::mbar_init(&v_free, 1);
        tc::mbar_init(&p_full[0], 4); tc::mbar_init(&p_full[1], 4); tc::mbar_init(&o_full, 1);  // p_full[h]: one arrival per softmax warp of key half h
        tc::fence_barrier_init();
    }
    if (warp == 8) tc::tmem_alloc(&tmem_slot, TMEM_COLS);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem_base = tc::warp_uniform(tmem_slot);
    tc::grid_dep_wait();     // PDL: the prologue above overlapped the previous kernel's tail
    tc::grid_dep_launch();   // persistent grid, fully resident
    auto stamp = [&](int tile, int ev) {  // events: 0-4 key half 0 (S ready, exp done, arrived, o_full seen, epilogue done), 6-8 key half 1,
                                          // 11 bounds done, 12-15 MMA warp (p_full seen, P V issued, S.H1 issued, top of iteration)
        if (PROF && a.tl != nullptr && blockIdx.x == 0 && lane == 0 && tile < 64) a.tl[tile * 16 + ev] = (unsigned long long)clock64();
    };

    if (warp == 9) {
        // ===================================================== TMA producer (all 32 lanes, one elected lane issues)
        int T = 0, it = 0;
        const uint32_t q_s = tc::smem_u32(Qs), k_s = tc::smem_u32(Ks), v_s = tc::smem_u32(Vs);
        auto load_q = [&](int qrow0, int col, int i) {
            const int b = T & 1;
            tc::mbar_wait(&q_free[b], (uint32_t)(((T >> 1) & 1) ^ 1));
            tc::mbar_expect_tx_w(&q_full[b], QBYTES);
            const bool tail = a.Lq - i * QT <= 32;  // tail tile: the same 32-row box goes to all four lane quarters
#pragma unroll
            for (int j = 0; j < QT / 32; ++j)
                tc::tma_load_2d_w(q_s + (uint32_t)(b * QBYTES + j * 4096), &tq, &q_full[b], col, qrow0 + i * QT + (tail ? 0 : j * 32));
            ++T;
        };
        // K is double-buffered across items and requested ONE ITEM AHEAD, before the wait for the single V buffer: the bounds warp
        // needs K(it+1) (row norms of ~280 keys, ~3000 clk) before the first tile of item it+1 can start its exp pass; requested
        // behind V(it) it arrived ~3000 clk later and the key-half-1 warps waited ~1300 clk at every item boundary
        // (tools/attn_timeline.py).
        auto request_k = [&](long long w, int item) {
            const int seq = (int)(w / a.heads);
            const int col = (int)(w - (long long)seq * a.heads) * HD;
            const int kvrow0 = (int)(seq * a.kv_sa);
            const int kb = item & 1;
            tc::mbar_wait(&k_free[kb], (uint32_t)(((item >> 1) & 1) ^ 1));
            tc::mbar_expect_tx_w(&k_full[kb], KP32 * 128);
            for (int j = 0; j < KP32 / 32; ++j)
                tc::tma_load_2d_w(k_s + (uint32_t)(kb * KP32 * 128 + j * 4096), &tk, &k_full[kb], col, kvrow0 + j * 32);
        };
        if ((long long)blockIdx.x < a.items) request_k(blockIdx.x, 0);
        for (long long w = blockIdx.x; w < a.items; w += gridDim.x, ++it) {
            const int seq = (int)(w / a.heads);
            const int col = (int)(w - (long long)seq * a.heads) * HD;
            const int qrow0 = (int)(seq * a.q_sa), kvrow0 = (int)(seq * a.kv_sa);  // < 2^31 rows (checked on the host)
            load_q(qrow0, col, 0);
            if (nq > 1) load_q(qrow0, col, 1);
            if (w + gridDim.x < a.items) request_k(w + gridDim.x, it + 1);
            tc::mbar_wait(&v_free, (uint32_t)((it & 1) ^ 1));
            tc::mbar_expect_tx_w(&v_full, KP32 * 128);
            for (int j = 0; j < KP32 / 32; ++j)
                tc::tma_load_2d_w(v_s + (uint32_t)(j * 4096), &tv, &v_full, col, kvrow0 + j * 32);
            for (int i = 2; i < nq; ++i) load_q(qrow0, col, i);
        }
    } else if (warp >= 10) {
        // ===================================================== bounds: m_i * log2(e) = |q_i| * max_j |k_j| * scale * log2(e)
        // TWO warps (10, 11: different SM sub-partitions) split the key rows and the query rows: one warp needed ~3300 clk for
        // the keys of an item and ~2600 for the 128 query rows of a tile (it shares its scheduler with two softmax warps),
        // which delayed the first tile of every item (tools/attn_timeline.py)
        const int nw = warp - 10;
        int T = 0, it = 0;
        const float sl2 = fabsf(a.sl2) * 1.02f;   // head room: 0.8 % for each of the two norms (row_sumsq is within 1.6 %) + sqrt.approx + the rounding of S
        auto sqrt_fast = [](float x) { float y; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; };
        for (long long w = blockIdx.x; w < a.items; w += gridDim.x, ++it) {
            const int kb = it & 1;
            tc::mbar_wait(&k_full[kb], (uint32_t)((it >> 1) & 1));
            if (nw == 0) stamp(T, 9);
            float km = 0.f;
            for (int r = lane + 32 * nw; r < a.Lkv; r += 64) km = fmaxf(km, row_sumsq(Ks + kb * KP32 * 128, r, lane));
            km = warp_max(km);
            if (lane == 0) kpart[kb][nw] = km;
            tc::named_bar_sync(3, 64);             // the two bounds warps (a warp is never more than one item ahead of the other)
            const float kmax = sqrt_fast(fmaxf(kpart[kb][0], kpart[kb][1]));
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(&k_free[kb]);
            if (nw == 0) stamp(T, 10);
            for (int i = 0; i < nq; ++i, ++T) {
                const int b = T & 1, slot = T & 3;
                tc::mbar_wait(&q_full[b], (uint32_t)((T >> 1) & 1));
                if (nw == 0) stamp(T, 5);
                const int rem = a.Lq - i * QT;
                const bool tail = rem <= 32;
                float mb = 0.f;
#pragma unroll
                for (int j = 2 * nw; j < 2 * nw + 2; ++j) {
                    const int r = j * 32 + lane;
                    const float m = sqrt_fast(row_sumsq(Qs + b * QBYTES, r, lane)) * kmax * sl2;
                    qn[slot][r] = m;
                    if ((tail ? lane : r) < rem) mb = fmaxf(mb, m);   // rows past the sequence hold other data
                }
                mb = warp_max(mb);
                if (lane == 0) exact_f[slot][nw] = (a.force_exact || !(mb <= BOUND_MAX)) ? 1 : 0;
                __syncwarp();
                if (lane == 0) { tc::mbar_arrive(&n_full[slot]); tc::mbar_arrive(&q_free[b]); }
                if (nw == 0) stamp(T, 11);
            }
        }
    } else if (warp == 8) {
        // ===================================================== MMA issuer (all 32 lanes, one elected lane issues)
        if (ntiles > 0) {
            const uint32_t q_s = tc::smem_u32(Qs), k_s = tc::smem_u32(Ks), v_s = tc::smem_u32(Vs), p_s = tc::smem_u32(Ps);
            auto issue_s = [&](int T, int it, int h) {  // S(T), key half h -> ring buffer (2T + h) % 3
                const int ncols = h ? KP - HB : HB;
                if (ncols > 0) {
                    const uint64_t da = tc::make_smem_desc_k128(q_s + (uint32_t)((T & 1) * QBYTES));
                    const uint64_t db = tc::make_smem_desc_k128(k_s + (uint32_t)((it & 1) * KP32 * 128 + h * HB * 128));
                    const uint32_t idesc = tc::make_idesc_bf16(QT, ncols);
                    const uint32_t d = tmem_base + S_COL + (uint32_t)((2 * T + h) % 3) * SBUF;
                    tc::umma_bf16_w4(d, da, db, idesc);   // the four K = 16 steps of head_dim 64 under ONE election
                }
                tc::umma_commit_w(&s_full[T & 1][h]);
            };
            // P V(T) is issued PER KEY HALF as soon as that half of P is complete: the key-half-0 warps arrive ~1000 clk before
            // the key-half-1 warps (they get their S earlier), so half of the P V issue time leaves the critical chain
            // P(T).H1 -> P V -> S(T+1).H1 -> exp.  The arrival of the key-half-0 warps also says "O(T-1) has been read back".
            constexpr uint32_t idesc_pv = tc::make_idesc_bf16_bmn(QT, HD);
            auto issue_pv_tail = [&]() {  // P in shared memory (rows 0..31 of every 64-key chunk), keys of all eight warps
                for (int j = 0; j < KP / 16; ++j) {
                    const uint64_t da = tc::make_smem_desc_k128(p_s + (uint32_t)((j >> 2) * QBYTES)) + (uint64_t)(2 * (j & 3));
                    const uint64_t db = tc::make_smem_desc_k128(v_s + (uint32_t)(j * 2048));  // 16 keys x 128 B, MN-major
                    tc::umma_bf16_w(tmem_base + O_COL, da, db, idesc_pv, j != 0);
                }
            };
            auto issue_pv_half = [&](int T, int h) {  // P in tensor memory, in place over the first half of the columns of its S half-buffer
                const int nj = (h ? KP - HB : HB) >> 4;
                const uint32_t pa = tmem_base + S_COL + (uint32_t)((2 * T + h) % 3) * SBUF;
                uint32_t acc = h ? 1u : 0u;   // half 0 starts the accumulator (it is never empty: HB >= 16)
                int j = 0;
                for (; j + 4 <= nj; j += 4) {   // four K = 16 steps per election
                    const uint64_t db = tc::make_smem_desc_k128(v_s + (uint32_t)((h * hb + j) * 2048));
                    tc::umma_bf16_ts_w4(tmem_base + O_COL, pa + (uint32_t)(8 * j), db, idesc_pv, acc);
                    acc = 1;
                }
                for (; j < nj; ++j) {
                    const uint64_t db = tc::make_smem_desc_k128(v_s + (uint32_t)((h * hb + j) * 2048));
                    tc::umma_bf16_ts_w(tmem_base + O_COL, pa + (uint32_t)(8 * j), db, idesc_pv, acc);
                    acc = 1;
                }
            };
            auto after_s = [&](int T, int it, int i) {  // both halves of S(T) issued: its Q buffer and (last tile) its K buffer retire with them
                tc::umma_commit_w(&q_free[T & 1]);
                if (i == nq - 1) tc::umma_commit_w(&k_free[it & 1]);
            };
            tc::mbar_wait(&k_full[0], 0);
            tc::mbar_wait(&q_full[0], 0);
            tc::fence_after_sync();
            issue_s(0, 0, 0); issue_s(0, 0, 1); after_s(0, 0, 0);
            int it = 0, i = 0;  // item / tile-in-item of tile T
            for (int T = 0; T < ntiles; ++T) {
                stamp(T, 15);
                const bool last_of_item = i + 1 == nq;
                const int it1 = last_of_item ? it + 1 : it, i1 = last_of_item ? 0 : i + 1;  // of tile T + 1
                if (T + 1 < ntiles) {
                    if (last_of_item) tc::mbar_wait(&k_full[it1 & 1], (uint32_t)((it1 >> 1) & 1));
                    tc::mbar_wait(&q_full[(T + 1) & 1], (uint32_t)(((T + 1) >> 1) & 1));
                    tc::fence_after_sync();
                    issue_s(T + 1, it1, 0);  // ring buffer of S(T-1).H1 = P(T-1).H1: P V(T-1) is ahead of it in the (in-order) MMA pipe
                }
                tc::mbar_wait(&p_full[0], (uint32_t)(T & 1));  // P(T).H0 written; O(T-1) read back
                if (i == 0) tc::mbar_wait(&v_full, (uint32_t)(it & 1));
                tc::fence_after_sync();
                if (a.Lq - i * QT <= 32) {   // tail tile: every warp holds keys of both halves
                    tc::mbar_wait(&p_full[1], (uint32_t)(T & 1));
                    tc::fence_after_sync();
                    stamp(T, 12);
                    issue_pv_tail();
                } else {
                    issue_pv_half(T, 0);
                    tc::mbar_wait(&p_full[1], (uint32_t)(T & 1));  // P(T).H1 written
                    tc::fence_after_sync();
                    stamp(T, 12);
                    issue_pv_half(T, 1);
                }
                tc::umma_commit_w(&o_full);
                stamp(T, 13);
                // S(T+1).H1 goes into the ring buffer of S(T).H0, which now holds P(T).H0: behind P V(T) in the pipe
                if (T + 1 < ntiles) { issue_s(T + 1, it1, 1); after_s(T + 1, it1, i1); }
                stamp(T, 14);
                if (last_of_item) tc::umma_commit_w(&v_free);
                it = it1; i = i1;
            }
        }
    } else {
        // ===================================================== softmax + epilogue warps
        // Critical path of a tile: P V(T) -> S(T+1).H1 -> exp pass of the key-half-1 warps -> P V(T+1).  The key-half-0 warps
        // get S(T+1).H0 before P V(T) even starts, so they are idle for most of that time: THEY read back all 64 columns of
        // O(T) (after their exp pass), the key-half-1 warps never touch O.
        const int q = warp & 3, h = warp >> 2, row = q * 32 + lane;
        const int c_begin = h ? HB : 0, c_end = h ? KP : HB;
        const uint32_t t_lane = tmem_base + ((uint32_t)(q * 32) << 16);
        uint8_t* prow_tail = Ps + (lane >> 3) * 1024 + (lane & 7) * 128;  // tail tiles: row = lane in every quarter
        int T = 0;
        // deferred epilogue of the previous tile (key-half-0 warps only)
        bool pend = false, pend_store = false, pend_active = false, pend_tail = false;
        __nv_bfloat16* pend_ptr = nullptr;
        float* pend_lse = nullptr;
        float pend_msc = 0.f;
        int pend_T = 0;
        auto epilogue = [&]() {
            tc::mbar_wait(&o_full, (uint32_t)(pend_T & 1));
            tc::fence_after_sync();
            if (q == 0) stamp(pend_T, 3);
            if (pend_active) {  // warp-uniform (tcgen05.ld is .sync.aligned); only the store is per-row
                const int sb = pend_T & 1;
                float sum;
                if (pend_tail) {  // the eight warps each hold a slice of the keys of row `lane`
                    sum = 0.f;
#pragma unroll
                    for (int j = 0; j < 4; ++j) sum += ssum[sb][0][j * 32 + lane] + ssum[sb][1][j * 32 + lane];
                } else {
                    sum = ssum[sb][0][row] + ssum[sb][1][row];
                }
                const float inv = 1.f / sum;
                if (pend_store && pend_lse != nullptr) *pend_lse = pend_msc + __log2f(sum);  // log2-domain log-sum-exp of the row
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    float o[32];
                    tc::tmem_ld32(t_lane + O_COL + half * 32, o);
                    if (pend_store) {
#pragma unroll
                        for (int j = 0; j < 32; j += 8)
                            *reinterpret_cast<uint4*>(pend_ptr + half * 32 + j) =
                                make_uint4(hv_pack_bf16(o[j] * inv, o[j + 1] * inv), hv_pack_bf16(o[j + 2] * inv, o[j + 3] * inv),
                                           hv_pack_bf16(o[j + 4] * inv, o[j + 5] * inv), hv_pack_bf16(o[j + 6] * inv, o[j + 7] * inv));
                    }
                }
            }
            tc::fence_before_sync();
            if (q == 0) stamp(pend_T, 4);
        };
        for (long long w = blockIdx.x; w < a.items; w += gridDim.x) {
            const long long seq = w / a.heads;
            const int head = (int)(w - seq * a.heads);
            for (int i = 0; i < nq; ++i, ++T) {
                const bool active = i * QT + q * 32 < a.Lq;  // warp-uniform: this warp has query rows in the tile
                const bool tail = a.Lq - i * QT <= 32;       // keys split eight ways, P through shared memory
                const int nch = (c_end - c_begin) >> 4;
                const int slot = T & 3;
                const uint32_t s_base = t_lane + S_COL + (uint32_t)((2 * T + h) % 3) * SBUF;  // column of key c_begin; P in place
                const uint32_t s_addr = s_base - (uint32_t)c_begin;
                tc::mbar_wait(&n_full[slot], (uint32_t)((T >> 2) & 1));
                const bool exact = (exact_f[slot][0] | exact_f[slot][1]) != 0;
                float msc = qn[slot][row];
                tc::mbar_wait(&s_full[T & 1][h], (uint32_t)((T >> 1) & 1));
                tc::fence_after_sync();
                if (q == 0) stamp(T, h * 6 + 0);
                if (exact) {
                    // ---- exact path: row maximum over this thread's keys, exchanged through shared memory
                    float mx = -INFINITY;
                    if (tail) {
                        uint32_t va[16];
                        for (int k = q; k < nch; k += 4) {
                            const int c = c_begin + (k << 4);
                            ld16_issue(s_addr + c, va);
                            ld16_wait(va);
#pragma unroll
                            for (int j = 0; j < 16; ++j) if (c + j < a.Lkv) mx = fmaxf(mx, __uint_as_float(va[j]));
                        }
                    } else if (active) {
                        uint32_t va[16];
                        for (int c = c_begin; c < c_end; c += 16) {
                            ld16_issue(s_addr + c, va);
                            ld16_wait(va);
#pragma unroll
                            for (int j = 0; j < 16; ++j) if (c + j < a.Lkv) mx = fmaxf(mx, __uint_as_float(va[j]));
                        }
                    }
                    smax[h][row] = mx;
                    tc::named_bar_sync(1, 256);
                    if (tail) {
#pragma unroll
                        for (int j = 0; j < 4; ++j) mx = fmaxf(mx, fmaxf(smax[0][j * 32 + lane], smax[1][j * 32 + lane]));
                    } else {
                        mx = fmaxf(mx, smax[h ^ 1][row]);
                    }
                    msc = mx * a.sl2;
                    tc::named_bar_sync(2, 256);  // smax is reused by the next exact tile
                }
                // ---- the only pass on the fast path: P = exp2(S*c - m*c) -> bf16, row sums
                float sum4[4] = {0.f, 0.f, 0.f, 0.f};
                auto expo = [&](uint32_t (&r)[16], int c, float (&v)[16]) {
                    if (c + 16 <= a.Lkv) {
#pragma unroll
                        for (int j = 0; j < 16; ++j) v[j] = tc::ex2_approx(fmaf(__uint_as_float(r[j]), a.sl2, -msc));
                    } else {
#pragma unroll
                        for (int j = 0; j < 16; ++j) v[j] = (c + j < a.Lkv) ? tc::ex2_approx(fmaf(__uint_as_float(r[j]), a.sl2, -msc)) : 0.f;
                    }
#pragma unroll
                    for (int j = 0; j < 4; ++j) sum4[j] += (v[j] + v[j + 4]) + (v[j + 8] + v[j + 12]);
                };
                if (tail) {
                    uint32_t va[16];
                    for (int k = q; k < nch; k += 4) {
                        const int c = c_begin + (k << 4);
                        ld16_issue(s_addr + c, va);
                        ld16_wait(va);
                        float v[16];
                        expo(va, c, v);
                        uint8_t* pc = prow_tail + (c >> 6) * QBYTES;
                        const int jc = (c & 63) >> 3;
                        *reinterpret_cast<uint4*>(pc + (((jc) ^ (lane & 7)) << 4)) =
                            make_uint4(hv_pack_bf16(v[0], v[1]), hv_pack_bf16(v[2], v[3]), hv_pack_bf16(v[4], v[5]), hv_pack_bf16(v[6], v[7]));
                        *reinterpret_cast<uint4*>(pc + (((jc + 1) ^ (lane & 7)) << 4)) =
                            make_uint4(hv_pack_bf16(v[8], v[9]), hv_pack_bf16(v[10], v[11]), hv_pack_bf16(v[12], v[13]), hv_pack_bf16(v[14], v[15]));
                    }
                } else if (active && c_begin < c_end) {
                    // P(c .. c+15) goes to columns (c - c_begin) / 2 .. +7 of the same half-buffer: always behind the reads.
                    // Software-pipelined by hand: one warp per SM sub-partition runs this loop most of the time (the other one is
                    // in a different phase), so nothing hides the 16 MUFU.EX2 of a chunk (8 clk each) but this warp's OWN
                    // independent work: the row sums, the bf16 packing and the P store of chunk i-1 are issued inside the MUFU
                    // phase of chunk i (measured before: 331 clk per chunk against the MUFU bound of 128, tools/attn_timeline.py).
                    auto ex = [&](const uint32_t (&r)[16], float (&e)[16]) {
#pragma unroll
                        for (int j = 0; j < 16; ++j) e[j] = tc::ex2_approx(fmaf(__uint_as_float(r[j]), a.sl2, -msc));
                    };
                    auto finish = [&](const float (&v)[16], int c) {
#pragma unroll
                        for (int j = 0; j < 4; ++j) sum4[j] += (v[j] + v[j + 4]) + (v[j + 8] + v[j + 12]);
                        uint32_t pk[8];
#pragma unroll
                        for (int j = 0; j < 8; ++j) pk[j] = hv_pack_bf16(v[2 * j], v[2 * j + 1]);
                        tc::tmem_st8(s_base + (uint32_t)((c - c_begin) >> 1), pk);
                    };
                    const int lk16 = a.Lkv & ~15;
                    const int cfe = c_end < lk16 ? c_end : lk16;   // end of the fully valid chunks of this thread's key range
                    int c = c_begin;
                    if (c < cfe) {
                        uint32_t va[16], vb[16];
                        float ea[16], eb[16];
                        ld16_issue(s_addr + c, va);
                        ld16_wait(va);
                        bool more = c + 16 < cfe;
                        if (more) ld16_issue(s_addr + c + 16, vb);
                        ex(va, ea);
                        bool in_b = false;   // the chunk that still has to be finished sits in eb (else ea)
                        while (more) {
                            c += 16;
                            ld16_wait(vb);
                            more = c + 16 < cfe;
                            if (more) ld16_issue(s_addr + c + 16, va);
                            ex(vb, eb);
                            finish(ea, c - 16);
                            in_b = true;
                            if (!more) break;
                            c += 16;
                            ld16_wait(va);
                            more = c + 16 < cfe;
                            if (more) ld16_issue(s_addr + c + 16, vb);
                            ex(va, ea);
                            finish(eb, c - 16);
                            in_b = false;
                        }
                        if (in_b) finish(eb, c); else finish(ea, c);
                        c += 16;
                    }
                    if (c < c_end) {  // the one partially valid chunk (Lkv not a multiple of 16), masked
                        uint32_t va[16];
                        ld16_issue(s_addr + c, va);
                        ld16_wait(va);
                        float v[16];
                        expo(va, c, v);
                        uint32_t pk[8];
#pragma unroll
                        for (int j = 0; j < 8; ++j) pk[j] = hv_pack_bf16(v[2 * j], v[2 * j + 1]);
                        tc::tmem_st8(s_base + (uint32_t)((c - c_begin) >> 1), pk);
                    }
                }
                ssum[T & 1][h][row] = (sum4[0] + sum4[1]) + (sum4[2] + sum4[3]);
                if (q == 0) stamp(T, h * 6 + 1);
                if (h == 0 && pend) epilogue();      // O(T-1), all 64 columns: off the critical path (see above)
                // ---- publish: P(T) complete (and O(T-1) read back) -> the MMA warp may issue P V(T)
                tc::tmem_st_wait();
                tc::fence_before_sync();
                if (tail) tc::fence_proxy_async();   // P of a tail tile: generic-proxy stores -> visible to the tensor core
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(&p_full[h]);
                if (q == 0) stamp(T, h * 6 + 2);
                if (h == 0) {
                    const int qr = i * QT + row;
                    pend = true;
                    pend_T = T;
                    pend_active = active;
                    pend_tail = tail;
                    pend_store = qr < a.Lq;
                    pend_ptr = a.out + (seq * a.q_sa + qr) * a.ldo + head * HD;
                    pend_lse = a.lse ? a.lse + (w * a.Lq + (qr < a.Lq ? qr : 0)) * 2 : nullptr;
                    pend_msc = msc;
                }
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

static unsigned long long* g_tl = nullptr;
// HV_ATTN_TIMELINE=1: copies the stamps of the last launches (64 tiles x 16 events of CTA 0) to host[1024] and clears them
extern "C" int hv_attn_timeline_read(unsigned long long* host) {
    if (!g_tl) return 0;
    cudaDeviceSynchronize();
    cudaMemcpy(host, g_tl, 1024 * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
    cudaMemset(g_tl, 0, 1024 * sizeof(unsigned long long));
    return 1;
}

// returns 1 if the shape is handled here (and launched), 0 if the caller should use another kernel, < 0 on error.
// HV_ATTN_TC2=0 switches this design off (the dispatcher then takes attn_tc.cu), HV_ATTN_EXACT=1 forces the maximum pass.
int hv_attention_tc2_bf16(const AttnArgs& a, int64_t seqs, cudaStream_t st) {
    static const bool off = [] { const char* e = getenv("HV_ATTN_TC2"); return e && atoi(e) == 0; }();
    static const int force_exact = [] { const char* e = getenv("HV_ATTN_EXACT"); return e ? atoi(e) : 0; }();
    if (off) return 0;
    const HvSeqMap &qm = a.qmap, &km = a.kvmap;
    const bool dense = qm.sdiv == 1 && qm.sc == 1 && km.sdiv == 1 && km.sc == 1;
    if (!dense || a.hd != HD || a.mask || a.Lkv < 16 || a.Lkv > KMAX || !(a.scale > 0.f)) return 0;
    if ((a.ldq % 8) || (a.ldk % 8) || (a.ldv % 8) || (a.ldo % 8)) return 0;
    auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
    if (!al16(a.q) || !al16(a.k) || !al16(a.v) || !al16(a.out)) return 0;
    const long long qrows = (seqs - 1) * qm.sa + a.Lq, kvrows = (seqs - 1) * km.sa + a.Lkv;
    if (qrows >= (1ll << 31) || kvrows >= (1ll << 31)) return 0;
    const long long tiles = seqs * a.heads * (long long)((a.Lq + QT - 1) / QT);
    if (tiles >= (1ll << 31)) return 0;  // per-CTA tile counters are 32-bit
    CUtensorMap tq, tk, tv;
    int rc;
    const int cols = a.heads * HD;
    if ((rc = hv_make_tmap_bf16(a.q, qrows, cols, a.ldq, 32, &tq)) != HV_OK) return rc;
    if ((rc = hv_make_tmap_bf16(a.k, kvrows, cols, a.ldk, 32, &tk)) != HV_OK) return rc;
    if ((rc = hv_make_tmap_bf16(a.v, kvrows, cols, a.ldv, 32, &tv)) != HV_OK) return rc;
    Tc2Args t;
    t.out = reinterpret_cast<__nv_bfloat16*>(a.out);
    t.ldo = a.ldo; t.q_sa = qm.sa; t.kv_sa = km.sa;
    t.heads = a.heads; t.Lq = a.Lq; t.Lkv = a.Lkv; t.KP = (a.Lkv + 15) & ~15;
    t.items = seqs * a.heads;
    t.sl2 = a.scale * 1.4426950408889634f;
    t.force_exact = force_exact;
    t.lse = a.lse_stats;
    static const bool timeline = [] { const char* e = getenv("HV_ATTN_TIMELINE"); return e && atoi(e) != 0; }();
    if (timeline && !g_tl && cudaMalloc(&g_tl, 1024 * sizeof(unsigned long long)) == cudaSuccess) cudaMemset(g_tl, 0, 1024 * sizeof(unsigned long long));
    t.tl = timeline ? g_tl : nullptr;
    const int KP32 = (t.KP + 31) & ~31, nchunks = (t.KP + 63) >> 6;
    const int last = a.Lq % QT;
    const bool has_tail = last >= 1 && last <= 32;
    const int smem = 2 * QBYTES + 3 * KP32 * 128 + (has_tail ? nchunks * QBYTES : 0) + 1024;
    static int configured_of[HV_MAX_DEVICES] = {};  // function attributes are per device
    int& configured = configured_of[hv_device_slot()];
    if (configured < smem) {
        HV_CUDA(cudaFuncSetAttribute(attn_tc2_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        HV_CUDA(cudaFuncSetAttribute(attn_tc2_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        configured = smem;
    }
    const int sms = hv_num_sms();
    const int grid = (int)(t.items < sms ? t.items : sms);
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cfg.attrs = at; cfg.numAttrs = hv_pdl_enabled() ? 1 : 0;
    if (t.tl) HV_CUDA(cudaLaunchKernelEx(&cfg, attn_tc2_kernel<true>, tq, tk, tv, t));
    else HV_CUDA(cudaLaunchKernelEx(&cfg, attn_tc2_kernel<false>, tq, tk, tv, t));
    HV_LAUNCH_CHECK();
    return 1;
}
