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
//       Also writes the log2-sum-exp per query row for the second kernel.  D comes from a small streaming pre-kernel
//       (one coalesced pass over O and dO; the row-scattered loads it replaces cost 4300 of 17000 cycles per tile).
//       Software pipeline: when dS~(T) is staged the MMA warp issues S(T+1) BEFORE dQ(T), and the compute warps read dQ(T)
//       back only after their pass 2 of tile T+1, so they never wait for the (slow, N = 64) dQ product.
//   dkv kernel, one 128-KEY tile at a time:  S^T = K_t Q^T (keys on the TMEM lanes, <= 288 query columns),
//       P^T = exp2(S^T c - lse_q) exact (lse from the dq kernel), kept in registers AND staged to shared memory,
//       dV_t = P^T dO (dO MN-major);  dP^T = V_t dO^T overwrites S^T;  dS^T = P^T (dP^T - D_q) scale -> the same staging
//       buffer once the dV MMA has retired (dS^T is formed in place in the P^T registers while that MMA runs) ->
//       dK_t = dS^T Q (Q MN-major).  Same pipeline: S^T(t+1) is issued before dK_t, the accumulators of tile t are read
//       back after pass 2 of tile t+1.
// No atomics anywhere: every output element is produced by exactly one CTA (bit-reproducible).
//
// Warp roles (384 threads): warps 0..7 compute (TMEM lane quarter = warp % 4, column half = warp / 4), warp 8 = MMA issuer +
// TMEM owner, warp 9 = TMA producer, warps 10-11 idle (they complete the third warpgroup: setmaxnreg moves registers from
// it to the compute warpgroups, 216 per compute thread — P lives in 72 of them).  CTAs are persistent over (sequence, head) pairs.
#include <cuda.h>
#include <cstdlib>
#include "attn.cuh"
#include "tc_ptx.cuh"

int hv_make_tmap_bf16(const void* ptr, long long rows, int K, long long ld, int box_rows, CUtensorMap* out);  // gemm_tc.cu

namespace {

constexpr int QT = 128, HD = 64, LMAX = 288, THREADS = 384;  // 12 warps = 3 warpgroups (setmaxnreg works per warpgroup)
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
    int have_lse;                // element 0 was written by the forward (hv_attention_lse): the dq kernel skips its maximum / sum passes
    unsigned long long* dbg;     // phase cycle counters (HV_ATTN_BWD_PROF=1) or nullptr: [kernel][cta][16]
};

// cycles since the previous tick of (warp 0, lane 0) -> counter `phase` of this CTA (only when profiling is on)
#define BWD_TICK(phase) do { if (prof) { const long long t_ = clock64(); a.dbg[(KERNEL_ID * 148 + (blockIdx.x % 148)) * 16 + (phase)] += (unsigned long long)(t_ - tk0); tk0 = t_; } } while (0)

__device__ __forceinline__ void regs_compute() { asm volatile("setmaxnreg.inc.sync.aligned.u32 216;"); }
__device__ __forceinline__ void regs_service() { asm volatile("setmaxnreg.dec.sync.aligned.u32 56;"); }

__device__ __forceinline__ void ld32_issue(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}
// issues the tcgen05.ld of columns [c, c + 32) (or [c, c + 16) when only 16 are left) and waits for it
__device__ __forceinline__ void ld_cols(uint32_t taddr, bool full32, uint32_t (&r)[32]) {
    if (full32) {
        ld32_issue(taddr, r);
    } else {
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
            "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
            : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
              "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
            : "r"(taddr)
            : "memory");
    }
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

__device__ __forceinline__ void ld16_issue(uint32_t taddr, uint32_t (&r)[16]);
__device__ __forceinline__ void ld16_wait(uint32_t (&r)[16]);
// Walks this thread's TMEM columns [c_begin, c_end) in 16-column chunks; the load of chunk k + 1 is in flight while f(k, c,
// values) works on chunk k.  TMEM -> register bandwidth is 64 B/clk/SM (one pass over a 128 x 288 fp32 tile = 2300 clk), so a
// pass is as fast as the hardware allows only when its arithmetic hides under the loads; the 32 extra live registers are
// what setmaxnreg pays for.  Fully unrolled: k is a compile-time constant inside f (register arrays are indexed statically).
template <typename F>
__device__ __forceinline__ void for_chunks(uint32_t s_addr, int c_begin, int c_end, F&& f) {
    uint32_t buf[2][16];
    ld16_issue(s_addr + c_begin, buf[0]);
#pragma unroll
    for (int k = 0; k < 9; ++k) {
        const int c = c_begin + 16 * k;
        if (c < c_end) {  // warp-uniform
            ld16_wait(buf[k & 1]);
            if (c + 16 < c_end) ld16_issue(s_addr + c + 16, buf[(k + 1) & 1]);
            f(k, c, buf[k & 1]);
        }
    }
}

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
// explicit st.shared (the staging pointers are carved out of the dynamic shared-memory block at run time; without the
// state space the compiler emits generic ST.E)
__device__ __forceinline__ void sts128(uint32_t saddr, const uint4 v) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(saddr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void stage16(uint8_t* base, int row, int c, const float (&v)[16]) {
    const uint32_t pc = tc::smem_u32(base) + (c >> 6) * TBYTES + (row >> 3) * 1024 + (row & 7) * 128;
    const int jc = (c & 63) >> 3;
    sts128(pc + ((jc ^ (row & 7)) << 4), make_uint4(hv_pack_bf16(v[0], v[1]), hv_pack_bf16(v[2], v[3]), hv_pack_bf16(v[4], v[5]), hv_pack_bf16(v[6], v[7])));
    sts128(pc + (((jc + 1) ^ (row & 7)) << 4), make_uint4(hv_pack_bf16(v[8], v[9]), hv_pack_bf16(v[10], v[11]), hv_pack_bf16(v[12], v[13]), hv_pack_bf16(v[14], v[15])));
}
__device__ __forceinline__ void stage16_packed(uint8_t* base, int row, int c, const uint4 lo, const uint4 hi) {  // 2 x 8 bf16
    const uint32_t pc = tc::smem_u32(base) + (c >> 6) * TBYTES + (row >> 3) * 1024 + (row & 7) * 128;
    const int jc = (c & 63) >> 3;
    sts128(pc + ((jc ^ (row & 7)) << 4), lo);
    sts128(pc + (((jc + 1) ^ (row & 7)) << 4), hi);
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
        tc::umma_bf16_w4(d_tmem + (uint32_t)r0, da, db, idesc);  // the four K = 16 steps of head_dim 64 under one election
    }
}
// PV-like product: D[128 x 64] = A[128 x kdim] B[kdim x 64], A staged K-major in 64-column chunks, B = [kdim rows][64] rows of
// 128 B as TMA delivers them (MN-major operand)
__device__ __forceinline__ void issue_pv_like(uint32_t d_tmem, const uint8_t* a_stage, const uint8_t* b_rows, int kdim) {
    constexpr uint32_t idesc = tc::make_idesc_bf16_bmn(QT, HD);
    const int n16 = kdim / 16;
    int j = 0;
    for (; j + 4 <= n16; j += 4) {  // one 64-column chunk of A = four K = 16 steps, one election (elect.sync costs ~70 clk per MMA)
        const uint64_t da = tc::make_smem_desc_k128(tc::smem_u32(a_stage + (j >> 2) * TBYTES));
        const uint64_t db = tc::make_smem_desc_k128(tc::smem_u32(b_rows + j * 2048));
        tc::umma_bf16_pv_w4(d_tmem, da, db, idesc, j != 0);
    }
    for (; j < n16; ++j) {
        const uint64_t da = tc::make_smem_desc_k128(tc::smem_u32(a_stage + (j >> 2) * TBYTES)) + (uint64_t)(2 * (j & 3));
        const uint64_t db = tc::make_smem_desc_k128(tc::smem_u32(b_rows + j * 2048));
        tc::umma_bf16_w(d_tmem, da, db, idesc, j != 0);
    }
}

// ------------------------------------------------------------------------------------------------ D = rowsum(dO o O)
// one thread per (sequence, query row, head), heads fastest: adjacent threads read adjacent 128-byte head slices
__global__ void __launch_bounds__(256) attn_bwd_delta_kernel(const __nv_bfloat16* __restrict__ o, long long ldo,
                                                             const __nv_bfloat16* __restrict__ dout, long long lddo, float* __restrict__ stats,
                                                             long long seqs, int heads, int Lq, long long q_sa) {
    const long long idx = blockIdx.x * 256ll + threadIdx.x;
    if (idx >= seqs * Lq * heads) return;
    const int head = (int)(idx % heads);
    const long long t = idx / heads, seq = t / Lq;
    const int row = (int)(t - seq * Lq);
    const uint4* po = reinterpret_cast<const uint4*>(o + (seq * q_sa + row) * ldo + head * HD);
    const uint4* pd = reinterpret_cast<const uint4*>(dout + (seq * q_sa + row) * lddo + head * HD);
    float d = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const uint4 x = __ldg(po + j), y = __ldg(pd + j);
        d += bf_lo(x.x) * bf_lo(y.x) + bf_hi(x.x) * bf_hi(y.x) + bf_lo(x.y) * bf_lo(y.y) + bf_hi(x.y) * bf_hi(y.y)
           + bf_lo(x.z) * bf_lo(y.z) + bf_hi(x.z) * bf_hi(y.z) + bf_lo(x.w) * bf_lo(y.w) + bf_hi(x.w) * bf_hi(y.w);
    }
    stats[(((seq * heads + head) * Lq) + row) * 2 + 1] = d;
}

// ------------------------------------------------------------------------------------------------ dQ (+ log-sum-exp)
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
    const uint32_t tmem_base = tc::warp_uniform(tmem_slot);  // uniform register: operand of warp-converged MMA issue

    if (warp >= 10) {
        regs_service();  // idle members of the service warpgroup
    } else if (warp == 9) {
        // ===================================================== TMA producer
        regs_service();
        {   // all 32 lanes, warp-uniform values; one elected lane issues (tc_ptx.cuh *_w)
            long long T = 0, it = 0;
            for (long long w = blockIdx.x; w < a.items; w += gridDim.x, ++it) {
                const long long seq = w / a.heads;
                const int col = (int)(w - seq * a.heads) * HD;
                const long long qrow0 = seq * a.q_sa, kvrow0 = seq * a.kv_sa;
                tc::mbar_wait(&kv_free, (uint32_t)((it & 1) ^ 1));
                tc::mbar_expect_tx_w(&kv_full, 2 * KP32 * 128);
                for (int j = 0; j < KP32 / 32; ++j) {
                    tc::tma_load_2d_w(tc::smem_u32(Ks + j * 4096), &tk, &kv_full, col, (int)(kvrow0 + j * 32));
                    tc::tma_load_2d_w(tc::smem_u32(Vs + j * 4096), &tv, &kv_full, col, (int)(kvrow0 + j * 32));
                }
                for (int i = 0; i < nq; ++i, ++T) {
                    tc::mbar_wait(&q_free, (uint32_t)((T & 1) ^ 1));
                    tc::mbar_expect_tx_w(&q_full, TBYTES);
                    for (int j = 0; j < 4; ++j) tc::tma_load_2d_w(tc::smem_u32(Qs + j * 4096), &tq, &q_full, col, (int)(qrow0 + i * QT + j * 32));
                    tc::mbar_wait(&do_free, (uint32_t)((T & 1) ^ 1));
                    tc::mbar_expect_tx_w(&do_full, TBYTES);
                    for (int j = 0; j < 4; ++j) tc::tma_load_2d_w(tc::smem_u32(dOs + j * 4096), &tdo, &do_full, col, (int)(qrow0 + i * QT + j * 32));
                }
            }
        }
    } else if (warp == 8) {
        // ===================================================== MMA issuer
        regs_service();
        {   // all 32 lanes, warp-uniform values; one elected lane issues (tc_ptx.cuh *_w)
            long long T = 0, it = 0;
            bool ahead = false;  // S of the current tile was already issued (before the previous tile's dQ)
            auto issue_s = [&](long long Tn) {
                tc::mbar_wait(&q_full, (uint32_t)(Tn & 1));
                tc::fence_after_sync();
                issue_s_like(tmem_base + S_COL, Qs, Ks, KP, HB);           // S = Q K^T
                tc::umma_commit_w(&s_full);
                tc::umma_commit_w(&q_free);
            };
            for (long long w = blockIdx.x; w < a.items; w += gridDim.x, ++it) {
                tc::mbar_wait(&kv_full, (uint32_t)(it & 1));
                for (int i = 0; i < nq; ++i, ++T) {
                    const uint32_t ph = (uint32_t)(T & 1);
                    if (!ahead) issue_s(T);
                    tc::mbar_wait(&do_full, ph);
                    tc::mbar_wait(&s_free, ph);                            // P~ is in registers, S may be overwritten
                    tc::fence_after_sync();
                    issue_s_like(tmem_base + S_COL, dOs, Vs, KP, HB);      // dP = dO V^T
                    tc::umma_commit_w(&dp_full);
                    tc::umma_commit_w(&do_free);
                    tc::mbar_wait(&ds_full, ph);                           // dS~ staged, dP consumed; the previous dQ was read back
                    tc::fence_after_sync();
                    ahead = i + 1 < nq;
                    if (ahead) issue_s(T + 1);                             // next tile of the same item first: the compute warps need it first
                    issue_pv_like(tmem_base + ACC_COL, dSs, Ks, KP);       // dQ~ = dS~ K
                    tc::umma_commit_w(&dq_full);
                    if (i == nq - 1) tc::umma_commit_w(&kv_free);
                }
            }
        }
    } else {
        // ===================================================== compute warps
        regs_compute();
        const int q = warp & 3, h = warp >> 2, row = q * 32 + lane;
        const int c_begin = h ? HB : 0, c_end = h ? KP : HB;
        const uint32_t t_lane = tmem_base + ((uint32_t)(q * 32) << 16);
        const uint32_t s_addr = t_lane + S_COL;
        long long T = 0;
        // deferred read-back of the previous tile's dQ
        bool pend = false, pend_valid = false;
        float pend_f = 0.f;
        __nv_bfloat16* pend_dst = nullptr;
        auto epilogue = [&]() {
            tc::mbar_wait(&dq_full, (uint32_t)((T - 1) & 1));
            tc::fence_after_sync();
            float o[32];
            tc::tmem_ld32(t_lane + ACC_COL + h * 32, o);
            if (pend_valid) {
                const float f = pend_f;
#pragma unroll
                for (int j = 0; j < 32; j += 8)
                    *reinterpret_cast<uint4*>(pend_dst + j) = make_uint4(hv_pack_bf16(o[j] * f, o[j + 1] * f), hv_pack_bf16(o[j + 2] * f, o[j + 3] * f),
                                                                         hv_pack_bf16(o[j + 4] * f, o[j + 5] * f), hv_pack_bf16(o[j + 6] * f, o[j + 7] * f));
            }
            tc::fence_before_sync();
        };
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
                float* const st = a.stats + (w * a.Lq + (valid ? qr : 0)) * 2;
                const float Dr = valid ? __ldg(st + 1) : 0.f;  // from attn_bwd_delta_kernel
                tc::mbar_wait(&s_full, ph);
                tc::fence_after_sync();
                BWD_TICK(1);   // wait S
                // ---- pass 1: row maximum over this thread's key half (skipped when the forward handed over the log-sum-exp: then
                // P = exp2(S c - lse) is already normalised and neither the maximum nor the row sum is needed)
                float mx = -INFINITY;
                if (!a.have_lse)
                for_chunks(s_addr, c_begin, c_end, [&](int, int c, const uint32_t (&va)[16]) {
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
                });
                float msc;
                if (a.have_lse) {
                    msc = valid ? __ldg(st) : 1e30f;   // rows past the sequence: P = 0
                    BWD_TICK(2);
                } else {
                    smax[h][row] = mx;
                    BWD_TICK(2);   // pass 1
                    tc::named_bar_sync(1, 256);
                    mx = fmaxf(mx, smax[h ^ 1][row]);
                    msc = mx * a.sl2;
                }
                BWD_TICK(3);   // max exchange
                // ---- pass 2: P~ = exp2(S c - max c) -> registers (packed bf16), row sum
                uint32_t preg[72];
                float sum = 0.f;
                for_chunks(s_addr, c_begin, c_end, [&](int k, int c, const uint32_t (&va)[16]) {
                    float p[16];
                    if (c + 16 <= a.Lkv) {   // every chunk but (at most) the last one: no per-element mask
#pragma unroll
                        for (int j = 0; j < 16; ++j) p[j] = tc::ex2_approx(fmaf(__uint_as_float(va[j]), a.sl2, -msc));
                    } else {
#pragma unroll
                        for (int j = 0; j < 16; ++j)
                            p[j] = (c + j < a.Lkv) ? tc::ex2_approx(fmaf(__uint_as_float(va[j]), a.sl2, -msc)) : 0.f;
                    }
#pragma unroll
                    for (int j = 0; j < 8; ++j) preg[8 * k + j] = hv_pack_bf16(p[2 * j], p[2 * j + 1]);
                    sum += ((p[0] + p[1]) + (p[2] + p[3])) + ((p[4] + p[5]) + (p[6] + p[7]))
                         + ((p[8] + p[9]) + (p[10] + p[11])) + ((p[12] + p[13]) + (p[14] + p[15]));
                });
                BWD_TICK(4);   // pass 2
                tc::fence_before_sync();
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(&s_free);
                float l = 1.f;
                if (!a.have_lse) {
                    ssum[h][row] = sum;
                    tc::named_bar_sync(1, 256);
                    l = sum + ssum[h ^ 1][row];
                    if (h == 0 && valid) st[0] = msc + __log2f(l);
                }
                BWD_TICK(5);   // sum exchange + stats
                // ---- the previous tile's dQ (its MMA was issued behind this tile's S and is long done)
                if (pend) epilogue();
                BWD_TICK(9);   // epilogue (deferred)
                // ---- pass 3: dS~ = P~ (dP - D) -> bf16 -> shared memory
                tc::mbar_wait(&dp_full, ph);
                tc::fence_after_sync();
                BWD_TICK(6);   // wait dP
                for_chunks(s_addr, c_begin, c_end, [&](int k, int c, const uint32_t (&va)[16]) {
                    float ds[16];
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        ds[2 * j] = bf_lo(preg[8 * k + j]) * (__uint_as_float(va[2 * j]) - Dr);
                        ds[2 * j + 1] = bf_hi(preg[8 * k + j]) * (__uint_as_float(va[2 * j + 1]) - Dr);
                    }
                    stage16(dSs, row, c, ds);
                });
                tc::fence_before_sync();
                tc::fence_proxy_async();
                __syncwarp();
                BWD_TICK(7);   // pass 3
                if (lane == 0) tc::mbar_arrive(&ds_full);
                pend = true;
                pend_valid = valid;
                pend_f = a.scale / l;
                pend_dst = a.dq + (seq * a.q_sa + qr) * a.lddq + head * HD + h * 32;
                if (prof) a.dbg[(KERNEL_ID * 148 + (blockIdx.x % 148)) * 16 + 15] += 1;
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

// ------------------------------------------------------------------------------------------------ dK, dV
__global__ void __launch_bounds__(THREADS, 1)
attn_bwd_dkv_tc_kernel(const __grid_constant__ CUtensorMap tq, const __grid_constant__ CUtensorMap tk,
                       const __grid_constant__ CUtensorMap tv, const __grid_constant__ CUtensorMap tdo, const BwdTc a) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t qdo_full, qdo_free, kt_full, kt_free, st_full, p_full, dpt_full, dv_done, ds_full, acc_full;
    __shared__ uint32_t tmem_slot;
    __shared__ __align__(16) float sLse[LMAX], sD[LMAX];

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
    const uint32_t tmem_base = tc::warp_uniform(tmem_slot);  // uniform register: operand of warp-converged MMA issue

    if (warp >= 10) {
        regs_service();  // idle members of the service warpgroup
    } else if (warp == 9) {
        // ===================================================== TMA producer
        regs_service();
        {   // all 32 lanes, warp-uniform values; one elected lane issues (tc_ptx.cuh *_w)
            long long T = 0, it = 0;
            for (long long w = blockIdx.x; w < a.items; w += gridDim.x, ++it) {
                const long long seq = w / a.heads;
                const int col = (int)(w - seq * a.heads) * HD;
                const long long qrow0 = seq * a.q_sa, kvrow0 = seq * a.kv_sa;
                tc::mbar_wait(&qdo_free, (uint32_t)((it & 1) ^ 1));
                tc::mbar_expect_tx_w(&qdo_full, 2 * QP32 * 128);
                for (int j = 0; j < QP32 / 32; ++j) {
                    tc::tma_load_2d_w(tc::smem_u32(Qs + j * 4096), &tq, &qdo_full, col, (int)(qrow0 + j * 32));
                    tc::tma_load_2d_w(tc::smem_u32(dOs + j * 4096), &tdo, &qdo_full, col, (int)(qrow0 + j * 32));
                }
                for (int t = 0; t < nk; ++t, ++T) {
                    tc::mbar_wait(&kt_free, (uint32_t)((T & 1) ^ 1));
                    tc::mbar_expect_tx_w(&kt_full, 2 * TBYTES);
                    for (int j = 0; j < 4; ++j) {
                        tc::tma_load_2d_w(tc::smem_u32(Kt + j * 4096), &tk, &kt_full, col, (int)(kvrow0 + t * QT + j * 32));
                        tc::tma_load_2d_w(tc::smem_u32(Vt + j * 4096), &tv, &kt_full, col, (int)(kvrow0 + t * QT + j * 32));
                    }
                }
            }
        }
    } else if (warp == 8) {
        // ===================================================== MMA issuer
        regs_service();
        {   // all 32 lanes, warp-uniform values; one elected lane issues (tc_ptx.cuh *_w)
            long long T = 0, it = 0;
            bool ahead = false;
            auto issue_st = [&](long long Tn) {
                tc::mbar_wait(&kt_full, (uint32_t)(Tn & 1));
                tc::fence_after_sync();
                issue_s_like(tmem_base + S_COL, Kt, Qs, QP, HB);            // S^T = K_t Q^T
                tc::umma_commit_w(&st_full);
            };
            for (long long w = blockIdx.x; w < a.items; w += gridDim.x, ++it) {
                tc::mbar_wait(&qdo_full, (uint32_t)(it & 1));
                for (int t = 0; t < nk; ++t, ++T) {
                    const uint32_t ph = (uint32_t)(T & 1);
                    if (!ahead) issue_st(T);
                    tc::mbar_wait(&p_full, ph);                             // P^T in registers + staged; accumulators of the previous tile read back
                    tc::fence_after_sync();
                    issue_s_like(tmem_base + S_COL, Vt, dOs, QP, HB);       // dP^T = V_t dO^T over the S^T columns
                    tc::umma_commit_w(&dpt_full);
                    tc::umma_commit_w(&kt_free);                              // K_t / V_t retire: the producer prefetches the next key tile
                    issue_pv_like(tmem_base + ACC_COL, Ps, dOs, QP);        // dV_t = P^T dO
                    tc::umma_commit_w(&dv_done);
                    tc::mbar_wait(&ds_full, ph);                            // dS^T staged (after dv_done), dP^T consumed
                    tc::fence_after_sync();
                    ahead = t + 1 < nk;
                    if (ahead) issue_st(T + 1);
                    issue_pv_like(tmem_base + ACC_COL + HD, Ps, Qs, QP);    // dK_t = dS^T Q
                    tc::umma_commit_w(&acc_full);
                    if (t == nk - 1) tc::umma_commit_w(&qdo_free);
                }
            }
        }
    } else {
        // ===================================================== compute warps: thread = (key row, query half)
        regs_compute();
        const int q = warp & 3, h = warp >> 2, row = q * 32 + lane;
        const int c_begin = h ? HB : 0, c_end = h ? QP : HB;
        const uint32_t t_lane = tmem_base + ((uint32_t)(q * 32) << 16);
        const uint32_t s_addr = t_lane + S_COL;
        long long T = 0;
        bool pend = false, pend_valid = false;
        __nv_bfloat16 *pend_v = nullptr, *pend_k = nullptr;
        auto epilogue = [&]() {
            tc::mbar_wait(&acc_full, (uint32_t)((T - 1) & 1));
            tc::fence_after_sync();
            float ov[32], ok[32];
            tc::tmem_ld32(t_lane + ACC_COL + h * 32, ov);
            tc::tmem_ld32(t_lane + ACC_COL + HD + h * 32, ok);
            if (pend_valid) {
#pragma unroll
                for (int j = 0; j < 32; j += 8) {
                    *reinterpret_cast<uint4*>(pend_v + j) = make_uint4(hv_pack_bf16(ov[j], ov[j + 1]), hv_pack_bf16(ov[j + 2], ov[j + 3]),
                                                                       hv_pack_bf16(ov[j + 4], ov[j + 5]), hv_pack_bf16(ov[j + 6], ov[j + 7]));
                    *reinterpret_cast<uint4*>(pend_k + j) = make_uint4(hv_pack_bf16(ok[j], ok[j + 1]), hv_pack_bf16(ok[j + 2], ok[j + 3]),
                                                                       hv_pack_bf16(ok[j + 4], ok[j + 5]), hv_pack_bf16(ok[j + 6], ok[j + 7]));
                }
            }
            tc::fence_before_sync();
        };
        for (long long w = blockIdx.x; w < a.items; w += gridDim.x) {
            const long long seq = w / a.heads;
            const int head = (int)(w - seq * a.heads);
            // per-query statistics -> shared memory (the previous item's readers are past their last use: every thread has
            // arrived on ds_full of its last tile after reading them)
            tc::named_bar_sync(1, 256);
            for (int i = threadIdx.x; i < QP; i += 256) {
                // padded query columns: lse = +1e30 makes exp2(S c - lse) exactly 0 without a per-element mask (S is finite there:
                // rows of the next sequence or TMA zero fill)
                const float2 s = i < a.Lq ? reinterpret_cast<const float2*>(a.stats)[w * a.Lq + i] : make_float2(1e30f, 0.f);
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
                for_chunks(s_addr, c_begin, c_end, [&](int k, int c, const uint32_t (&va)[16]) {
                    float ls[16];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {  // broadcast reads
                        const float4 v4 = *reinterpret_cast<const float4*>(sLse + c + 4 * j);
                        ls[4 * j] = v4.x; ls[4 * j + 1] = v4.y; ls[4 * j + 2] = v4.z; ls[4 * j + 3] = v4.w;
                    }
                    float p[16];
#pragma unroll
                    for (int j = 0; j < 16; ++j) p[j] = tc::ex2_approx(fmaf(__uint_as_float(va[j]), a.sl2, -ls[j]));
#pragma unroll
                    for (int j = 0; j < 8; ++j) preg[8 * k + j] = hv_pack_bf16(p[2 * j], p[2 * j + 1]);
                });
                BWD_TICK(4);   // pass 2 (registers)
                // ---- the previous tile's dV / dK leave TMEM, and its dK MMA has finished READING the staging buffer, before this
                // tile's P^T goes there and this tile's dV MMA (issued on p_full) overwrites the accumulators
                if (pend) epilogue();
                BWD_TICK(9);   // epilogue (deferred)
#pragma unroll
                for (int k = 0; k < 9; ++k) {
                    const int c = c_begin + 16 * k;
                    if (c < c_end)
                        stage16_packed(Ps, row, c, make_uint4(preg[8 * k], preg[8 * k + 1], preg[8 * k + 2], preg[8 * k + 3]),
                                       make_uint4(preg[8 * k + 4], preg[8 * k + 5], preg[8 * k + 6], preg[8 * k + 7]));
                }
                BWD_TICK(3);   // stage P^T
                tc::fence_before_sync();
                tc::fence_proxy_async();
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(&p_full);
                // ---- dS^T = P^T (dP^T - D_q) scale, formed in place in the P^T registers while the dV MMA reads the staging buffer
                tc::mbar_wait(&dpt_full, ph);
                tc::fence_after_sync();
                BWD_TICK(6);   // wait dP^T
                for_chunks(s_addr, c_begin, c_end, [&](int k, int c, const uint32_t (&va)[16]) {
                    float dd[16];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const float4 v4 = *reinterpret_cast<const float4*>(sD + c + 4 * j);
                        dd[4 * j] = v4.x; dd[4 * j + 1] = v4.y; dd[4 * j + 2] = v4.z; dd[4 * j + 3] = v4.w;
                    }
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const float d0 = bf_lo(preg[8 * k + j]) * (__uint_as_float(va[2 * j]) - dd[2 * j]) * a.scale;
                        const float d1 = bf_hi(preg[8 * k + j]) * (__uint_as_float(va[2 * j + 1]) - dd[2 * j + 1]) * a.scale;
                        preg[8 * k + j] = hv_pack_bf16(d0, d1);
                    }
                });
                BWD_TICK(7);   // pass 3 (registers)
                tc::mbar_wait(&dv_done, ph);   // the dV MMA has read P^T: the staging buffer may take dS^T
                BWD_TICK(8);   // wait dV
#pragma unroll
                for (int k = 0; k < 9; ++k) {
                    const int c = c_begin + 16 * k;
                    if (c < c_end)
                        stage16_packed(Ps, row, c, make_uint4(preg[8 * k], preg[8 * k + 1], preg[8 * k + 2], preg[8 * k + 3]),
                                       make_uint4(preg[8 * k + 4], preg[8 * k + 5], preg[8 * k + 6], preg[8 * k + 7]));
                }
                tc::fence_before_sync();
                tc::fence_proxy_async();
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(&ds_full);
                BWD_TICK(5);   // stage dS^T
                pend = true;
                pend_valid = valid;
                pend_v = a.dv + (seq * a.kv_sa + kr) * a.lddv + head * HD + h * 32;
                pend_k = a.dk + (seq * a.kv_sa + kr) * a.lddk + head * HD + h * 32;
                if (prof) a.dbg[(KERNEL_ID * 148 + (blockIdx.x % 148)) * 16 + 15] += 1;
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
                             const HvSeqMap& qm, const HvSeqMap& km, float scale, float* stats, int have_lse, cudaStream_t st) {
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
    a.have_lse = have_lse;
    a.dbg = bwd_prof_buffer();
    const int KP32 = (a.KP + 31) & ~31, QP32 = (a.QP + 31) & ~31;
    const int smem_dq = 2 * TBYTES + 2 * KP32 * 128 + ((a.KP + 63) >> 6) * TBYTES + 1024;
    const int smem_dkv = 2 * QP32 * 128 + 2 * TBYTES + ((a.QP + 63) >> 6) * TBYTES + 1024;
    static int conf_dq_of[HV_MAX_DEVICES] = {}, conf_dkv_of[HV_MAX_DEVICES] = {};  // function attributes are per device
    int &conf_dq = conf_dq_of[hv_device_slot()], &conf_dkv = conf_dkv_of[hv_device_slot()];
    if (conf_dq < smem_dq) { HV_CUDA(cudaFuncSetAttribute(attn_bwd_dq_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_dq)); conf_dq = smem_dq; }
    if (conf_dkv < smem_dkv) { HV_CUDA(cudaFuncSetAttribute(attn_bwd_dkv_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_dkv)); conf_dkv = smem_dkv; }
    const int sms = hv_num_sms();
    const int grid = (int)(a.items < sms ? a.items : sms);
    const long long nd = seqs * Lq * heads;
    attn_bwd_delta_kernel<<<(unsigned)((nd + 255) / 256), 256, 0, st>>>(a.o, ldo, a.dout, lddo, stats, seqs, heads, Lq, qm.sa);
    HV_LAUNCH_CHECK();
    attn_bwd_dq_tc_kernel<<<grid, THREADS, smem_dq, st>>>(tq, tk, tv, tdo, a);
    HV_LAUNCH_CHECK();
    attn_bwd_dkv_tc_kernel<<<grid, THREADS, smem_dkv, st>>>(tq, tk, tv, tdo, a);
    HV_LAUNCH_CHECK();
    return 1;
}
