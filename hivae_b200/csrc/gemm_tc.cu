// bf16 tensor-core GEMM for sm_100a:  out = epilogue(A[M,K] @ W[N,K]^T), fp32 accumulation.
//
//   * operands staged by TMA (cp.async.bulk.tensor, SWIZZLE_128B) into a shared-memory ring,
//   * tcgen05.mma issued by one thread, accumulators in TMEM.  Large problems run as CTA PAIRS
//     (cta_group::2, one 256 x 256 x 16 MMA over two SMs): each CTA stages 128 rows of A and only
//     HALF of the B tile, which halves the shared-memory operand traffic per SM and leaves room for
//     the epilogue's staging traffic; small problems use one CTA per 128 x BN tile (cta_group::1),
//   * two TMEM accumulator stages so the epilogue of tile i overlaps the main loop of tile i+1,
//   * persistent CTAs walking tiles N-fastest so the weight matrix stays in L2 and every activation
//     tile is streamed from HBM once,
//   * 8 epilogue warps read TMEM with tcgen05.ld, apply bias / GELU-tanh / per-head QK LayerNorm in the
//     "thread = row" layout, and transpose 32x32 blocks through shared memory so that global accesses
//     are row-contiguous.  The gated residual epilogue (res += gate * (acc + bias)) never reads the
//     residual on the SM: the staged fp32 block is added into HBM/L2 by a TMA reduce
//     (cp.reduce.async.bulk.tensor .add) through a 3-D map (column, row in frame, frame) whose
//     out-of-range clipping handles the frame boundaries of the (frames, frame_rows, D) buffer.
//
// Warp roles (320 threads): warp 0 = TMA producer, warp 1 = MMA issuer + TMEM owner,
// warps 2..9 = epilogue (TMEM lane quarter = warp_idx % 4, column half = (warp_idx - 2) / 4).
#include <cuda.h>
#include <stdlib.h>
#include <string.h>
#include <mutex>
#include <string>
#include <unordered_map>
#include "epilogue.cuh"
#include "tc_ptx.cuh"

namespace {

constexpr int BM = 128, BK = 64;                              // rows of A / accumulator lanes per CTA, K per stage
constexpr int EPI_WARPS = 8, THREADS = 64 + 32 * EPI_WARPS;  // warp 0 TMA, warp 1 MMA, warps 2..9 epilogue
constexpr int SMEM_LIMIT = 232448 - 512;                     // 227 KB per CTA minus the static barriers

// One 32x32 fp32 transpose tile per epilogue warp; the residual (TMA reduce) epilogue keeps 32 zero rows behind
// each tile so that a reduce may start at any row of the tile (see epi_block_reduce).
// MODE_DUAL: internal flag on top of HV_EPI_GELU / HV_EPI_QKNORM — two bf16 outputs (pre-activation + activation) leave
// through two alternating staging tiles per warp, so a tile is only rewritten when the store issued two back has been read
constexpr int MODE_DUAL = 8;
template <int MODE> constexpr int staging_floats_per_warp() { return (MODE & ~MODE_DUAL) == HV_EPI_RESIDUAL ? 64 * 32 : ((MODE & MODE_DUAL) ? 2 * 32 * 32 : 32 * 32); }
template <int MODE> constexpr int staging_bytes() { return EPI_WARPS * staging_floats_per_warp<MODE>() * 4; }
// CTAS = 1: one CTA per 128 x BN tile.  CTAS = 2: a CTA pair per 256 x BN tile, BN/2 rows of B per CTA.
template <int BN, int CTAS> constexpr int stage_bytes() { return (BM + BN / CTAS) * BK * 2; }
template <int BN, int MODE, int CTAS> constexpr int num_stages() {
    return (SMEM_LIMIT - 1024 - staging_bytes<MODE>()) / stage_bytes<BN, CTAS>() > 8 ? 8 : (SMEM_LIMIT - 1024 - staging_bytes<MODE>()) / stage_bytes<BN, CTAS>();
}
template <int BN, int MODE, int CTAS> constexpr int smem_bytes() {
    return num_stages<BN, MODE, CTAS>() * stage_bytes<BN, CTAS>() + staging_bytes<MODE>() + 1024;
}

__device__ __forceinline__ float gelu_tanh_fast(float x) {
    // tanh.approx (MUFU): 2^-11 relative error, far below the bf16 rounding of the stored value
    const float k0 = 0.7978845608028654f, k1 = 0.044715f;
    float u = k0 * fmaf(k1 * x * x, x, x), t;
    asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(u));
    return 0.5f * x * (1.0f + t);
}

// 32x32 fp32 tile, 16-byte chunk c of row r at chunk c ^ (r & 7): conflict-free for "thread = row" float4
// writes and for "8 lanes = one row" float4 reads.
__device__ __forceinline__ float4* stage_ptr(float* stage, int r, int c) {
    return reinterpret_cast<float4*>(stage + r * 32 + ((c ^ (r & 7)) << 2));
}

// Epilogue of one 32-row x 32-column block owned by one warp.  `acc` holds the columns [col0, col0+32) of
// row (row0 + lane), bias / activation / QK-norm already applied.  The block is transposed through shared
// memory so that global accesses are row-contiguous (128 B per row for fp32, 64 B for bf16).
// Gated residual block through a TMA reduce: res[rows, col0:col0+32) += gate * acc.  The 32-row block may
// straddle one frame boundary of the residual buffer (L >= 32).  Rows of frame f: one reduce at row i0, the
// rows >= L are clipped by the tensor map (dims: N columns x L rows x frames).  Rows of frame f+1: a second
// reduce at row 0 whose SOURCE starts s = L - i0 rows into the staging tile (TMA's 128 B swizzle is a function
// of the absolute shared-memory address, so any row offset is consistent); its last s rows read the zero rows
// kept behind the tile and add 0.  (Negative start coordinates are not an option: they fault on sm_100.)
__device__ __forceinline__ void epi_block_reduce(const HvEpiDev& d, const CUtensorMap* tmap_r, float* stage, int lane,
                                                 long long row0, long long r, long long M, int col0, int N,
                                                 float (&acc)[32], long long gate_off) {
    const HvEpilogue& e = d.e;
    if (e.gate && r < M) {
        const float* g = e.gate + gate_off + col0;
        if (col0 + 32 <= N) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
                const float4 gg = __ldg(reinterpret_cast<const float4*>(g + j));
                acc[j] *= gg.x; acc[j + 1] *= gg.y; acc[j + 2] *= gg.z; acc[j + 3] *= gg.w;
            }
        } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) if (col0 + j < N) acc[j] *= __ldg(g + j);
        }
    }
    tc::bulk_wait_read0();  // the previous reduce has finished reading the staging tile (all lanes: only the issuing lane has groups)
    __syncwarp();
#pragma unroll
    for (int k = 0; k < 8; ++k)
        *stage_ptr(stage, lane, k) = make_float4(acc[4 * k], acc[4 * k + 1], acc[4 * k + 2], acc[4 * k + 3]);
    tc::fence_proxy_async();
    __syncwarp();
    {   // warp-converged: every lane forms the (warp-uniform) coordinates, one elected lane issues the reduce(s) and commits
        const long long f0 = row0 / e.L;
        const int i0 = (int)(row0 - f0 * e.L), s = e.L - i0;
        const uint32_t st = tc::warp_uniform(tc::smem_u32(stage));
        const int c0 = (int)tc::warp_uniform((uint32_t)col0), f = (int)tc::warp_uniform((uint32_t)f0);
        const int ii = (int)tc::warp_uniform((uint32_t)i0), ss = (int)tc::warp_uniform((uint32_t)s);
        const uint32_t do1 = tc::warp_uniform(row0 < M ? 1u : 0u), do2 = tc::warp_uniform((s < 32 && row0 + s < M) ? 1u : 0u);
        tc::tma_reduce_add_3d_x2_commit_w(tmap_r, st, c0, ii, f, do1, st + (uint32_t)(ss * 32 * 4), c0, 0, f + 1, do2);
    }
}

// bf16 output block of 32 rows x 64 columns through a TMA store: every thread packs the 64 columns of its row
// (128 B) into the warp's staging tile in the 128 B swizzle layout (conflict-free 16-byte writes), one lane
// issues cp.async.bulk.tensor.  No shared-memory read-back and no per-lane global stores; rows >= M and
// columns >= N are clipped by the tensor map.
template <int PENDING = 0>
__device__ __forceinline__ void epi_block_tma_store(const CUtensorMap* tmap_c, float* stage, int lane, long long row0,
                                                    int col0, const float (&lo)[32], const float (&hi)[32]) {
    // the previous store from THIS staging tile has finished reading it (all lanes: only the issuing lane has groups)
    if (PENDING == 0) tc::bulk_wait_read0(); else tc::bulk_wait_read1();
    __syncwarp();
    uint8_t* rowp = reinterpret_cast<uint8_t*>(stage) + lane * 128;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        *reinterpret_cast<uint4*>(rowp + ((k ^ (lane & 7)) << 4)) =
            make_uint4(hv_pack_bf16(lo[8 * k], lo[8 * k + 1]), hv_pack_bf16(lo[8 * k + 2], lo[8 * k + 3]),
                       hv_pack_bf16(lo[8 * k + 4], lo[8 * k + 5]), hv_pack_bf16(lo[8 * k + 6], lo[8 * k + 7]));
        *reinterpret_cast<uint4*>(rowp + (((k + 4) ^ (lane & 7)) << 4)) =
            make_uint4(hv_pack_bf16(hi[8 * k], hi[8 * k + 1]), hv_pack_bf16(hi[8 * k + 2], hi[8 * k + 3]),
                       hv_pack_bf16(hi[8 * k + 4], hi[8 * k + 5]), hv_pack_bf16(hi[8 * k + 6], hi[8 * k + 7]));
    }
    tc::fence_proxy_async();
    __syncwarp();
    tc::tma_store_2d_commit_w(tmap_c, tc::warp_uniform(tc::smem_u32(stage)), (int)tc::warp_uniform((uint32_t)col0),
                              (int)tc::warp_uniform((uint32_t)row0));  // one elected lane stores and commits
}

template <int MODE>
__device__ __forceinline__ void epi_block_out(const HvEpiDev& d, float* stage, int lane, long long row0, long long M,
                                              int col0, int N, const float (&acc)[32], long long res_off, long long gate_off) {
    const HvEpilogue& e = d.e;
#pragma unroll
    for (int k = 0; k < 8; ++k)
        *stage_ptr(stage, lane, k) = make_float4(acc[4 * k], acc[4 * k + 1], acc[4 * k + 2], acc[4 * k + 3]);
    __syncwarp();
    const bool full = d.vec_ok && (col0 + 32 <= N);
    if (MODE == HV_EPI_RESIDUAL) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int rr = 4 * i + (lane >> 3), cc = lane & 7;
            const long long ro = __shfl_sync(0xffffffffu, res_off, rr);
            const long long go = __shfl_sync(0xffffffffu, gate_off, rr);
            const float4 v = *stage_ptr(stage, rr, cc);
            const int col = col0 + 4 * cc;
            if (row0 + rr < M) {
                float* o = e.res + ro + col;
                const float* g = e.gate ? e.gate + go + col : nullptr;
                if (full) {
                    float4 x = *reinterpret_cast<float4*>(o);
                    const float4 gg = g ? __ldg(reinterpret_cast<const float4*>(g)) : make_float4(1.f, 1.f, 1.f, 1.f);
                    x.x = fmaf(gg.x, v.x, x.x); x.y = fmaf(gg.y, v.y, x.y); x.z = fmaf(gg.z, v.z, x.z); x.w = fmaf(gg.w, v.w, x.w);
                    *reinterpret_cast<float4*>(o) = x;
                } else {
                    const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                    for (int j = 0; j < 4; ++j) if (col + j < N) o[j] = fmaf(g ? g[j] : 1.f, vv[j], o[j]);
                }
            }
        }
    } else if (e.out_dtype == HV_F32) {
        float* out = reinterpret_cast<float*>(e.out);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int rr = 4 * i + (lane >> 3), cc = lane & 7;
            const float4 v = *stage_ptr(stage, rr, cc);
            const int col = col0 + 4 * cc;
            if (row0 + rr < M) {
                float* o = out + (row0 + rr) * e.ldo + col;
                if (full) {
                    *reinterpret_cast<float4*>(o) = v;
                } else {
                    const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                    for (int j = 0; j < 4; ++j) if (col + j < N) o[j] = vv[j];
                }
            }
        }
    } else {
        __nv_bfloat16* out = reinterpret_cast<__nv_bfloat16*>(e.out);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int rr = 8 * i + (lane >> 2), cc = lane & 3;
            const float4 v0 = *stage_ptr(stage, rr, 2 * cc), v1 = *stage_ptr(stage, rr, 2 * cc + 1);
            const int col = col0 + 8 * cc;
            if (row0 + rr < M) {
                __nv_bfloat16* o = out + (row0 + rr) * e.ldo + col;
                if (full) {
                    *reinterpret_cast<uint4*>(o) = make_uint4(hv_pack_bf16(v0.x, v0.y), hv_pack_bf16(v0.z, v0.w),
                                                              hv_pack_bf16(v1.x, v1.y), hv_pack_bf16(v1.z, v1.w));
                } else {
                    const float vv[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
                    for (int j = 0; j < 8; ++j) if (col + j < N) o[j] = __float2bfloat16(vv[j]);
                }
            }
        }
    }
    __syncwarp();  // the staging tile is reused by the next block
}

__device__ __forceinline__ void epi_add_bias(const HvEpiDev& d, int col0, int N, float (&acc)[32]) {
    const float* bias = d.e.bias;
    if (!bias) return;
    if (d.vec_ok && col0 + 32 <= N) {
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
            const float4 b = __ldg(reinterpret_cast<const float4*>(bias + col0 + j));
            acc[j] += b.x; acc[j + 1] += b.y; acc[j + 2] += b.z; acc[j + 3] += b.w;
        }
    } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) if (col0 + j < N) acc[j] += __ldg(bias + col0 + j);
    }
}

template <int BN, int MODEX, int CTAS>
__global__ void __launch_bounds__(THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
               const __grid_constant__ CUtensorMap tmap_r, const __grid_constant__ CUtensorMap tmap_r2, long long M, int N, int K,
               int use_reduce, int tn, const __grid_constant__ HvEpiDev epi) {
    // tn != 0: "transposed-operand" product for weight gradients, out[M x N] = At^T Bt with At (K x M) and Bt (K x N) both
    // row-major (K = tokens): the operands are MN-major in shared memory — TMA boxes of [64 tokens][64 columns = 128 B] —
    // so dW = dY^T X needs no transposed copies of dY and X (round 1 paid three bf16 transposes per Linear)
    constexpr int MODE = MODEX & ~MODE_DUAL;
    constexpr bool DUAL = (MODEX & MODE_DUAL) != 0;  // second (pre-activation) output through tmap_r2; TMA-store path only
    constexpr int STAGES = num_stages<BN, MODEX, CTAS>();
    constexpr int STAGE_BYTES = stage_bytes<BN, CTAS>();
    constexpr int BNL = BN / CTAS;     // rows of the B tile staged by this CTA
    constexpr int TILE_M = BM * CTAS;  // rows of the output tile of the CTA (pair)
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[STAGES], empty_bar[STAGES], tmem_full_bar[2], tmem_empty_bar[2];
    __shared__ uint32_t tmem_slot;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = CTAS == 2 ? tc::cluster_ctarank() : 0u;  // 0 = leader CTA of the pair
    const long long unit = blockIdx.x / CTAS, units = gridDim.x / CTAS;
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    auto smem_a = [&](int s) { return smem + s * STAGE_BYTES; };
    auto smem_b = [&](int s) { return smem + s * STAGE_BYTES + BM * BK * 2; };

    const int n_blocks = (N + BN - 1) / BN;
    const long long m_blocks = (M + TILE_M - 1) / TILE_M;
    const long long num_tiles = m_blocks * n_blocks;
    const int num_kb = (K + BK - 1) / BK;
    // split-K (reduce-add residual epilogue only; the host guarantees every split is non-empty): work item t is split
    // t / num_tiles of tile t % num_tiles and accumulates k blocks [ks * kb_per, min(num_kb, (ks + 1) * kb_per))
    const int ksplit = (MODE == HV_EPI_RESIDUAL && epi.e.k_split > 1) ? epi.e.k_split : 1;
    const int kb_per = (num_kb + ksplit - 1) / ksplit;
    const long long num_items = num_tiles * ksplit;
    constexpr uint32_t TMEM_COLS = 2 * BN;  // 256 or 512: power of two

    if (threadIdx.x == 0) {
        tc::prefetch_tmap(&tmap_a);
        tc::prefetch_tmap(&tmap_b);
        if (use_reduce) tc::prefetch_tmap(&tmap_r);
        if (DUAL) tc::prefetch_tmap(&tmap_r2);
#pragma unroll
        for (int s = 0; s < STAGES; ++s) { tc::mbar_init(&full_bar[s], 1); tc::mbar_init(&empty_bar[s], 1); }
#pragma unroll
        for (int a = 0; a < 2; ++a) { tc::mbar_init(&tmem_full_bar[a], 1); tc::mbar_init(&tmem_empty_bar[a], EPI_WARPS * CTAS); }
        tc::fence_barrier_init();
    }
    if (warp == 1) {
        if (CTAS == 2) tc::tmem_alloc_pair(&tmem_slot, TMEM_COLS); else tc::tmem_alloc(&tmem_slot, TMEM_COLS);
    }
    tc::fence_before_sync();
    if (CTAS == 2) tc::cluster_sync(); else __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem_base = tmem_slot;
    tc::grid_dep_wait();     // PDL: everything above overlapped the previous kernel's tail; global memory is touched only below
    tc::grid_dep_launch();   // persistent grid, fully resident: the next kernel's CTAs may take the SMs as these CTAs exit

    if (warp == 0) {
        // ===================================================== TMA producer (one per CTA).  All 32 lanes run the loop with
        // warp-uniform values and one elected lane issues (tc_ptx.cuh *_w): under `if (lane == 0)` ptxas wraps every TMA /
        // MMA instruction in an ELECT + R2UR.BROADCAST loop (270 clk per MMA against 65, tools/probes/mma_probe.cu)
        {
            int s = 0; uint32_t ph = 0;
            const uint32_t sbase = tc::smem_u32(smem);
            for (long long it = unit; it < num_items; it += units) {
                const long long t = it % num_tiles;
                const int kb0 = (int)tc::warp_uniform((uint32_t)((int)(it / num_tiles) * kb_per)), kb1 = kb0 + kb_per < num_kb ? kb0 + kb_per : num_kb;
                const int m0 = (int)tc::warp_uniform((uint32_t)((int)(t / n_blocks) * TILE_M + (int)rank * BM));
                const int n0 = (int)tc::warp_uniform((uint32_t)((int)(t % n_blocks) * BN + (int)rank * BNL));
                for (int kb = kb0; kb < kb1; ++kb) {
                    tc::mbar_wait(&empty_bar[s], ph ^ 1);
                    const uint32_t sa = sbase + (uint32_t)(s * STAGE_BYTES), sb = sa + BM * BK * 2;
                    if (tn) {  // boxes of [64 k rows][64 columns]: 2 for the 128 rows of A, BNL / 64 for this CTA's share of B
                        if (CTAS == 2) {
                            const uint32_t bar = tc::mapa(tc::smem_u32(&full_bar[s]), 0);
                            if (rank == 0) tc::mbar_expect_tx_w(&full_bar[s], 2 * STAGE_BYTES);
#pragma unroll
                            for (int b = 0; b < BM / 64; ++b) tc::tma_load_2d_pair_w(sa + b * 8192, &tmap_a, bar, m0 + 64 * b, kb * BK);
#pragma unroll
                            for (int b = 0; b < BNL / 64; ++b) tc::tma_load_2d_pair_w(sb + b * 8192, &tmap_b, bar, n0 + 64 * b, kb * BK);
                        } else {
                            tc::mbar_expect_tx_w(&full_bar[s], STAGE_BYTES);
#pragma unroll
                            for (int b = 0; b < BM / 64; ++b) tc::tma_load_2d_w(sa + b * 8192, &tmap_a, &full_bar[s], m0 + 64 * b, kb * BK);
#pragma unroll
                            for (int b = 0; b < BNL / 64; ++b) tc::tma_load_2d_w(sb + b * 8192, &tmap_b, &full_bar[s], n0 + 64 * b, kb * BK);
                        }
                    } else if (CTAS == 2) {
                        // both CTAs' bytes complete on the leader's barrier, which the leader armed for the pair
                        const uint32_t bar = tc::mapa(tc::smem_u32(&full_bar[s]), 0);
                        if (rank == 0) tc::mbar_expect_tx_w(&full_bar[s], 2 * STAGE_BYTES);
                        tc::tma_load_2d_pair_w(sa, &tmap_a, bar, kb * BK, m0);
                        tc::tma_load_2d_pair_w(sb, &tmap_b, bar, kb * BK, n0);
                    } else {
                        tc::mbar_expect_tx_w(&full_bar[s], STAGE_BYTES);
                        tc::tma_load_2d_w(sa, &tmap_a, &full_bar[s], kb * BK, m0);
                        tc::tma_load_2d_w(sb, &tmap_b, &full_bar[s], kb * BK, n0);
                    }
                    if (++s == STAGES) { s = 0; ph ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================================================== MMA issuer (leader CTA only; warp-converged, one elected lane issues)
        if (rank == 0) {
            const uint32_t idesc = tc::make_idesc_bf16(TILE_M, BN) | (tn ? (3u << 15) : 0u);  // bits 15 / 16: A / B are MN-major
            const uint32_t tbase = tc::warp_uniform(tmem_base), sbase = tc::smem_u32(smem);
            int s = 0; uint32_t ph = 0; int a = 0; uint32_t aph = 0;
            for (long long it = unit; it < num_items; it += units) {
                const int kb0 = (int)tc::warp_uniform((uint32_t)((int)(it / num_tiles) * kb_per)), kb1 = kb0 + kb_per < num_kb ? kb0 + kb_per : num_kb;
                if (CTAS == 2) tc::mbar_wait_cluster(&tmem_empty_bar[a], aph ^ 1); else tc::mbar_wait(&tmem_empty_bar[a], aph ^ 1);
                tc::fence_after_sync();
                const uint32_t d_tmem = tbase + (uint32_t)(a * BN);
                for (int kb = kb0; kb < kb1; ++kb) {
                    tc::mbar_wait(&full_bar[s], ph);
                    tc::fence_after_sync();
                    const uint32_t sa = sbase + (uint32_t)(s * STAGE_BYTES), sb = sa + BM * BK * 2;
                    const uint64_t da = tn ? tc::make_smem_desc_mn128(sa, 8192) : tc::make_smem_desc_k128(sa);
                    const uint64_t db = tn ? tc::make_smem_desc_mn128(sb, 8192) : tc::make_smem_desc_k128(sb);
                    const uint64_t kstep = tn ? 128 : 2;  // K = 16 step: 16 rows x 128 B (MN-major) / 32 B inside the swizzle row (K-major), in 16-byte units
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k) {
                        if (CTAS == 2) tc::umma_bf16_pair_w(d_tmem, da + kstep * k, db + kstep * k, idesc, ((kb - kb0) | k) != 0);
                        else tc::umma_bf16_w(d_tmem, da + kstep * k, db + kstep * k, idesc, ((kb - kb0) | k) != 0);
                    }
                    // frees the smem stage (in both CTAs) once these MMAs retire
                    if (CTAS == 2) tc::umma_commit_pair_w(&empty_bar[s]); else tc::umma_commit_w(&empty_bar[s]);
                    if (++s == STAGES) { s = 0; ph ^= 1; }
                }
                if (CTAS == 2) tc::umma_commit_pair_w(&tmem_full_bar[a]); else tc::umma_commit_w(&tmem_full_bar[a]);  // accumulator complete
                if ((a ^= 1) == 0) aph ^= 1;
            }
        }
    } else {
        // ===================================================== epilogue warps
        const int ew = warp - 2;
        const int q = warp & 3;    // TMEM lane quarter this warp may access (hardware rule: warp id % 4)
        const int half = ew >> 2;  // which half of the tile's columns
        constexpr int HALF = BN / 2;
        float* stage = reinterpret_cast<float*>(smem + STAGES * STAGE_BYTES) + ew * staging_floats_per_warp<MODEX>();
        const HvEpilogue& e = epi.e;
        const bool reduce = MODE == HV_EPI_RESIDUAL && use_reduce;
        const bool tma_store = MODE != HV_EPI_RESIDUAL && use_reduce;  // tmap_r is then the bf16 OUTPUT map (box 64 x 32)
        if (MODE == HV_EPI_RESIDUAL) {  // the zero rows behind the tile
#pragma unroll
            for (int k = 0; k < 8; ++k) reinterpret_cast<float4*>(stage + 32 * 32)[k * 32 + lane] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        int a = 0; uint32_t aph = 0;
        for (long long it = unit; it < num_items; it += units) {
            const long long t = it % num_tiles;
            const long long m0 = (t / n_blocks) * TILE_M + (long long)rank * BM;
            const int n0 = (int)(t % n_blocks) * BN + half * HALF;
            const long long row0 = m0 + q * 32, r = row0 + lane;
            long long res_off = 0, gate_off = 0;
            if (MODE == HV_EPI_RESIDUAL && r < M) {
                HvRowMap map{e.L, (long long)e.frame_rows, e.row_off};
                int frame, abs_i;
                res_off = map.res_row(r, frame, abs_i) * N;
                if (e.gate) {
                    const bool img = (abs_i >= e.img_start) && (abs_i < e.img_start + e.img_len);
                    gate_off = (long long)((frame / e.cond_div) * e.cond_mul) * e.gate_ld + (img ? e.gate_col_img : e.gate_col_mot);
                }
            }
            if (MODE == HV_EPI_RESIDUAL && !reduce) {
                // pull this warp's 32 x HALF residual block into L2 while the tile's MMAs are still running
#pragma unroll
                for (int i = 0; i < HALF / 32; ++i) {
                    const int li = lane + 32 * i, rr = li / (HALF / 32), seg = li % (HALF / 32);
                    const long long ro = __shfl_sync(0xffffffffu, res_off, rr);
                    if (row0 + rr < M && n0 + seg * 32 < N)
                        asm volatile("prefetch.global.L2 [%0];" ::"l"(e.res + ro + n0 + seg * 32));
                }
            }
            tc::mbar_wait(&tmem_full_bar[a], aph);
            tc::fence_after_sync();
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(a * BN + half * HALF);
            if (MODE == HV_EPI_QKNORM) {
#pragma unroll 1
                for (int c = 0; c < HALF; c += 64) {
                    const int col0 = n0 + c;
                    if (col0 >= N) break;  // warp-uniform
                    float lo[32], hi[32];
                    tc::tmem_ld32(taddr + c, lo);
                    tc::tmem_ld32(taddr + c + 32, hi);
                    epi_add_bias(epi, col0, N, lo);
                    epi_add_bias(epi, col0 + 32, N, hi);
                    if (DUAL) epi_block_tma_store<1>(&tmap_r2, stage + 32 * 32, lane, row0, col0, lo, hi);  // pre-norm q / k (and v)
                    const float *w = nullptr, *b = nullptr;
                    if (col0 < e.q_cols) { w = e.qn_w; b = e.qn_b; }
                    else if (col0 < e.q_cols + e.k_cols) { w = e.kn_w; b = e.kn_b; }
                    if (w) {  // per-head LayerNorm over the 64 columns of this thread's row
                        float s = 0.f;
#pragma unroll
                        for (int j = 0; j < 32; ++j) s += lo[j] + hi[j];
                        const float mean = s * (1.f / 64.f);
                        float v = 0.f;
#pragma unroll
                        for (int j = 0; j < 32; ++j) { float t0 = lo[j] - mean, t1 = hi[j] - mean; v = fmaf(t0, t0, v); v = fmaf(t1, t1, v); }
                        const float rstd = rsqrtf(v * (1.f / 64.f) + e.qk_eps);
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            lo[j] = (lo[j] - mean) * rstd * __ldg(w + j) + __ldg(b + j);
                            hi[j] = (hi[j] - mean) * rstd * __ldg(w + 32 + j) + __ldg(b + 32 + j);
                        }
                    }
                    if (tma_store) {
                        epi_block_tma_store<DUAL ? 1 : 0>(&tmap_r, stage, lane, row0, col0, lo, hi);
                    } else {
                        epi_block_out<MODE>(epi, stage, lane, row0, M, col0, N, lo, 0, 0);
                        epi_block_out<MODE>(epi, stage, lane, row0, M, col0 + 32, N, hi, 0, 0);
                    }
                }
            } else if (tma_store) {  // STORE / GELU with a bf16 output: 64 columns per TMA store
#pragma unroll 1
                for (int c = 0; c < HALF; c += 64) {
                    const int col0 = n0 + c;
                    if (col0 >= N) break;  // warp-uniform
                    float lo[32], hi[32];
                    tc::tmem_ld32(taddr + c, lo);
                    tc::tmem_ld32(taddr + c + 32, hi);
                    epi_add_bias(epi, col0, N, lo);
                    epi_add_bias(epi, col0 + 32, N, hi);
                    if (DUAL) epi_block_tma_store<1>(&tmap_r2, stage + 32 * 32, lane, row0, col0, lo, hi);  // pre-GELU activations
                    if (MODE == HV_EPI_GELU) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) { lo[j] = gelu_tanh_fast(lo[j]); hi[j] = gelu_tanh_fast(hi[j]); }
                    }
                    epi_block_tma_store<DUAL ? 1 : 0>(&tmap_r, stage, lane, row0, col0, lo, hi);
                }
            } else {
#pragma unroll 1
                for (int c = 0; c < HALF; c += 32) {
                    const int col0 = n0 + c;
                    if (col0 >= N) break;  // warp-uniform
                    float acc[32];
                    tc::tmem_ld32(taddr + c, acc);
                    epi_add_bias(epi, col0, N, acc);
                    if (MODE == HV_EPI_GELU) {
                        if (e.out_dtype == HV_BF16) {
#pragma unroll
                            for (int j = 0; j < 32; ++j) acc[j] = gelu_tanh_fast(acc[j]);
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; ++j) acc[j] = hv_gelu_tanh(acc[j]);
                        }
                    }
                    if (reduce) epi_block_reduce(epi, &tmap_r, stage, lane, row0, r, M, col0, N, acc, gate_off);
                    else epi_block_out<MODE>(epi, stage, lane, row0, M, col0, N, acc, res_off, gate_off);
                }
            }
            tc::fence_before_sync();
            __syncwarp();
            if (lane == 0) {
                if (CTAS == 2) tc::mbar_arrive_cluster(tc::mapa(tc::smem_u32(&tmem_empty_bar[a]), 0));
                else tc::mbar_arrive(&tmem_empty_bar[a]);
            }
            if ((a ^= 1) == 0) aph ^= 1;
        }
        if (reduce || tma_store) tc::bulk_wait0();  // every bulk store / reduce of this warp has landed before the CTA retires
    }
    tc::fence_before_sync();
    if (CTAS == 2) tc::cluster_sync(); else __syncthreads();
    if (warp == 1) {
        tc::fence_after_sync();
        if (CTAS == 2) tc::tmem_dealloc_pair(tmem_base, TMEM_COLS); else tc::tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

// ------------------------------------------------------------------------------------------ host
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    });
    return fn;
}

struct MapKey {
    const void* ptr; long long rows, ld; int K, box_rows;
    bool operator==(const MapKey& o) const { return ptr == o.ptr && rows == o.rows && ld == o.ld && K == o.K && box_rows == o.box_rows; }
};
struct MapKeyHash {
    size_t operator()(const MapKey& k) const {
        size_t h = reinterpret_cast<size_t>(k.ptr);
        h ^= (size_t)k.rows * 0x9E3779B97F4A7C15ull + (size_t)k.ld * 0xC2B2AE3D27D4EB4Full + (size_t)k.K * 31 + (size_t)k.box_rows;
        return h;
    }
};

}  // namespace

// K-major bf16 matrix [rows, K] with row stride ld -> 2-D tensor map, box = 64 (K) x box_rows, 128 B swizzle
// (shared with attn_tc.cu)
int hv_make_tmap_bf16(const void* ptr, long long rows, int K, long long ld, int box_rows, CUtensorMap* out) {
    static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> cache;
    static std::mutex mu;
    MapKey key{ptr, rows, ld, K, box_rows};
    {
        std::lock_guard<std::mutex> g(mu);
        auto it = cache.find(key);
        if (it != cache.end()) { *out = it->second; return HV_OK; }
    }
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) { hv_set_error("cuTensorMapEncodeTiled entry point not available (driver too old?)"); return HV_ECUDA; }
    cuuint64_t gdim[2] = {(cuuint64_t)K, (cuuint64_t)rows};
    cuuint64_t gstride[1] = {(cuuint64_t)ld * 2};
    cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult rc = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), gdim, gstride, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (rc != CUDA_SUCCESS) { hv_set_error("cuTensorMapEncodeTiled failed with CUresult %d (rows=%lld K=%d ld=%lld)", (int)rc, rows, K, ld); return HV_ECUDA; }
    std::lock_guard<std::mutex> g(mu);
    if (cache.size() > 8192) cache.clear();
    cache.emplace(key, *out);
    return HV_OK;
}

// fp32 residual rows [row_off, row_off+L) of every frame of a (frames, frame_rows, N) buffer as a 3-D map
// (N columns, L rows, frames); box = 32 columns x 32 rows x 1 frame, 128 B swizzle (= the staging layout)
static int make_tmap_residual(float* res, int N, int L, long long frames, long long frame_rows, int row_off, CUtensorMap* out) {
    struct Key { const void* p; int N, L, off; long long frames, fr; };
    static std::unordered_map<std::string, CUtensorMap> cache;
    static std::mutex mu;
    Key key;
    memset(&key, 0, sizeof(key));  // padding bytes are part of the cache key
    key.p = res; key.N = N; key.L = L; key.off = row_off; key.frames = frames; key.fr = frame_rows;
    std::string ks(reinterpret_cast<const char*>(&key), sizeof(key));
    {
        std::lock_guard<std::mutex> g(mu);
        auto it = cache.find(ks);
        if (it != cache.end()) { *out = it->second; return HV_OK; }
    }
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) { hv_set_error("cuTensorMapEncodeTiled entry point not available (driver too old?)"); return HV_ECUDA; }
    cuuint64_t gdim[3] = {(cuuint64_t)N, (cuuint64_t)L, (cuuint64_t)frames};
    cuuint64_t gstride[2] = {(cuuint64_t)N * 4, (cuuint64_t)frame_rows * N * 4};
    cuuint32_t box[3] = {32, 32, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult rc = fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, res + (long long)row_off * N, gdim, gstride, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (rc != CUDA_SUCCESS) { hv_set_error("cuTensorMapEncodeTiled (residual) failed with CUresult %d", (int)rc); return HV_ECUDA; }
    std::lock_guard<std::mutex> g(mu);
    if (cache.size() > 8192) cache.clear();
    cache.emplace(ks, *out);
    return HV_OK;
}

namespace {

// CTA pairs that can be co-resident (one CTA per SM, clusters of 2 inside a TPC)
template <int BN, int MODE>
int max_pairs() {
    static int pairs_of[HV_MAX_DEVICES] = {};  // 0 = not computed yet, per device
    int& pairs = pairs_of[hv_device_slot()];
    if (pairs <= 0) {
        cudaLaunchConfig_t cfg = {};
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.gridDim = dim3(2 * (hv_num_sms() / 2)); cfg.blockDim = dim3(THREADS);
        cfg.dynamicSmemBytes = smem_bytes<BN, MODE, 2>(); cfg.attrs = at; cfg.numAttrs = 1;
        int n = 0;
        if (cudaOccupancyMaxActiveClusters(&n, gemm_tc_kernel<BN, MODE, 2>, &cfg) != cudaSuccess || n <= 0) {
            cudaGetLastError();
            n = hv_num_sms() / 2;
        }
        pairs = n < hv_num_sms() / 2 ? n : hv_num_sms() / 2;
    }
    return pairs;
}

template <int BN, int MODE, int CTAS>
int launch(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tr, const CUtensorMap& tr2, int use_reduce, int64_t M, int N,
           int K, const HvEpiDev& d, cudaStream_t st, int tn = 0) {
    auto kern = gemm_tc_kernel<BN, MODE, CTAS>;
    static bool configured_of[HV_MAX_DEVICES] = {};  // function attributes are per device
    bool& configured = configured_of[hv_device_slot()];
    if (!configured) {
        HV_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes<BN, MODE, CTAS>()));
        configured = true;
    }
    const long long tiles = ((M + BM * CTAS - 1) / (BM * CTAS)) * ((N + BN - 1) / BN) * (d.e.k_split > 1 ? d.e.k_split : 1);
    if constexpr (CTAS == 1) {
        const int sms = hv_num_sms();
        const int grid = (int)(tiles < sms ? tiles : sms);
        cudaLaunchConfig_t cfg = {};
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.gridDim = dim3(grid); cfg.blockDim = dim3(THREADS);
        cfg.dynamicSmemBytes = smem_bytes<BN, MODE, CTAS>(); cfg.stream = st; cfg.attrs = at; cfg.numAttrs = hv_pdl_enabled() ? 1 : 0;
        HV_CUDA(cudaLaunchKernelEx(&cfg, kern, ta, tb, tr, tr2, (long long)M, N, K, use_reduce, tn, d));
    } else {
        const int pairs = max_pairs<BN, MODE>();
        const int units = (int)(tiles < pairs ? tiles : pairs);
        cudaLaunchConfig_t cfg = {};
        cudaLaunchAttribute at[2];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[1].val.programmaticStreamSerializationAllowed = 1;
        cfg.gridDim = dim3(2 * units); cfg.blockDim = dim3(THREADS);
        cfg.dynamicSmemBytes = smem_bytes<BN, MODE, CTAS>(); cfg.stream = st; cfg.attrs = at; cfg.numAttrs = hv_pdl_enabled() ? 2 : 1;
        HV_CUDA(cudaLaunchKernelEx(&cfg, kern, ta, tb, tr, tr2, (long long)M, N, K, use_reduce, tn, d));
    }
    HV_LAUNCH_CHECK();
    return HV_OK;
}

template <int MODE>
int launch_mode(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tr, const CUtensorMap& tr2, int use_reduce, int64_t M,
                int N, int K, const HvEpiDev& d, cudaStream_t st, int BN, int ctas, int tn = 0) {
    if (ctas == 2) return launch<256, MODE, 2>(ta, tb, tr, tr2, use_reduce, M, N, K, d, st, tn);
    if (BN == 256) return launch<256, MODE, 1>(ta, tb, tr, tr2, use_reduce, M, N, K, d, st, tn);
    if constexpr ((MODE & MODE_DUAL) != 0) { hv_set_error("hv_linear_bf16: dual output needs BN = 256"); return HV_EINVAL; }
    else return launch<128, MODE, 1>(ta, tb, tr, tr2, use_reduce, M, N, K, d, st, tn);
}

// HV_GEMM_CTAS=1 forces the single-CTA kernel, HV_GEMM_REDUCE=0 the read-modify-write residual epilogue (A/B testing)
int env_int(const char* name, int dflt) {
    const char* v = getenv(name);
    return v && *v ? atoi(v) : dflt;
}

}  // namespace

extern "C" int hv_linear_bf16(const void* A, int64_t lda, const void* W, int64_t ldw,
                              int64_t M, int N, int K, const HvEpilogue* epi, void* stream) {
    HV_REQUIRE(M >= 0 && N > 0 && K > 0 && lda >= K && ldw >= K, "hv_linear_bf16: bad sizes");
    HV_REQUIRE(K % 8 == 0 && lda % 8 == 0 && ldw % 8 == 0, "hv_linear_bf16: K, lda, ldw must be multiples of 8 (16-byte TMA strides)");
    HV_REQUIRE((reinterpret_cast<uintptr_t>(A) & 15) == 0 && (reinterpret_cast<uintptr_t>(W) & 15) == 0, "hv_linear_bf16: A and W must be 16-byte aligned");
    HV_REQUIRE(M < (1ll << 31), "hv_linear_bf16: M too large");
    HvEpiDev d;
    int rc = hv_epilogue_prepare(epi, M, N, d, "hv_linear_bf16");
    if (rc) return rc;
    if (M == 0) return HV_OK;
    static const int want_ctas = env_int("HV_GEMM_CTAS", 2), want_reduce = env_int("HV_GEMM_REDUCE", 1),
                     want_tma_store = env_int("HV_GEMM_TMA_STORE", 1);
    const int BN = N > 128 ? 256 : 128;
    // a CTA pair pays off once there are enough 256 x 256 tiles to occupy the pairs
    const int ctas = (want_ctas == 2 && N > 128 && M > 256) ? 2 : 1;
    CUtensorMap ta, tb, tr;
    if ((rc = hv_make_tmap_bf16(A, M, K, lda, BM, &ta)) != HV_OK) return rc;
    if ((rc = hv_make_tmap_bf16(W, N, K, ldw, BN / ctas, &tb)) != HV_OK) return rc;
    int use_reduce = 0;
    if (d.e.mode == HV_EPI_RESIDUAL && want_reduce && d.vec_ok && d.e.L >= 32) {
        if ((rc = make_tmap_residual(d.e.res, N, d.e.L, M / d.e.L, d.e.frame_rows, d.e.row_off, &tr)) != HV_OK) return rc;
        use_reduce = 1;
    } else if (d.e.mode != HV_EPI_RESIDUAL && want_tma_store && d.vec_ok && d.e.out_dtype == HV_BF16 && BN == 256) {
        // bf16 outputs leave through TMA stores: 64-column x 32-row boxes of the output matrix (128 B swizzle)
        if ((rc = hv_make_tmap_bf16(d.e.out, M, N, d.e.ldo, 32, &tr)) != HV_OK) return rc;
        use_reduce = 1;
    } else {
        tr = ta;  // unused
    }
    // split-K needs the atomic TMA reduce-add epilogue and nothing that must be applied exactly once
    if (d.e.k_split > 1) {
        HV_REQUIRE(d.e.mode == HV_EPI_RESIDUAL && !d.e.bias && !d.e.gate, "hv_linear_bf16: k_split needs a plain residual epilogue (no bias, no gate)");
        const int num_kb = (K + 63) / 64;
        int S = (use_reduce && d.e.k_split <= num_kb) ? d.e.k_split : (use_reduce ? num_kb : 1);
        const int per = (num_kb + S - 1) / S;
        d.e.k_split = (num_kb + per - 1) / per;  // every split owns at least one k block
    } else {
        d.e.k_split = 1;
    }
    cudaStream_t st = hv_stream(stream);
    if (d.e.out2) {
        // second (pre-activation) output: fused into the TMA-store epilogue when that path applies, otherwise an extra
        // plain-store launch of the same GEMM
        const bool tma2 = use_reduce && d.e.mode != HV_EPI_RESIDUAL && d.vec_ok2;
        if (tma2) {
            CUtensorMap tr2;
            if ((rc = hv_make_tmap_bf16(d.e.out2, M, N, d.e.ldo2, 32, &tr2)) != HV_OK) return rc;
            if (d.e.mode == HV_EPI_GELU) return launch_mode<HV_EPI_GELU | MODE_DUAL>(ta, tb, tr, tr2, use_reduce, M, N, K, d, st, BN, ctas);
            return launch_mode<HV_EPI_QKNORM | MODE_DUAL>(ta, tb, tr, tr2, use_reduce, M, N, K, d, st, BN, ctas);
        }
        HvEpilogue pre = *epi;
        pre.mode = HV_EPI_STORE; pre.out = epi->out2; pre.ldo = epi->ldo2; pre.out2 = nullptr;
        if ((rc = hv_linear_bf16(A, lda, W, ldw, M, N, K, &pre, stream)) != HV_OK) return rc;
        d.e.out2 = nullptr;
    }
    switch (d.e.mode) {
        case HV_EPI_STORE: return launch_mode<HV_EPI_STORE>(ta, tb, tr, tr, use_reduce, M, N, K, d, st, BN, ctas);
        case HV_EPI_GELU: return launch_mode<HV_EPI_GELU>(ta, tb, tr, tr, use_reduce, M, N, K, d, st, BN, ctas);
        case HV_EPI_RESIDUAL: return launch_mode<HV_EPI_RESIDUAL>(ta, tb, tr, tr, use_reduce, M, N, K, d, st, BN, ctas);
        default: return launch_mode<HV_EPI_QKNORM>(ta, tb, tr, tr, use_reduce, M, N, K, d, st, BN, ctas);
    }
}

// out[M x N] (+)= At^T Bt with At (K x M, row stride ldat) and Bt (K x N, row stride ldbt) both row-major bf16: the weight
// gradient dW = dY^T X of nn.Linear (autograd of the Linears at model/modules.py:699-719) without transposed operand copies.
// The contraction dimension K (tokens) may be anything (TMA zero-fills past the end); M, N, ldat, ldbt multiples of 8,
// 16-byte aligned bases.  Epilogues: HV_EPI_STORE (fp32 / bf16) or HV_EPI_RESIDUAL without bias / gate (accumulate into an
// fp32 M x N matrix, optionally split over K with `k_split` work items per tile, reduce-added in arrival order).
extern "C" int hv_linear_bf16_tn(const void* At, int64_t ldat, const void* Bt, int64_t ldbt, int64_t M, int N, int64_t K,
                                 const HvEpilogue* epi, void* stream) {
    HV_REQUIRE(M > 0 && N > 0 && K > 0 && ldat >= M && ldbt >= N && K < (1ll << 31) && M < (1ll << 31), "hv_linear_bf16_tn: bad sizes");
    HV_REQUIRE(M % 8 == 0 && N % 8 == 0 && ldat % 8 == 0 && ldbt % 8 == 0, "hv_linear_bf16_tn: M, N, ldat, ldbt must be multiples of 8 (16-byte TMA strides)");
    HV_REQUIRE((reinterpret_cast<uintptr_t>(At) & 15) == 0 && (reinterpret_cast<uintptr_t>(Bt) & 15) == 0, "hv_linear_bf16_tn: operands must be 16-byte aligned");
    HvEpiDev d;
    int rc = hv_epilogue_prepare(epi, M, N, d, "hv_linear_bf16_tn");
    if (rc) return rc;
    HV_REQUIRE((d.e.mode == HV_EPI_STORE || d.e.mode == HV_EPI_RESIDUAL) && !d.e.out2, "hv_linear_bf16_tn: STORE or RESIDUAL epilogue only");
    static const int want_ctas = env_int("HV_GEMM_CTAS", 2), want_reduce = env_int("HV_GEMM_REDUCE", 1);
    const int BN = N > 128 ? 256 : 128;
    const int ctas = (want_ctas == 2 && N > 128 && M > 256) ? 2 : 1;
    CUtensorMap ta, tb, tr;
    // the operands as K-row matrices: inner (contiguous) dimension = the output row / column index, boxes of 64 x 64
    if ((rc = hv_make_tmap_bf16(At, K, (int)M, ldat, 64, &ta)) != HV_OK) return rc;
    if ((rc = hv_make_tmap_bf16(Bt, K, N, ldbt, 64, &tb)) != HV_OK) return rc;
    int use_reduce = 0;
    if (d.e.mode == HV_EPI_RESIDUAL && want_reduce && d.vec_ok && d.e.L >= 32) {
        if ((rc = make_tmap_residual(d.e.res, N, d.e.L, M / d.e.L, d.e.frame_rows, d.e.row_off, &tr)) != HV_OK) return rc;
        use_reduce = 1;
    } else {
        tr = ta;
    }
    if (d.e.k_split > 1) {
        HV_REQUIRE(d.e.mode == HV_EPI_RESIDUAL && !d.e.bias && !d.e.gate, "hv_linear_bf16_tn: k_split needs a plain residual epilogue (no bias, no gate)");
        const int num_kb = (int)((K + 63) / 64);
        int S = (use_reduce && d.e.k_split <= num_kb) ? d.e.k_split : (use_reduce ? num_kb : 1);
        const int per = (num_kb + S - 1) / S;
        d.e.k_split = (num_kb + per - 1) / per;
    } else {
        d.e.k_split = 1;
    }
    cudaStream_t st = hv_stream(stream);
    if (d.e.mode == HV_EPI_STORE) return launch_mode<HV_EPI_STORE>(ta, tb, tr, tr, use_reduce, M, N, (int)K, d, st, BN, ctas, 1);
    return launch_mode<HV_EPI_RESIDUAL>(ta, tb, tr, tr, use_reduce, M, N, (int)K, d, st, BN, ctas, 1);
}
