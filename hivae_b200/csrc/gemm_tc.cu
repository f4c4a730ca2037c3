// bf16 tensor-core GEMM for sm_100a:  out = epilogue(A[M,K] @ W[N,K]^T), fp32 accumulation.
//
//   * operands staged by TMA (cp.async.bulk.tensor, SWIZZLE_128B) into a 4-stage smem ring,
//   * tcgen05.mma (cta_group::1, 128 x BN x 16) issued by one thread, accumulators in TMEM,
//   * two TMEM accumulator stages so the epilogue of tile i overlaps the main loop of tile i+1,
//   * persistent CTAs (one per SM) walking tiles N-fastest so the weight matrix stays in L2 and
//     every activation tile is streamed from HBM once,
//   * epilogue warps read TMEM with tcgen05.ld and apply the fused epilogue (bias / GELU-tanh /
//     gated residual scatter / per-head QK LayerNorm) straight from registers.
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = MMA issuer + TMEM owner,
// warps 2..5 = epilogue (TMEM lane quarter = warp_idx % 4).
#include <cuda.h>
#include <mutex>
#include <unordered_map>
#include "epilogue.cuh"
#include "tc_ptx.cuh"

namespace {

constexpr int BM = 128, BK = 64, STAGES = 4, THREADS = 192;

template <int BN> constexpr int stage_bytes() { return (BM + BN) * BK * 2; }
template <int BN> constexpr int smem_bytes() { return STAGES * stage_bytes<BN>() + 1024; }

template <int BN, bool EPI64>
__global__ void __launch_bounds__(THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
               long long M, int N, int K, const __grid_constant__ HvEpiDev epi) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[STAGES], empty_bar[STAGES], tmem_full_bar[2], tmem_empty_bar[2];
    __shared__ uint32_t tmem_slot;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    auto smem_a = [&](int s) { return smem + s * stage_bytes<BN>(); };
    auto smem_b = [&](int s) { return smem + s * stage_bytes<BN>() + BM * BK * 2; };

    const int n_blocks = (N + BN - 1) / BN;
    const long long m_blocks = (M + BM - 1) / BM;
    const long long num_tiles = m_blocks * n_blocks;
    const int num_kb = (K + BK - 1) / BK;
    constexpr uint32_t TMEM_COLS = 2 * BN;  // 256 or 512: power of two

    if (threadIdx.x == 0) {
        tc::prefetch_tmap(&tmap_a);
        tc::prefetch_tmap(&tmap_b);
#pragma unroll
        for (int s = 0; s < STAGES; ++s) { tc::mbar_init(&full_bar[s], 1); tc::mbar_init(&empty_bar[s], 1); }
#pragma unroll
        for (int a = 0; a < 2; ++a) { tc::mbar_init(&tmem_full_bar[a], 1); tc::mbar_init(&tmem_empty_bar[a], 4); }
        tc::fence_barrier_init();
    }
    if (warp == 1) tc::tmem_alloc(&tmem_slot, TMEM_COLS);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem_base = tmem_slot;

    if (warp == 0) {
        // ===================================================== TMA producer
        if (lane == 0) {
            int s = 0; uint32_t ph = 0;
            for (long long t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                const int m0 = (int)(t / n_blocks) * BM, n0 = (int)(t % n_blocks) * BN;
                for (int kb = 0; kb < num_kb; ++kb) {
                    tc::mbar_wait(&empty_bar[s], ph ^ 1);
                    tc::mbar_expect_tx(&full_bar[s], stage_bytes<BN>());
                    tc::tma_load_2d(smem_a(s), &tmap_a, &full_bar[s], kb * BK, m0);
                    tc::tma_load_2d(smem_b(s), &tmap_b, &full_bar[s], kb * BK, n0);
                    if (++s == STAGES) { s = 0; ph ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================================================== MMA issuer
        if (lane == 0) {
            constexpr uint32_t idesc = tc::make_idesc_bf16(BM, BN);
            int s = 0; uint32_t ph = 0; int a = 0; uint32_t aph = 0;
            for (long long t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                tc::mbar_wait(&tmem_empty_bar[a], aph ^ 1);
                tc::fence_after_sync();
                const uint32_t d_tmem = tmem_base + (uint32_t)(a * BN);
                for (int kb = 0; kb < num_kb; ++kb) {
                    tc::mbar_wait(&full_bar[s], ph);
                    tc::fence_after_sync();
                    const uint64_t da = tc::make_smem_desc_k128(tc::smem_u32(smem_a(s)));
                    const uint64_t db = tc::make_smem_desc_k128(tc::smem_u32(smem_b(s)));
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k)  // +32 B per K=16 step inside the 128 B swizzle row
                        tc::umma_bf16(d_tmem, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (kb | k) != 0);
                    tc::umma_commit(&empty_bar[s]);  // frees the smem stage once these MMAs retire
                    if (++s == STAGES) { s = 0; ph ^= 1; }
                }
                tc::umma_commit(&tmem_full_bar[a]);  // accumulator complete
                if ((a ^= 1) == 0) aph ^= 1;
            }
        }
    } else {
        // ===================================================== epilogue warps
        const int q = warp & 3;  // TMEM lane quarter this warp may access
        int a = 0; uint32_t aph = 0;
        constexpr int NC = EPI64 ? 64 : 32;
        for (long long t = blockIdx.x; t < num_tiles; t += gridDim.x) {
            const long long m0 = (t / n_blocks) * BM;
            const int n0 = (int)(t % n_blocks) * BN;
            tc::mbar_wait(&tmem_full_bar[a], aph);
            tc::fence_after_sync();
            const long long r = m0 + q * 32 + lane;
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(a * BN);
#pragma unroll 1
            for (int c = 0; c < BN; c += NC) {
                if (n0 + c >= N) break;  // warp-uniform
                float acc[NC];
                if constexpr (EPI64) {
                    float lo[32], hi[32];
                    tc::tmem_ld32(taddr + c, lo);
                    tc::tmem_ld32(taddr + c + 32, hi);
#pragma unroll
                    for (int j = 0; j < 32; ++j) { acc[j] = lo[j]; acc[32 + j] = hi[j]; }
                } else {
                    tc::tmem_ld32(taddr + c, acc);
                }
                if (r < M) hv_epilogue_row<NC>(epi, r, n0 + c, N, acc);
            }
            tc::fence_before_sync();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(&tmem_empty_bar[a]);
            if ((a ^= 1) == 0) aph ^= 1;
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 1) {
        tc::fence_after_sync();
        tc::tmem_dealloc(tmem_base, TMEM_COLS);
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

// K-major bf16 matrix [rows, K] with row stride ld -> 2-D tensor map, box = 64 (K) x box_rows, 128 B swizzle
int make_tmap(const void* ptr, long long rows, int K, long long ld, int box_rows, CUtensorMap* out) {
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

template <int BN, bool EPI64>
int launch(const CUtensorMap& ta, const CUtensorMap& tb, int64_t M, int N, int K, const HvEpiDev& d, cudaStream_t st) {
    auto kern = gemm_tc_kernel<BN, EPI64>;
    static bool configured = false;
    if (!configured) {
        HV_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes<BN>()));
        configured = true;
    }
    const long long tiles = ((M + BM - 1) / BM) * ((N + BN - 1) / BN);
    const int sms = hv_num_sms();
    const int grid = (int)(tiles < sms ? tiles : sms);
    kern<<<grid, THREADS, smem_bytes<BN>(), st>>>(ta, tb, (long long)M, N, K, d);
    HV_LAUNCH_CHECK();
    return HV_OK;
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
    const int BN = N > 128 ? 256 : 128;
    CUtensorMap ta, tb;
    if ((rc = make_tmap(A, M, K, lda, BM, &ta)) != HV_OK) return rc;
    if ((rc = make_tmap(W, N, K, ldw, BN, &tb)) != HV_OK) return rc;
    cudaStream_t st = hv_stream(stream);
    const bool e64 = d.e.mode == HV_EPI_QKNORM;
    if (BN == 256) return e64 ? launch<256, true>(ta, tb, M, N, K, d, st) : launch<256, false>(ta, tb, M, N, K, d, st);
    return e64 ? launch<128, true>(ta, tb, M, N, K, d, st) : launch<128, false>(ta, tb, M, N, K, d, st);
}
