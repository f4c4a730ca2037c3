// tcgen05.mma issue / execution rate probe for the attention shapes (sm_100a): M = 128, cta_group::1, K = 16 per instruction.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I../../hivae_b200/csrc -o mma_probe mma_probe.cu && ./mma_probe
// Warp 8 issues `iters` MMAs (operands: zeroed shared memory, SWIZZLE_128B K-major; or A from TMEM), commits, waits.
//   mode "lane0":     issued under `if (lane == 0)` with per-thread operand registers (what ptxas turns into an
//                     ELECT / R2UR.BROADCAST loop per instruction)
//   mode "converged": all 32 lanes run the loop with warp-uniform operands, one elected lane issues (tc_ptx.cuh *_w)
// Warps 0..7 optionally load TMEM with tcgen05.ld (+ 16 ex2 per 16 columns, + tcgen05.st) like softmax warps do.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "tc_ptx.cuh"

#define LD16(r, addr) asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];" \
    : "=r"(r[0]),"=r"(r[1]),"=r"(r[2]),"=r"(r[3]),"=r"(r[4]),"=r"(r[5]),"=r"(r[6]),"=r"(r[7]),"=r"(r[8]),"=r"(r[9]),"=r"(r[10]),"=r"(r[11]),"=r"(r[12]),"=r"(r[13]),"=r"(r[14]),"=r"(r[15]) : "r"(addr) : "memory")

struct Cfg { int N, iters, with_ld, ts, bmn, converged; };

__global__ void __launch_bounds__(288, 1) probe(Cfg c, unsigned long long* out, float* sink) {
    extern __shared__ uint8_t raw[];
    __shared__ uint32_t slot;
    __shared__ __align__(8) uint64_t bar;
    __shared__ volatile int done;
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* As = smem;            // 128 x 64 bf16
    uint8_t* Bs = smem + 16384;    // 256 x 64 bf16 (only the addresses matter)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < (16384 + 32768) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
    if (threadIdx.x == 0) { tc::mbar_init(&bar, 1); tc::fence_barrier_init(); done = 0; }
    if (warp == 8) tc::tmem_alloc(&slot, 512);
    tc::fence_proxy_async();
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tb = slot;
    float facc = 0.f;
    if (warp == 8) {
        const uint32_t idesc = c.bmn ? tc::make_idesc_bf16_bmn(128, c.N) : tc::make_idesc_bf16(128, c.N);
        const uint64_t da = tc::make_smem_desc_k128(tc::smem_u32(As)), db = tc::make_smem_desc_k128(tc::smem_u32(Bs));
        const uint64_t bstep = c.bmn ? 128 : 2;
        long long t0 = 0, t1 = 0, t2 = 0;
        if (c.converged) {
            const uint32_t tbu = tc::warp_uniform(tb);
            t0 = clock64();
            for (int i = 0; i < c.iters; i += 4) {
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    if (c.ts) tc::umma_bf16_ts_w(tbu, tbu + 448 + (uint32_t)(8 * u), db + bstep * (uint64_t)u, idesc, 1);
                    else tc::umma_bf16_w(tbu, da + (uint64_t)(2 * u), db + bstep * (uint64_t)u, idesc, 1);
                }
            }
            t1 = clock64();
            tc::umma_commit_w(&bar);
            tc::mbar_wait(&bar, 0);
            t2 = clock64();
        } else if (lane == 0) {
            t0 = clock64();
            for (int i = 0; i < c.iters; ++i) tc::umma_bf16(tb, da + (uint64_t)(2 * (i & 3)), db + bstep * (uint64_t)(i & 3), idesc, 1);
            t1 = clock64();
            tc::umma_commit(&bar);
            tc::mbar_wait(&bar, 0);
            t2 = clock64();
        }
        if (lane == 0) {
            out[blockIdx.x * 2] = (unsigned long long)(t1 - t0);
            out[blockIdx.x * 2 + 1] = (unsigned long long)(t2 - t0);
            done = 1;
        }
    } else if (c.with_ld) {
        const uint32_t base = tb + ((uint32_t)((warp & 3) * 32) << 16) + 256u + (uint32_t)(warp >> 2) * 64u;
        uint32_t a[16];
        int n = 0;
        while (!done) {
            LD16(a, base + (n & 1) * 16);
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            facc += __uint_as_float(a[0]);
            if (c.with_ld >= 2) {  // + the MUFU work of the softmax loop
#pragma unroll
                for (int j = 0; j < 16; ++j) facc += tc::ex2_approx(__uint_as_float(a[j]) * 0.1f - 3.f);
            }
            if (c.with_ld == 3) {  // + the in-place P store
                uint32_t pk[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) pk[j] = a[2 * j] ^ a[2 * j + 1];
                tc::tmem_st8(base + 32 + (n & 1) * 8, pk);
            }
            ++n;
        }
    }
    sink[blockIdx.x * blockDim.x + threadIdx.x] = facc;
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 8) { tc::fence_after_sync(); tc::tmem_dealloc(tb, 512); }
}

int main() {
    unsigned long long* out; float* sink;
    cudaMalloc(&out, 148 * 16); cudaMalloc(&sink, 148 * 288 * 4);
    const int smem = 16384 + 32768 + 1024;
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    const int iters = 2048;
    const char* lds[] = {"idle", "tcgen05.ld", "ld + 16 ex2", "ld + ex2 + st"};
    printf("M = 128, K = 16 per MMA, 148 CTAs; floor = N / 2 clk per MMA\n");
    for (int conv = 0; conv < 2; ++conv)
        for (int ts = 0; ts < 2; ++ts) {
            if (!conv && ts) continue;
            for (int with_ld = 0; with_ld < 4; ++with_ld) {
                if (!conv && with_ld > 1) continue;
                for (int N : {64, 128, 144, 256}) {
                    Cfg c{N, iters, with_ld, ts, N == 64 ? 1 : 0, conv};
                    probe<<<148, 288, smem>>>(c, out, sink);
                    probe<<<148, 288, smem>>>(c, out, sink);
                    unsigned long long h[296];
                    if (cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost) != cudaSuccess) { printf("error %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
                    double ti = 0, te = 0;
                    for (int i = 0; i < 148; ++i) { ti += (double)h[2 * i]; te += (double)h[2 * i + 1]; }
                    printf("%-9s %-9s other warps: %-14s N %3d: issue %6.1f clk/MMA, executed %6.1f clk/MMA (floor %5.1f) = %3.0f%% of peak\n",
                           conv ? "converged" : "lane0", ts ? "A in TMEM" : "A in smem", lds[with_ld], N, ti / 148 / iters, te / 148 / iters, N / 2.0,
                           100.0 * (N / 2.0) / (te / 148 / iters));
                }
            }
        }
    return 0;
}
