// TMEM read / MUFU throughput probe for the attention softmax loop (sm_100a).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tmem_probe tmem_probe.cu && ./tmem_probe
// One CTA per SM, `warps` warps each reading its own lane quarter; cycles per iteration from clock64 around the loop.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define LD16(r, addr) asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];" \
    : "=r"(r[0]),"=r"(r[1]),"=r"(r[2]),"=r"(r[3]),"=r"(r[4]),"=r"(r[5]),"=r"(r[6]),"=r"(r[7]),"=r"(r[8]),"=r"(r[9]),"=r"(r[10]),"=r"(r[11]),"=r"(r[12]),"=r"(r[13]),"=r"(r[14]),"=r"(r[15]) : "r"(addr) : "memory")
#define WAIT16(r) asm volatile("tcgen05.wait::ld.sync.aligned;" : "+r"(r[0]),"+r"(r[1]),"+r"(r[2]),"+r"(r[3]),"+r"(r[4]),"+r"(r[5]),"+r"(r[6]),"+r"(r[7]),"+r"(r[8]),"+r"(r[9]),"+r"(r[10]),"+r"(r[11]),"+r"(r[12]),"+r"(r[13]),"+r"(r[14]),"+r"(r[15]) :: "memory")
#define WAITALL() asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory")
#define ST8(addr, r) asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" :: "r"(addr),"r"(r[0]),"r"(r[1]),"r"(r[2]),"r"(r[3]),"r"(r[4]),"r"(r[5]),"r"(r[6]),"r"(r[7]) : "memory")

__device__ __forceinline__ float ex2(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

__global__ void __launch_bounds__(512, 1) probe(int mode, int iters, unsigned long long* out, float* sink) {
    __shared__ uint32_t slot;
    __shared__ unsigned long long t0s;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&slot)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t base = slot + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)((warp >> 2) & 3) * 128u;
    uint32_t a[16], b[16];
    uint32_t acc = 0;
    float facc = 0.f, x = (float)lane * -0.01f;
    __syncthreads();
    if (threadIdx.x == 0) t0s = clock64();
    __syncthreads();
    if (mode == 0) {          // x16, one in flight
        for (int i = 0; i < iters; ++i) { LD16(a, base + (i & 7) * 16); WAIT16(a); acc ^= a[0] ^ a[15]; }
    } else if (mode == 1) {   // x16, two in flight
        LD16(a, base);
        for (int i = 0; i < iters; i += 2) {
            WAIT16(a); LD16(b, base + 16); acc ^= a[0] ^ a[15];
            WAIT16(b); LD16(a, base + 32); acc ^= b[0] ^ b[15];
        }
        WAITALL();
    } else if (mode == 2) {   // MUFU only: 16 independent ex2 per iteration
        for (int i = 0; i < iters; ++i) {
#pragma unroll
            for (int j = 0; j < 16; ++j) facc += ex2(x + (float)j);
            x += 1e-6f;
        }
    } else if (mode == 3) {   // the softmax loop without stores: double-buffered x16 + 16 (fma, ex2, add)
        LD16(a, base);
        for (int i = 0; i < iters; i += 2) {
            WAIT16(a); LD16(b, base + 16);
#pragma unroll
            for (int j = 0; j < 16; ++j) facc += ex2(fmaf(__uint_as_float(a[j]), 0.18f, -x));
            WAIT16(b); LD16(a, base + 32);
#pragma unroll
            for (int j = 0; j < 16; ++j) facc += ex2(fmaf(__uint_as_float(b[j]), 0.18f, -x));
        }
        WAITALL();
    } else if (mode == 4) {   // mode 3 + bf16 pack + tcgen05.st x8 (P in place)
        LD16(a, base);
        for (int i = 0; i < iters; i += 2) {
            uint32_t pk[8];
            WAIT16(a); LD16(b, base + 16);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float v0 = ex2(fmaf(__uint_as_float(a[2 * j]), 0.18f, -x)), v1 = ex2(fmaf(__uint_as_float(a[2 * j + 1]), 0.18f, -x));
                facc += v0 + v1;
                asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(pk[j]) : "f"(v1), "f"(v0));
            }
            ST8(base + 64, pk);
            WAIT16(b); LD16(a, base + 32);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float v0 = ex2(fmaf(__uint_as_float(b[2 * j]), 0.18f, -x)), v1 = ex2(fmaf(__uint_as_float(b[2 * j + 1]), 0.18f, -x));
                facc += v0 + v1;
                asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(pk[j]) : "f"(v1), "f"(v0));
            }
            ST8(base + 72, pk);
        }
        WAITALL();
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) out[blockIdx.x] = clock64() - t0s;
    sink[blockIdx.x * blockDim.x + threadIdx.x] = facc + (float)acc;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(slot), "r"(512u) : "memory");
}

int main() {
    unsigned long long* out; float* sink;
    cudaMalloc(&out, 148 * 8); cudaMalloc(&sink, 148 * 512 * 4);
    const char* names[] = {"x16, 1 in flight", "x16, 2 in flight", "MUFU only (16 ex2)", "ld x16 dbuf + 16 ex2", "ld + ex2 + pack + st x8"};
    const int iters = 4096;
    for (int mode = 0; mode < 5; ++mode)
        for (int warps = 4; warps <= 16; warps *= 2) {
            probe<<<148, warps * 32>>>(mode, iters, out, sink);
            probe<<<148, warps * 32>>>(mode, iters, out, sink);
            unsigned long long h[148];
            if (cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost) != cudaSuccess) { printf("error %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
            double avg = 0; for (int i = 0; i < 148; ++i) avg += (double)h[i]; avg /= 148;
            const double per_iter = avg / iters;                       // cycles per 16-column chunk per warp (all warps concurrently)
            const double bytes = (mode == 2) ? 0 : 2048.0 * warps;     // TMEM bytes read per iteration over the SM
            printf("%-26s warps %2d: %7.1f clk / 16-col chunk / warp   TMEM %6.1f B/clk/SM   ex2 %5.1f /clk/SM\n", names[mode], warps, per_iter,
                   bytes / per_iter, (mode >= 2 ? 16.0 * 32 * warps / per_iter : 0.0));
        }
    return 0;
}
