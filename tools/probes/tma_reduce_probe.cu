// Probe: which cp.reduce.async.bulk.tensor configurations run on this GPU (fp32 add; swizzle; negative / OOB coords).
// nvcc -gencode arch=compute_100a,code=sm_100a -o tma_reduce_probe tma_reduce_probe.cu -lcuda
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>
#include <cmath>

__global__ void probe(const __grid_constant__ CUtensorMap tm, int c0, int c1, int c2, int rank, int swz, int srow) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    float* st = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int lane = threadIdx.x;
    for (int k = 0; k < 8; ++k) {  // row = lane, chunk k (4 floats): value = 1000*row + col
        int kk = swz ? (k ^ (lane & 7)) : k;
        for (int j = 0; j < 4; ++j) st[lane * 32 + kk * 4 + j] = 1000.f * lane + (k * 4 + j);
    }
    for (int i = lane; i < 32 * 32; i += 32) st[32 * 32 + i] = 0.f;  // rows 32..63 = 0
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncwarp();
    if (lane == 0) {
        unsigned s = (unsigned)__cvta_generic_to_shared(st + srow * 32);
        if (rank == 3)
            asm volatile("cp.reduce.async.bulk.tensor.3d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3, %4}], [%1];"
                         ::"l"(reinterpret_cast<unsigned long long>(&tm)), "r"(s), "r"(c0), "r"(c1), "r"(c2) : "memory");
        else
            asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];"
                         ::"l"(reinterpret_cast<unsigned long long>(&tm)), "r"(s), "r"(c0), "r"(c1) : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
    void* p = nullptr; cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    EncodeFn enc = (EncodeFn)p;
    const int N = 128, L = 40, FR = 50, frames = 6, off = 3;
    float* d; cudaMalloc(&d, sizeof(float) * frames * FR * N);
    struct Case { int rank, swz, c0, c1, c2; const char* name; int srow; } cases[] = {
        {2, 0, 0, 0, 0, "2d noswz in-bounds"}, {2, 1, 0, 0, 0, "2d swz128 in-bounds"},
        {3, 0, 32, 0, 1, "3d noswz in-bounds"}, {3, 1, 32, 0, 1, "3d swz128 in-bounds"},
        {3, 1, 32, 20, 1, "3d swz128 rows past L clipped"},
        {3, 1, 96, 8, 5, "3d swz128 last frame"}, {3, 1, 0, 8, 6, "3d swz128 frame == frames (all OOB)"},
        {3, 1, 32, 0, 2, "3d swz128 source offset 8 rows", 8}, {3, 1, 32, 0, 2, "3d swz128 source offset 5 rows", 5},
        {3, 1, 64, 0, 3, "3d swz128 source offset 31 rows", 31}, {3, 0, 64, 0, 3, "3d noswz source offset 13 rows", 13},
        {3, 1, 32, -12, 2, "3d swz128 negative row start"}};
    for (auto& c : cases) {
        cudaMemset(d, 0, sizeof(float) * frames * FR * N);
        CUtensorMap tm;
        cuuint64_t gdim[3] = {(cuuint64_t)N, (cuuint64_t)L, (cuuint64_t)frames};
        cuuint64_t gstr[2] = {(cuuint64_t)N * 4, (cuuint64_t)FR * N * 4};
        cuuint32_t box[3] = {32, 32, 1}, es[3] = {1, 1, 1};
        if (c.rank == 2) { gdim[1] = (cuuint64_t)frames * FR - off; }
        CUresult rc = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, c.rank, d + off * N, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          c.swz ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (rc != CUDA_SUCCESS) { printf("%-40s encode failed %d\n", c.name, (int)rc); continue; }
        probe<<<1, 32, 16384>>>(tm, c.c0, c.c1, c.c2, c.rank, c.swz, c.srow);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("%-40s RUNTIME ERROR: %s\n", c.name, cudaGetErrorString(e)); return 1; }
        std::vector<float> h(frames * FR * N);
        cudaMemcpy(h.data(), d, h.size() * 4, cudaMemcpyDeviceToHost);
        long bad = 0, hit = 0;
        for (int f = 0; f < frames; ++f) for (int r = 0; r < FR; ++r) for (int col = 0; col < N; ++col) {
            float want = 0.f;
            if (c.rank == 3) {
                int i = r - off - c.c1, j = col - c.c0;
                if (f == c.c2 && r - off >= 0 && r - off < L && i >= 0 && i < 32 && j >= 0 && j < 32 && i + c.srow < 32) want = 1000.f * (i + c.srow) + j;
            } else {
                long row = (long)f * FR + r - off - c.c1; int j = col - c.c0;
                if (row >= 0 && row < 32 && j >= 0 && j < 32) want = 1000.f * row + j;
            }
            float got = h[((size_t)f * FR + r) * N + col];
            if (want != 0.f) ++hit;
            if (got != want) ++bad;
        }
        printf("%-40s ok=%d (elements written %ld, mismatches %ld)\n", c.name, bad == 0, hit, bad);
    }
    return 0;
}
