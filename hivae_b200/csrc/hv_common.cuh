// Shared device/host helpers for the hivae_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include "../../include/hivae_b200.h"

// ---------------------------------------------------------------- host-side error plumbing
void hv_set_error(const char* fmt, ...);
void hv_count_launch();

#define HV_REQUIRE(cond, ...)                  \
    do {                                       \
        if (!(cond)) {                         \
            hv_set_error(__VA_ARGS__);         \
            return HV_EINVAL;                  \
        }                                      \
    } while (0)

#define HV_CUDA(call)                                                                   \
    do {                                                                                \
        cudaError_t e__ = (call);                                                       \
        if (e__ != cudaSuccess) {                                                       \
            hv_set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
            return HV_ECUDA;                                                            \
        }                                                                               \
    } while (0)

#define HV_LAUNCH_CHECK()                                                               \
    do {                                                                                \
        hv_count_launch();                                                              \
        cudaError_t e__ = cudaGetLastError();                                           \
        if (e__ != cudaSuccess) {                                                       \
            hv_set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(e__), __FILE__, __LINE__); \
            return HV_ECUDA;                                                            \
        }                                                                               \
    } while (0)

static inline cudaStream_t hv_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }
int hv_num_sms();
bool hv_pdl_enabled();  // HV_PDL=1 switches programmatic dependent launch on (measured slower under graph replay, see capi.cu)
int hv_device_slot();  // current device ordinal clamped to [0, HV_MAX_DEVICES): index of the per-device host caches
constexpr int HV_MAX_DEVICES = 64;

// launch with programmatic dependent launch allowed (tc_ptx.cuh grid_dep_wait): the kernel MUST execute griddepcontrol.wait
// before its first global-memory access
template <typename... KArgs, typename... Args>
inline cudaError_t hv_launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cfg.attrs = at; cfg.numAttrs = hv_pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

// ---------------------------------------------------------------- device helpers
__device__ __forceinline__ float hv_warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float hv_warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

__device__ __forceinline__ float hv_gelu_tanh(float x) {
    // F.gelu(approximate="tanh"): 0.5 x (1 + tanh(sqrt(2/pi) (x + 0.044715 x^3)))
    const float k0 = 0.7978845608028654f, k1 = 0.044715f;
    float u = k0 * (x + k1 * x * x * x);
    return 0.5f * x * (1.0f + tanhf(u));
}
__device__ __forceinline__ float hv_silu_f(float x) { return x / (1.0f + __expf(-x)); }

__device__ __forceinline__ uint32_t hv_pack_bf16(float a, float b) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
}

// typed scalar load/store through a dtype code (HV_F32 / HV_BF16)
template <typename T> __device__ __forceinline__ float hv_ld(const T* p);
template <> __device__ __forceinline__ float hv_ld<float>(const float* p) { return *p; }
template <> __device__ __forceinline__ float hv_ld<__nv_bfloat16>(const __nv_bfloat16* p) { return __bfloat162float(*p); }
template <typename T> __device__ __forceinline__ void hv_st(T* p, float v);
template <> __device__ __forceinline__ void hv_st<float>(float* p, float v) { *p = v; }
template <> __device__ __forceinline__ void hv_st<__nv_bfloat16>(__nv_bfloat16* p, float v) { *p = __float2bfloat16(v); }

// residual-row addressing shared by the row kernels and the GEMM epilogues
struct HvRowMap {
    int L;
    long long frame_rows;
    int row_off;
    __device__ __forceinline__ long long res_row(long long r, int& frame, int& abs_i) const {
        frame = (int)(r / L);
        int i = (int)(r - (long long)frame * L);
        abs_i = row_off + i;
        return (long long)frame * frame_rows + abs_i;
    }
};
