// Train-step tail of train_amd.py:489-496 on flat fp32 arenas (one buffer each for parameters, gradients and
// the two AdamW moments): mean-squared-error loss (model/loss.py:13-14), global gradient norm
// (accelerator.clip_grad_norm_ -> torch.nn.utils.clip_grad_norm_) and the fused clip + AdamW update
// (torch.optim.AdamW defaults, train_amd.py:255).  All three are HBM-bound streaming kernels: 128-bit
// accesses, grid = SMs x resident CTAs, no host synchronisation anywhere (the clip coefficient and the
// loss stay on the device).
//
// Algorithmic bytes: mse 8 B/element; grad norm 4 B/parameter; clip+AdamW 28 B/parameter
// (read g, p, m, v = 16 B, write p, m, v = 12 B) — the clip scaling costs no extra pass.
#include "hv_common.cuh"

namespace {

constexpr int TPB = 256;

__device__ __forceinline__ double block_sum(double v, double* sh) {  // blockDim.x == TPB
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) sh[w] = v;
    __syncthreads();
    v = (threadIdx.x < TPB / 32) ? sh[threadIdx.x] : 0.0;
    if (w == 0) {
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    }
    return v;  // valid in thread 0
}

// partial[b] = sum over the CTA's grid-stride range of f(x, y); fp32 per-thread accumulation over short
// runs, fp64 across threads / CTAs (deterministic: fixed grid, fixed order)
template <int OP>  // 0: (x - y)^2, 1: x^2
__global__ void __launch_bounds__(TPB) sumsq_partial_kernel(const float* __restrict__ x, const float* __restrict__ y,
                                                            long long n, double* __restrict__ partial) {
    __shared__ double sh[TPB / 32];
    const long long n4 = n >> 2;
    const float4* x4 = reinterpret_cast<const float4*>(x);
    const float4* y4 = reinterpret_cast<const float4*>(y);
    double acc = 0.0;
    for (long long i = blockIdx.x * (long long)TPB + threadIdx.x; i < n4; i += (long long)gridDim.x * TPB) {
        float4 a = __ldcs(x4 + i);
        if (OP == 0) { const float4 b = __ldcs(y4 + i); a.x -= b.x; a.y -= b.y; a.z -= b.z; a.w -= b.w; }
        acc += (double)((a.x * a.x + a.y * a.y) + (a.z * a.z + a.w * a.w));
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {  // ragged tail
        const long long i = (n4 << 2) + threadIdx.x;
        const float d = OP == 0 ? x[i] - y[i] : x[i];
        acc += (double)(d * d);
    }
    const double s = block_sum(acc, sh);
    if (threadIdx.x == 0) partial[blockIdx.x] = s;
}

// out[0] = OP 0: sum / n (mean squared error)   OP 1: sqrt(sum) (l2 norm); out[1] (OP 1) = clip coefficient
template <int OP>
__global__ void __launch_bounds__(TPB) sumsq_final_kernel(const double* __restrict__ partial, int parts, double denom,
                                                          float max_norm, float* __restrict__ out) {
    __shared__ double sh[TPB / 32];
    double acc = 0.0;
    for (int i = threadIdx.x; i < parts; i += TPB) acc += partial[i];
    const double s = block_sum(acc, sh);
    if (threadIdx.x == 0) {
        if (OP == 0) {
            out[0] = (float)(s / denom);
        } else {
            const float norm = (float)(sqrt(s) * denom);  // denom = |grad_scale| here: ||s g|| = |s| ||g||
            out[0] = norm;
            // torch.nn.utils.clip_grad_norm_: clip_coef = max_norm / (total_norm + 1e-6), clamped to 1
            const float coef = max_norm / (norm + 1e-6f);
            out[1] = coef < 1.0f ? coef : 1.0f;
        }
    }
}

struct AdamArgs {  // every scalar is formed in double on the host (like torch's Python scalars) and rounded once
    float decay;          // 1 - lr * weight_decay
    float omb1, beta2, omb2, eps;  // 1 - beta1, beta2, 1 - beta2, eps
    float step_size;      // lr / (1 - beta1^t)
    float bc2_sqrt;       // sqrt(1 - beta2^t)
    float grad_scale;     // folded into the gradient before the clip coefficient (1/world for summed all-reduces)
};

// operation order of torch.optim.adamw._single_tensor_adamw (fp32):
//   p *= 1 - lr*wd;  m = lerp(m, g, 1-b1);  v = v*b2 + (1-b2)*g*g;  p += -step_size * m / (sqrt(v)/bc2_sqrt + eps)
__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, const AdamArgs& a, float coef) {
    g = g * a.grad_scale * coef;
    p = p * a.decay;
    m = m + a.omb1 * (g - m);
    v = v * a.beta2 + a.omb2 * g * g;
    const float denom = sqrtf(v) / a.bc2_sqrt + a.eps;
    p = p + (-a.step_size) * (m / denom);
}

// device-resident optimizer state for replayed CUDA graphs (hv_adamw_clip_devstate): state[0] = step count, state[1] = lr;
// one thread advances the step and forms the step-dependent scalars in double exactly as the host path does
__global__ void adam_state_kernel(double* __restrict__ state, float* __restrict__ derived, double beta1, double beta2, double weight_decay) {
    const double step = state[0] + 1.0, lr = state[1];
    state[0] = step;
    derived[0] = (float)(1.0 - lr * weight_decay);          // decay
    derived[1] = (float)(lr / (1.0 - pow(beta1, step)));    // step_size
    derived[2] = (float)sqrt(1.0 - pow(beta2, step));       // bc2_sqrt
}

__global__ void __launch_bounds__(TPB) adamw_clip_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                                         float* __restrict__ v, long long n, AdamArgs a,
                                                         const float* __restrict__ clip_coef, const float* __restrict__ derived = nullptr) {
    if (derived) { a.decay = __ldg(derived); a.step_size = __ldg(derived + 1); a.bc2_sqrt = __ldg(derived + 2); }
    const float coef = clip_coef ? __ldg(clip_coef) : 1.0f;
    const long long n4 = n >> 2;
    float4* p4 = reinterpret_cast<float4*>(p);
    const float4* g4 = reinterpret_cast<const float4*>(g);
    float4* m4 = reinterpret_cast<float4*>(m);
    float4* v4 = reinterpret_cast<float4*>(v);
    for (long long i = blockIdx.x * (long long)TPB + threadIdx.x; i < n4; i += (long long)gridDim.x * TPB) {
        float4 pp = __ldcs(p4 + i), mm = __ldcs(m4 + i), vv = __ldcs(v4 + i);
        const float4 gg = __ldcs(g4 + i);
        adam_one(pp.x, gg.x, mm.x, vv.x, a, coef); adam_one(pp.y, gg.y, mm.y, vv.y, a, coef);
        adam_one(pp.z, gg.z, mm.z, vv.z, a, coef); adam_one(pp.w, gg.w, mm.w, vv.w, a, coef);
        __stcs(p4 + i, pp); __stcs(m4 + i, mm); __stcs(v4 + i, vv);
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
        const long long i = (n4 << 2) + threadIdx.x;
        adam_one(p[i], g[i], m[i], v[i], a, coef);
    }
}

__global__ void __launch_bounds__(TPB) scale_by_device_kernel(float* __restrict__ x, long long n, const float* __restrict__ coef, float extra) {
    const float c = __ldg(coef) * extra;
    for (long long i = blockIdx.x * (long long)TPB + threadIdx.x; i < n; i += (long long)gridDim.x * TPB) x[i] *= c;
}

int stream_grid(long long n4) {
    long long need = (n4 + TPB - 1) / TPB;
    const long long cap = (long long)hv_num_sms() * 8;
    if (need < 1) need = 1;
    return (int)(need < cap ? need : cap);
}

}  // namespace

// number of fp64 partial sums the reduction entry points need in `workspace`
extern "C" int hv_reduce_workspace_doubles() { return hv_num_sms() * 8; }

// out[0] = mean((pre - gt)^2)   (model/loss.py:13-14, F.mse_loss reduction='mean'); no host sync
extern "C" int hv_mse_loss(const float* pre, const float* gt, int64_t n, double* workspace, float* out, void* stream) {
    HV_REQUIRE(n > 0 && pre && gt && workspace && out, "hv_mse_loss: bad arguments");
    HV_REQUIRE(((reinterpret_cast<uintptr_t>(pre) | reinterpret_cast<uintptr_t>(gt)) & 15) == 0, "hv_mse_loss: inputs must be 16-byte aligned");
    const int grid = stream_grid(n >> 2);
    cudaStream_t st = hv_stream(stream);
    sumsq_partial_kernel<0><<<grid, TPB, 0, st>>>(pre, gt, n, workspace);
    HV_LAUNCH_CHECK();
    sumsq_final_kernel<0><<<1, TPB, 0, st>>>(workspace, grid, (double)n, 0.f, out);
    HV_LAUNCH_CHECK();
    return HV_OK;
}

// ---- hv_mse_loss_static: hv_mse_loss for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that
// forbid cudaMalloc).  No counterpart in the reference.  The fp64 partial sums
// live in a file-scope device array instead of a caller-provided workspace: HV_MSE_STATIC_MAX_PARTIALS doubles = 8 KB, held by
// every GPU that loads the library until the process exits.  In host code `g_mse_static_scratch` is only a host-side
// stand-in for the symbol, NOT a device address: the device address is looked up once per device, with that device
// current, through cudaGetSymbolAddress and cached; that pointer is what the kernels get.
// Every call shares the array: calls that overlap on the GPU (different streams, or concurrent graph launches) race.
// Nothing has to be cleared between calls: the first kernel overwrites partials [0, grid) and the second reads exactly
// those, so stale values of an earlier (larger) call are never read.
constexpr int HV_MSE_STATIC_MAX_PARTIALS = 1024;
static __device__ __align__(256) double g_mse_static_scratch[HV_MSE_STATIC_MAX_PARTIALS];

extern "C" int hv_mse_loss_static_max_elements() { return HV_MSE_STATIC_MAX_PARTIALS * TPB * 4; }

extern "C" int hv_mse_loss_static(const float* pre, const float* gt, int64_t n, float* out, void* stream) {
    HV_REQUIRE(n > 0 && pre && gt && out, "hv_mse_loss_static: bad arguments");
    HV_REQUIRE(((reinterpret_cast<uintptr_t>(pre) | reinterpret_cast<uintptr_t>(gt)) & 15) == 0, "hv_mse_loss_static: inputs must be 16-byte aligned");
    const int grid = stream_grid(n >> 2);  // same launch geometry as hv_mse_loss: bit-identical result
    if (grid > HV_MSE_STATIC_MAX_PARTIALS) {
        hv_set_error("hv_mse_loss_static: %lld elements need %d partial sums, the static scratch holds %d (at most %d elements; "
                     "use hv_mse_loss with a caller-provided workspace for larger inputs)",
                     (long long)n, grid, HV_MSE_STATIC_MAX_PARTIALS, HV_MSE_STATIC_MAX_PARTIALS * TPB * 4);
        return HV_EINVAL;
    }
    constexpr int MAX_DEVICES = 64;
    static double* scratch_of[MAX_DEVICES] = {};  // device address of the symbol, per device
    int dev = -1;
    HV_CUDA(cudaGetDevice(&dev));
    HV_REQUIRE(dev >= 0 && dev < MAX_DEVICES, "hv_mse_loss_static: device ordinal out of range");
    if (scratch_of[dev] == nullptr) {  // set-up, once per device, with that device current (do it before a graph capture)
        double* ptr = nullptr;
        HV_CUDA(cudaGetSymbolAddress((void**)&ptr, g_mse_static_scratch));
        HV_REQUIRE(ptr != nullptr, "hv_mse_loss_static: cudaGetSymbolAddress returned a null address");
        scratch_of[dev] = ptr;
    }
    double* ws = scratch_of[dev];
    cudaStream_t st = hv_stream(stream);
    sumsq_partial_kernel<0><<<grid, TPB, 0, st>>>(pre, gt, n, ws);
    HV_LAUNCH_CHECK();
    sumsq_final_kernel<0><<<1, TPB, 0, st>>>(ws, grid, (double)n, 0.f, out);
    HV_LAUNCH_CHECK();
    return HV_OK;
}

// out[0] = ||grad_scale * g||_2 over the flat gradient arena, out[1] = min(1, max_norm / (out[0] + 1e-6))
// (torch.nn.utils.clip_grad_norm_, called through accelerator.clip_grad_norm_ at train_amd.py:492)
extern "C" int hv_grad_norm_clip_coef(const float* g, int64_t n, float grad_scale, float max_norm, double* workspace,
                                      float* out, void* stream) {
    HV_REQUIRE(n > 0 && g && workspace && out, "hv_grad_norm_clip_coef: bad arguments");
    HV_REQUIRE((reinterpret_cast<uintptr_t>(g) & 15) == 0, "hv_grad_norm_clip_coef: g must be 16-byte aligned");
    const int grid = stream_grid(n >> 2);
    cudaStream_t st = hv_stream(stream);
    sumsq_partial_kernel<1><<<grid, TPB, 0, st>>>(g, nullptr, n, workspace);
    HV_LAUNCH_CHECK();
    sumsq_final_kernel<1><<<1, TPB, 0, st>>>(workspace, grid, (double)fabsf(grad_scale), max_norm, out);
    HV_LAUNCH_CHECK();
    return HV_OK;
}

// fused clip + AdamW step t (1-based) on flat arenas; clip_coef: device scalar from hv_grad_norm_clip_coef or NULL
extern "C" int hv_adamw_clip(float* p, const float* g, float* m, float* v, int64_t n, double lr, double beta1, double beta2,
                             double eps, double weight_decay, int64_t step, float grad_scale, const float* clip_coef, void* stream) {
    HV_REQUIRE(n > 0 && p && g && m && v && step >= 1, "hv_adamw_clip: bad arguments");
    HV_REQUIRE(((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) | reinterpret_cast<uintptr_t>(v)) & 15) == 0,
               "hv_adamw_clip: arenas must be 16-byte aligned");
    AdamArgs a;
    a.decay = (float)(1.0 - lr * weight_decay); a.omb1 = (float)(1.0 - beta1); a.beta2 = (float)beta2;
    a.omb2 = (float)(1.0 - beta2); a.eps = (float)eps; a.grad_scale = grad_scale;
    // bias corrections in double like Python's float arithmetic in torch.optim (1 - beta ** step)
    const double bc1 = 1.0 - pow(beta1, (double)step), bc2 = 1.0 - pow(beta2, (double)step);
    a.step_size = (float)(lr / bc1);
    a.bc2_sqrt = (float)sqrt(bc2);
    adamw_clip_kernel<<<stream_grid(n >> 2), TPB, 0, hv_stream(stream)>>>(p, g, m, v, n, a, clip_coef);
    HV_LAUNCH_CHECK();
    return HV_OK;
}

// hv_adamw_clip with the step count and the learning rate in DEVICE memory (state: 2 doubles {steps done, lr}; derived:
// 3 floats of scratch), so that a captured CUDA graph of the whole train step can be replayed: each launch advances
// state[0] by one and uses bias corrections of the new step; the host changes lr by writing state[1].
extern "C" int hv_adamw_clip_devstate(float* p, const float* g, float* m, float* v, int64_t n, double* state, float* derived,
                                      double beta1, double beta2, double eps, double weight_decay, float grad_scale,
                                      const float* clip_coef, void* stream) {
    HV_REQUIRE(n > 0 && p && g && m && v && state && derived, "hv_adamw_clip_devstate: bad arguments");
    HV_REQUIRE(((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) | reinterpret_cast<uintptr_t>(v)) & 15) == 0,
               "hv_adamw_clip_devstate: arenas must be 16-byte aligned");
    AdamArgs a;
    a.decay = 1.f; a.step_size = 0.f; a.bc2_sqrt = 1.f;  // replaced by `derived` in the kernel
    a.omb1 = (float)(1.0 - beta1); a.beta2 = (float)beta2; a.omb2 = (float)(1.0 - beta2); a.eps = (float)eps; a.grad_scale = grad_scale;
    cudaStream_t st = hv_stream(stream);
    adam_state_kernel<<<1, 1, 0, st>>>(state, derived, beta1, beta2, weight_decay);
    HV_LAUNCH_CHECK();
    adamw_clip_kernel<<<stream_grid(n >> 2), TPB, 0, st>>>(p, g, m, v, n, a, clip_coef, derived);
    HV_LAUNCH_CHECK();
    return HV_OK;
}

// x *= coef[0] * extra (materialises the clip in place when a caller wants torch's observable behaviour)
extern "C" int hv_scale_by_device(float* x, int64_t n, const float* coef, float extra, void* stream) {
    HV_REQUIRE(n >= 0 && x && coef, "hv_scale_by_device: bad arguments");
    if (n == 0) return HV_OK;
    scale_by_device_kernel<<<stream_grid(n), TPB, 0, hv_stream(stream)>>>(x, n, coef, extra);
    HV_LAUNCH_CHECK();
    return HV_OK;
}
