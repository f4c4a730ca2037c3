// Library-level plumbing of the C ABI: error string, device check, launch counter.
#include <cstdlib>
#include "hv_common.cuh"
#include <atomic>
#include <string.h>

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

void hv_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
void hv_count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

bool hv_pdl_enabled() {
    // opt-in: measured on one B200, same box, graph replay (profiles/r02_l_pdl_ab.txt): 11.28 -> 10.97 clips/s and 97.3 -> 99.0 ms
    // per train step WITH programmatic dependent launch; only the eager train step gained (113.5 -> 111.3 ms)
    static const bool on = [] { const char* e = getenv("HV_PDL"); return e && atoi(e) != 0; }();
    return on;
}

int hv_device_slot() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0) dev = 0;
    return dev < HV_MAX_DEVICES ? dev : HV_MAX_DEVICES - 1;
}

int hv_num_sms() {
    static int sms[HV_MAX_DEVICES] = {};  // per device: one process may drive several GPUs
    const int slot = hv_device_slot();
    if (sms[slot] == 0) {
        int dev = 0, n = 0;
        if (cudaGetDevice(&dev) == cudaSuccess &&
            cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0)
            sms[slot] = n;
        else
            sms[slot] = 148;
    }
    return sms[slot];
}

extern "C" const char* hv_last_error(void) { return g_err; }
extern "C" int hv_version(void) { return 100; }
extern "C" int64_t hv_launch_count(void) { return (int64_t)g_launches.load(); }

extern "C" int hv_check_device(void) {
    int dev = 0, major = 0;
    HV_CUDA(cudaGetDevice(&dev));
    HV_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
    if (major != 10) {
        hv_set_error("hivae_b200 requires an sm_100 (B200) device, found compute capability %d.x", major);
        return HV_EARCH;
    }
    return HV_OK;
}
