// Shared argument block of the attention kernels.
#pragma once
#include "hv_common.cuh"

struct AttnArgs {
    const void *q, *k, *v;
    void* out;
    long long ldq, ldk, ldv, ldo;
    int heads, hd, Lq, Lkv;
    HvSeqMap qmap, kvmap;
    float scale;
    const float* mask;
};

__device__ __forceinline__ long long seq_row(const HvSeqMap& m, long long s, int tok) {
    return (s / m.sdiv) * m.sa + (s % m.sdiv) * m.sb + (long long)tok * m.sc;
}

