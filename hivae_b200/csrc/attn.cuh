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
    float* lse_stats = nullptr;  // optional: [seq * heads + head][Lq][2] floats, element 0 of a row <- log2-domain log-sum-exp
                                 // (what hv_attention_bwd keeps in `stats`); only attn_tc2.cu writes it (hv_attention_lse)
};

__device__ __forceinline__ long long seq_row(const HvSeqMap& m, long long s, int tok) {
    return (s / m.sdiv) * m.sa + (s % m.sdiv) * m.sb + (long long)tok * m.sc;
}

