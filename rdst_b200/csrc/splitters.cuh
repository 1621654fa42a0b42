// splitters.cuh — the two small kernels behind full-resolution splitters for the sharded
// sort: sampling key prefixes and counting records against a set of
// splitter values.  The reference has no distributed path; these serve the widening
// BASELINE.json configs[4] names (skewed keys sharded over the GPUs of one box), where one
// digit of the top level can hold most of the keys and digit-boundary splitters cannot
// balance the ranks.
//
// Key prefix = the record's top min(LEVELS, 8) levels as one order-preserving 64-bit integer
// (level LEVELS-1 most significant), i.e. what RadixKey::get_level (reference
// src/radix_key.rs:1-5) yields for those levels, concatenated.
#pragma once
#include "onesweep.cuh"

namespace b200 {

struct PrefixSrc {
    const uint32_t* src;
    uint64_t n;
    uint32_t words;     // record size in 32-bit words
    uint32_t is_float;
    SplitParams sp;     // word indices, xor, mask (+ splitters for the count kernel)
};

__device__ __forceinline__ uint64_t prefix_of(const PrefixSrc& ps, uint64_t rec) {
    const uint32_t* r = ps.src + rec * ps.words;
    const uint32_t lo = r[ps.sp.lo_word];
    const uint32_t hi = ps.sp.hi_word < ps.words ? r[ps.sp.hi_word] : 0u;
    uint64_t k = (uint64_t(hi) << 32) | lo;
    if (ps.is_float) {
        const uint32_t neg = (r[ps.sp.sign_word] >> ps.sp.sign_shift) & 1u;
        k ^= neg ? ~0ull : ps.sp.xorv;
    } else {
        k ^= ps.sp.xorv;
    }
    return k & ps.sp.mask;
}

// out[i] = prefix of record i * stride, i < count
static __global__ void prefix_sample_kernel(const PrefixSrc ps, uint64_t stride, uint64_t count,
                                            unsigned long long* out) {
    const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < count) out[i] = prefix_of(ps, i * stride);
}

// less[g] = #records with prefix < split[g], equal[g] = #records with prefix == split[g].
// The splitters ascend, so a record is described by lo = #splitters < prefix and
// up = #splitters <= prefix: it is "equal" to splitters lo .. up-1 and "less" than up .. nsplit-1.
// Each CTA histograms (lo, up) in shared memory (two 16-bin tables) and prefix-sums at the end.
constexpr int COUNT_THREADS = 256;
static __global__ void __launch_bounds__(COUNT_THREADS) splitter_count_kernel(const PrefixSrc ps,
                                                                             unsigned long long* less,
                                                                             unsigned long long* equal) {
    __shared__ uint32_t h_lo[MAX_SPLITTERS + 1], h_up[MAX_SPLITTERS + 1];  // a CTA sees < 2^32 records
    if (threadIdx.x <= MAX_SPLITTERS) h_lo[threadIdx.x] = h_up[threadIdx.x] = 0;
    __syncthreads();
    const uint64_t stride = uint64_t(gridDim.x) * COUNT_THREADS;
    auto count = [&](uint64_t k) {
        uint32_t lo = 0, up = 0;
#pragma unroll 1
        for (uint32_t g = 0; g < ps.sp.nsplit; ++g) {
            const uint64_t sv = ps.sp.split[g];
            lo += sv < k ? 1u : 0u;
            up += sv <= k ? 1u : 0u;
        }
        atomicAdd(&h_lo[lo], 1u);
        atomicAdd(&h_up[up], 1u);
    };
    uint64_t rec = uint64_t(blockIdx.x) * COUNT_THREADS + threadIdx.x;
    for (; rec + 3 * stride < ps.n; rec += 4 * stride) {  // four loads in flight per thread
        uint64_t k[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) k[u] = prefix_of(ps, rec + u * stride);
#pragma unroll
        for (int u = 0; u < 4; ++u) count(k[u]);
    }
    for (; rec < ps.n; rec += stride) count(prefix_of(ps, rec));
    __syncthreads();
    if (threadIdx.x < ps.sp.nsplit) {
        // less[g] = #(up <= g); equal[g] = #(lo <= g) - #(up <= g)
        unsigned long long cu = 0, cl = 0;
        for (uint32_t x = 0; x <= threadIdx.x; ++x) {
            cu += h_up[x];
            cl += h_lo[x];
        }
        if (cu) atomicAdd(&less[threadIdx.x], cu);
        if (cl - cu) atomicAdd(&equal[threadIdx.x], cl - cu);
    }
}

}  // namespace b200
