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

// less[g] = #records with prefix < split[g], equal[g] = #records with prefix == split[g]
constexpr int COUNT_THREADS = 256;
static __global__ void __launch_bounds__(COUNT_THREADS) splitter_count_kernel(const PrefixSrc ps,
                                                                             unsigned long long* less,
                                                                             unsigned long long* equal) {
    uint32_t lt[MAX_SPLITTERS], eq[MAX_SPLITTERS];
#pragma unroll
    for (int g = 0; g < MAX_SPLITTERS; ++g) lt[g] = eq[g] = 0;
    const uint64_t stride = uint64_t(gridDim.x) * COUNT_THREADS;
    // a thread sees at most n / stride + 1 records: 32-bit counters hold for any n the engine accepts per launch
    for (uint64_t rec = uint64_t(blockIdx.x) * COUNT_THREADS + threadIdx.x; rec < ps.n; rec += stride) {
        const uint64_t k = prefix_of(ps, rec);
#pragma unroll
        for (int g = 0; g < MAX_SPLITTERS; ++g) {
            if (uint32_t(g) < ps.sp.nsplit) {
                lt[g] += k < ps.sp.split[g] ? 1u : 0u;
                eq[g] += k == ps.sp.split[g] ? 1u : 0u;
            }
        }
    }
    __shared__ unsigned long long s_lt[MAX_SPLITTERS], s_eq[MAX_SPLITTERS];
    if (threadIdx.x < MAX_SPLITTERS) s_lt[threadIdx.x] = s_eq[threadIdx.x] = 0;
    __syncthreads();
#pragma unroll
    for (int g = 0; g < MAX_SPLITTERS; ++g) {
        if (uint32_t(g) < ps.sp.nsplit) {
            uint32_t a = lt[g], b = eq[g];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                a += __shfl_down_sync(0xffffffffu, a, o);
                b += __shfl_down_sync(0xffffffffu, b, o);
            }
            if ((threadIdx.x & 31) == 0) {
                atomicAdd(&s_lt[g], (unsigned long long)a);
                atomicAdd(&s_eq[g], (unsigned long long)b);
            }
        }
    }
    __syncthreads();
    if (threadIdx.x < ps.sp.nsplit) {
        atomicAdd(&less[threadIdx.x], s_lt[threadIdx.x]);
        atomicAdd(&equal[threadIdx.x], s_eq[threadIdx.x]);
    }
}

}  // namespace b200
