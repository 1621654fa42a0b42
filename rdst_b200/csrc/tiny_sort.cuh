// tiny_sort.cuh — K3-T, the sort of segments of up to 256 records: one WARP per segment.
//
// rdst ends its recursion with a comparison sort for chunks of up to 128 records (reference
// src/sorter.rs:33-45, src/sorts/comparative_sort.rs:29-119: `sort_by` on the level bytes for the stable
// case, `sort_unstable_by` otherwise).  The segment sorts K3 / K3-U give a whole CTA — barriers, a digit
// table to clear and scan — to every segment, which costs microseconds however small the segment is
// (profiles/r9_segments_bench_before_k3t.jsonl: 1 M segments of 64 u64 keys take 8.3 ms, 8 Grec/s; with K3-T 1.9 ms, of which ~0.8 ms is the upload of the segment list).  A warp needs
// none of that: the lanes load the segment (<= 8 records each), publish the 64-bit key images in shared
// memory, and every record counts the records that sort before it — key image smaller, or equal and
// earlier in the segment, so the order is the stable one — which is its position.  O(cnt^2) compares, all
// of them broadcast reads of shared memory; no barrier wider than the warp, nothing to clear.
#pragma once
#include "onesweep.cuh"

namespace b200 {

constexpr int TS_CAP = 256, TS_THREADS = 256, TS_WARPS = TS_THREADS / 32;

struct TinySortParams {
    uint32_t* data;           // sorted in place, segment by segment
    const uint2* seg_pairs;   // [nseg] {first record, records}, 2 <= records <= TS_CAP
    uint32_t nseg;
    SplitParams key;          // 64-bit order-preserving image of the whole key (LEVELS <= 8)
};

template <int WORDS, bool FLOATK, int IT>
__device__ __forceinline__ void tiny_sort_segment(uint32_t* g, uint32_t cnt, const SplitParams& key, unsigned long long* s_key, int lane) {
    Rec<WORDS> r[IT];
    unsigned long long k[IT];
    uint32_t rank[IT];
#pragma unroll
    for (int j = 0; j < IT; ++j) {
        const uint32_t e = uint32_t(j * 32 + lane);
        rank[j] = 0;
        k[j] = ~0ull;
        if (e < cnt) {
            r[j] = load_rec<WORDS>(g, e);
            k[j] = rec_prefix<WORDS, FLOATK>(key, r[j]);
            s_key[e] = k[j];
        }
    }
    __syncwarp();
    for (uint32_t i = 0; i < cnt; ++i) {
        const unsigned long long ki = s_key[i];  // same address in every lane: a broadcast
#pragma unroll
        for (int j = 0; j < IT; ++j) {
            const uint32_t e = uint32_t(j * 32 + lane);
            rank[j] += (ki < k[j] || (ki == k[j] && i < e)) ? 1u : 0u;
        }
    }
    __syncwarp();  // every lane is done with s_key, and every record of the segment sits in a register
#pragma unroll
    for (int j = 0; j < IT; ++j) {
        const uint32_t e = uint32_t(j * 32 + lane);
        if (e < cnt) store_rec<WORDS>(g, uint64_t(rank[j]), r[j]);
    }
}

template <int WORDS, bool FLOATK>
__global__ void __launch_bounds__(TS_THREADS) tiny_sort_kernel(const TinySortParams p) {
    __shared__ unsigned long long s_keys[TS_WARPS][TS_CAP];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t nw = gridDim.x * TS_WARPS;
    for (uint32_t seg = blockIdx.x * TS_WARPS + warp; seg < p.nseg; seg += nw) {
        const uint2 pr = p.seg_pairs[seg];
        uint32_t* g = p.data + size_t(pr.x) * WORDS;
        const uint32_t cnt = pr.y;  // warp-uniform
        if (cnt <= 32) tiny_sort_segment<WORDS, FLOATK, 1>(g, cnt, p.key, s_keys[warp], lane);
        else if (cnt <= 64) tiny_sort_segment<WORDS, FLOATK, 2>(g, cnt, p.key, s_keys[warp], lane);
        else if (cnt <= 128) tiny_sort_segment<WORDS, FLOATK, 4>(g, cnt, p.key, s_keys[warp], lane);
        else tiny_sort_segment<WORDS, FLOATK, 8>(g, cnt, p.key, s_keys[warp], lane);
        __syncwarp();
    }
}

}  // namespace b200
