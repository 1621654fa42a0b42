// heavy.cuh — oversized segments of the key-only hybrid route: the heavy-duplicate case.
//
// After the two MSD passes a segment normally fits one SM (<= 16 384 records).  Skewed inputs break that:
// a key that occurs a million times makes its whole segment a million records long (BASELINE.json
// configs[4], "skewed Zipf"; rdst meets the same shape with its skew-aware tuner rules,
// reference src/tuners/standard_tuner.rs:20-44, and with regions_sort / ska_sort's in-place swaps that cost
// nothing for records already in place, src/sorts/ska_sort.rs:13-57).  Such a segment is one dominant key
// plus a few thousand others, so it is not sorted but SPLIT three ways around a pivot key — less, equal,
// greater: the equal part (nearly everything) is final as it stands, equal keys being identical records;
// the two small parts go to the segment sort (K3-U).  Parts that are still too large take another round
// (another pivot); after a few rounds whatever is left sends the whole array back to the LSD passes.
//
// Kernels, all over "ranges" (oversized segments or parts) cut into tiles:
//   pivot_kernel      a pivot per range: the value seen twice among five samples, else their median
//   split_ends_kernel ONE read of a range: records smaller than the pivot are appended at the front of the range's
//                     slot in the other buffer, larger ones at its back (positions handed out by two atomic
//                     counters per range; order inside a part is free); equal records are not moved at all
//   fill_kernel       the equal part is WRITTEN, not copied: equal keys are identical records, so the middle of
//                     the range in `data` is filled with the pivot record
//   copy_ranges       a list of ranges copied between the two buffers (the give-up path only)
#pragma once
#include "local_count.cuh"

namespace b200 {

constexpr int HV_THREADS = 256, HV_ITEMS = 16, HV_TILE = HV_THREADS * HV_ITEMS;  // 4096 records per tile

struct HeavyRanges {
    const uint32_t* src;              // the buffer the ranges live in
    uint32_t* dst;                    // the other buffer: receives the smaller / larger records
    const uint32_t* range_off;        // [nranges] first record
    const uint32_t* range_cnt;        // [nranges]
    const uint32_t* range_tile_start; // [nranges + 1] first tile of every range
    uint32_t nranges, ntiles;
    unsigned long long* pivot;        // [nranges] the pivot RECORD (raw bits)
    uint32_t* ends;                   // [nranges][2] records appended at the front / at the back so far (zeroed before)
    unsigned long long xorv;          // key image = raw ^ xorv (integers)
};

template <int WORDS>
__device__ __forceinline__ Rec<WORDS> rec_of_bits(unsigned long long v) {
    Rec<WORDS> r;
    r.w[0] = uint32_t(v);
    if constexpr (WORDS == 2) r.w[1] = uint32_t(v >> 32);
    return r;
}
template <int WORDS>
__device__ __forceinline__ unsigned long long bits_of_rec(const Rec<WORDS>& r) {
    if constexpr (WORDS == 2) return (unsigned long long)r.w[0] | ((unsigned long long)r.w[1] << 32);
    return r.w[0];
}

template <int WORDS, bool FLOATK>
__global__ void pivot_kernel(const HeavyRanges p) {
    using KT = typename KeyImage<WORDS>::T;
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= p.nranges) return;
    const uint32_t off = p.range_off[r], cnt = p.range_cnt[r];
    KT k[5];
    unsigned long long raw[5];
#pragma unroll
    for (int i = 0; i < 5; ++i) {
        const uint32_t at = off + uint32_t((uint64_t(cnt) * (2 * i + 1)) / 10);
        const Rec<WORDS> x = load_rec<WORDS>(p.src, at);
        k[i] = key_image<WORDS, FLOATK>(x, KT(p.xorv));
        raw[i] = bits_of_rec<WORDS>(x);
    }
    // sort the five by key image (a dominant key shows up at least twice, adjacent after sorting)
#pragma unroll
    for (int a = 0; a < 5; ++a)
#pragma unroll
        for (int b = 0; b + 1 < 5 - a; ++b)
            if (k[b + 1] < k[b]) {
                const KT t = k[b];
                k[b] = k[b + 1];
                k[b + 1] = t;
                const unsigned long long u = raw[b];
                raw[b] = raw[b + 1];
                raw[b + 1] = u;
            }
    unsigned long long pv = raw[2];
    int best = 1;
    int run = 1;
#pragma unroll
    for (int i = 1; i < 5; ++i) {
        run = k[i] == k[i - 1] ? run + 1 : 1;
        if (run > best) best = run, pv = raw[i];
    }
    p.pivot[r] = pv;
}

__device__ __forceinline__ uint32_t hv_find_range(const uint32_t* tile_start, uint32_t nranges, uint32_t t) {
    uint32_t lo = 0, hi = nranges;  // tile_start[lo] <= t < tile_start[hi]
    while (hi - lo > 1) {
        const uint32_t mid = (lo + hi) >> 1;
        if (__ldg(tile_start + mid) <= t) lo = mid; else hi = mid;
    }
    return lo;
}

template <int WORDS, bool FLOATK>
__global__ void __launch_bounds__(HV_THREADS) split_ends_kernel(const HeavyRanges p) {
    using KT = typename KeyImage<WORDS>::T;
    __shared__ uint32_t s_w[2][HV_THREADS / 32];  // per warp: smaller / larger records of the tile
    __shared__ uint32_t s_base[2];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (uint32_t t = blockIdx.x; t < p.ntiles; t += gridDim.x) {
        const uint32_t r = hv_find_range(p.range_tile_start, p.nranges, t);
        const uint32_t off = __ldg(p.range_off + r), cnt = __ldg(p.range_cnt + r);
        const uint32_t first = (t - __ldg(p.range_tile_start + r)) * HV_TILE;
        const uint32_t valid = min(uint32_t(HV_TILE), cnt - first);
        const uint32_t* g = p.src + size_t(off + first) * WORDS;
        const KT pv = key_image<WORDS, FLOATK>(rec_of_bits<WORDS>(p.pivot[r]), KT(p.xorv));
        // the tile's records and their classes (0 smaller, 1 equal, 2 larger, 3 none) stay in registers
        Rec<WORDS> x[HV_ITEMS];
        uint32_t cls = 0;  // two bits per record
        uint32_t nlt = 0, ngt = 0;
#pragma unroll
        for (int j = 0; j < HV_ITEMS; ++j) {
            const uint32_t e = uint32_t(j * HV_THREADS + tid);
            uint32_t c = 3;
            if (e < valid) {
                x[j] = load_rec<WORDS>(g, e);
                const KT k = key_image<WORDS, FLOATK>(x[j], KT(p.xorv));
                c = k < pv ? 0u : (k == pv ? 1u : 2u);
            }
            cls |= c << (2 * j);
            nlt += c == 0 ? 1u : 0u;
            ngt += c == 2 ? 1u : 0u;
        }
        // thread ranks inside the tile: warp scan + warp totals (nearly every tile has nothing to move)
        uint32_t ilt = nlt, igt = ngt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t a = __shfl_up_sync(0xffffffffu, ilt, o), b = __shfl_up_sync(0xffffffffu, igt, o);
            if (lane >= o) ilt += a, igt += b;
        }
        if (lane == 31) s_w[0][warp] = ilt, s_w[1][warp] = igt;
        __syncthreads();
        uint32_t blt = 0, bgt = 0, tlt = 0, tgt = 0;
#pragma unroll
        for (int w = 0; w < HV_THREADS / 32; ++w) {
            const uint32_t a = s_w[0][w], b = s_w[1][w];
            if (w < warp) blt += a, bgt += b;
            tlt += a, tgt += b;
        }
        if (tid == 0) {
            s_base[0] = tlt ? atomicAdd(p.ends + 2 * size_t(r), tlt) : 0u;
            s_base[1] = tgt ? atomicAdd(p.ends + 2 * size_t(r) + 1, tgt) : 0u;
        }
        __syncthreads();
        if (nlt | ngt) {
            uint32_t at_lt = off + s_base[0] + blt + ilt - nlt;                  // front of the slot, growing upwards
            uint32_t at_gt = off + cnt - 1 - (s_base[1] + bgt + igt - ngt);      // back of the slot, growing downwards
#pragma unroll
            for (int j = 0; j < HV_ITEMS; ++j) {
                const uint32_t c = (cls >> (2 * j)) & 3u;
                if (c == 0) store_rec<WORDS>(p.dst, uint64_t(at_lt++), x[j]);
                if (c == 2) store_rec<WORDS>(p.dst, uint64_t(at_gt--), x[j]);
            }
        }
        __syncthreads();  // s_w / s_base are reused by the next tile
    }
}

// chunks[i] = {first record, records} (<= 65 536 records each), owner[i] = the range whose pivot record fills it
template <int WORDS>
__global__ void __launch_bounds__(256) fill_kernel(uint32_t* data, const uint2* chunks, const uint32_t* owner, uint32_t nchunks,
                                                   const unsigned long long* pivot) {
    for (uint32_t c = blockIdx.x; c < nchunks; c += gridDim.x) {
        const uint2 ch = chunks[c];
        const Rec<WORDS> pv = rec_of_bits<WORDS>(pivot[owner[c]]);
        for (uint32_t i = threadIdx.x; i < ch.y; i += blockDim.x) store_rec<WORDS>(data, uint64_t(ch.x) + i, pv);
    }
}

// chunks[i] = {first record, records} (<= 65 536 records each): dst[first .. ) = src[first .. )
template <int WORDS>
__global__ void __launch_bounds__(256) copy_ranges_kernel(const uint32_t* src, uint32_t* dst, const uint2* chunks, uint32_t nchunks) {
    for (uint32_t c = blockIdx.x; c < nchunks; c += gridDim.x) {
        const uint2 ch = chunks[c];
        for (uint32_t i = threadIdx.x; i < ch.y; i += blockDim.x)
            store_rec<WORDS>(dst, uint64_t(ch.x) + i, load_rec<WORDS>(src, ch.x + i));
    }
}

}  // namespace b200
