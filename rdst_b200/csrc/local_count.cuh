// local_count.cuh — K3-U, the shared-memory segment sort of the hybrid route for records that are
// nothing but their key (every byte is a level: u32 / u64 / i64 / f32 / f64 ...).
//
// rdst finishes the small buckets of its MSB-first recursion in cache (reference src/sorter.rs:121-157,
// src/sorts/lsb_sort.rs:38-128) and, below 128 records, with a comparison sort
// (src/sorts/comparative_sort.rs:29-119).  K3-U is both at once, for a segment that fits one SM:
//
//   count    ONE counting-sort level on a WIDE digit — the top b bits of what is still unsorted, b
//            chosen per segment so that the table has 2..4 counters per record (b <= 15; 16-bit
//            counters packed two per word, RED.ADD without a return value);
//   scan     every thread owns a contiguous run of counters (walked in a rotated order: no bank
//            conflicts), one block scan, counters := exclusive bases;
//   scatter  one ATOMS.ADD per record returns its position; the record goes to the staging buffer;
//            afterwards counter d holds the END of digit d's run (= the start of digit d + 1's);
//   fix-up + write-out, fused: records are read back in position order; a record whose digit run
//            holds c > 1 records ranks itself among the other c - 1 by the whole key image
//            (a comparison sort of the run, one record per thread, neighbouring addresses) and is
//            stored straight to its final place in global memory.
//
// A wide digit leaves runs of expected length < 0.5, so the radix part is one level instead of the
// stable engine's two (local_sort.cuh) and nothing has to be stable: records with equal keys are
// identical.  Runs longer than LC_MAXRUN are legal only if all their keys are equal (heavy
// duplicates: checked by a warp, nothing to do then); otherwise the segment is appended to
// `slow_list` and the stable segment sort (local_sort_kernel in list mode) handles it afterwards.
#pragma once
#include "pass_atomic.cuh"

namespace b200 {

constexpr int LC_ITEMS = 32, LC_MAXRUN = 16;

struct LocalCountParams {
    uint32_t* data;           // sorted in place, segment by segment
    const uint64_t* seg_off;  // [nseg + 1] record offsets of the segments
    uint32_t nseg;
    uint32_t* ticket;         // zeroed before the launch
    SplitParams key;          // 64-bit order-preserving image of the whole key
    uint32_t top_bit;         // image bits [0, top_bit) are unsorted inside a segment (>= 1)
    uint32_t bmax;            // cap on the digit width (<= CBITS of the instantiation)
    uint32_t* slow_list;      // [0] = count, [1 ..] = ids of the segments left to the stable segment sort
};

template <int WORDS, int THREADS, int CBITS>
struct LocalCountSmem {
    static constexpr int CAP = THREADS * LC_ITEMS;
    static constexpr size_t OUT_BYTES = size_t(CAP) * WORDS * 4;
    static constexpr size_t CNT_BYTES = (size_t(1) << CBITS) * 2;
    static constexpr size_t MISC_BYTES = 256;
    static constexpr size_t TOTAL = OUT_BYTES + CNT_BYTES + MISC_BYTES;
};

// The order-preserving image of a record that is nothing but its key (1 or 2 words, key at offset 0, every
// byte a level): raw ^ xorv for integers (xorv = the sign bit for signed keys), IEEE total order for floats
// (reference src/radix_key_impl.rs:162-185).  Equal to rec_prefix() for these records, without its run-time
// word selection.
template <int WORDS>
struct KeyImage;
template <>
struct KeyImage<1> {
    using T = uint32_t;
};
template <>
struct KeyImage<2> {
    using T = uint64_t;
};
template <int WORDS, bool FLOATK>
__device__ __forceinline__ typename KeyImage<WORDS>::T key_image(const Rec<WORDS>& r, typename KeyImage<WORDS>::T xorv) {
    if constexpr (WORDS == 1) {
        if constexpr (FLOATK) return r.w[0] ^ (uint32_t(int32_t(r.w[0]) >> 31) | 0x80000000u);
        return r.w[0] ^ xorv;
    } else {
        const uint64_t k = (uint64_t(r.w[1]) << 32) | r.w[0];
        if constexpr (FLOATK) return k ^ (uint64_t(int64_t(k) >> 63) | 0x8000000000000000ull);
        return k ^ xorv;
    }
}

template <int WORDS, bool FLOATK, int THREADS, int CBITS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) local_count_kernel(const LocalCountParams p) {
    using SM = LocalCountSmem<WORDS, THREADS, CBITS>;
    using KT = typename KeyImage<WORDS>::T;
    constexpr int ITEMS = LC_ITEMS, WARPS = THREADS / 32;
    extern __shared__ __align__(128) uint8_t smem_raw[];
    uint32_t* s_out = reinterpret_cast<uint32_t*>(smem_raw);
    uint32_t* s_cnt = reinterpret_cast<uint32_t*>(smem_raw + SM::OUT_BYTES);
    uint8_t* misc = smem_raw + SM::OUT_BYTES + SM::CNT_BYTES;
    volatile uint32_t* s_seg = reinterpret_cast<volatile uint32_t*>(misc);  // [0] first id, [1] next id, [2] long run seen, [3] slow
    uint32_t* s_wtot = reinterpret_cast<uint32_t*>(misc + 64);              // WARPS warp totals

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t cnt_base = smem_u32(s_cnt);
    const KT xorv = KT(p.key.xorv);
    for (int i = tid; i < int(SM::CNT_BYTES / 16); i += THREADS) reinterpret_cast<uint4*>(s_cnt)[i] = make_uint4(0, 0, 0, 0);

    uint32_t cnt = 0;
    uint32_t* g = nullptr;
    Rec<WORDS> r[ITEMS];
    auto fetch = [&](uint32_t sg) {  // header + records of segment sg (order is free: coalesced element order)
        cnt = 0;
        if (sg >= p.nseg) return;
        const uint64_t base = p.seg_off[sg];
        cnt = uint32_t(p.seg_off[sg + 1] - base);
        g = p.data + base * WORDS;
        if (cnt <= 1) return;
#pragma unroll
        for (int j = 0; j < ITEMS; ++j) {
            const uint32_t e = uint32_t(j * THREADS + tid);
            if (e < cnt) r[j] = load_rec<WORDS>(g, e);
        }
    };
    if (tid == 0) s_seg[0] = atomicAdd(p.ticket, 1u);
    __syncthreads();
    uint32_t seg = s_seg[0];
    fetch(seg);
    while (seg < p.nseg) {
        if (tid == 0) {
            s_seg[1] = atomicAdd(p.ticket, 1u);
            s_seg[2] = 0;
            s_seg[3] = 0;
        }
        __syncthreads();  // (0) previous segment done, counters clean, next id visible
        const uint32_t next_seg = s_seg[1];
        if (cnt <= 1) {
            seg = next_seg;
            fetch(seg);
            __syncthreads();  // everyone has read the next id before thread 0 claims another one
            continue;
        }
        const uint32_t cur_cnt = cnt, cur_seg = seg;
        uint32_t* const cur_g = g;
        // records this thread holds: elements tid, tid + THREADS, ...
        const int mine = int((cur_cnt - uint32_t(tid) + THREADS - 1) / THREADS) * int(uint32_t(tid) < cur_cnt);
        // digit width of this segment: 2..4 counters per record, within [1, min(bmax, top_bit)]
        uint32_t b = 33u - uint32_t(__clz(int(cur_cnt - 1)));  // ceil(log2 cnt) + 1 for cnt >= 2
        b = b < p.bmax ? b : p.bmax;
        b = b < p.top_bit ? b : p.top_bit;
        const uint32_t shift = p.top_bit - b;
        const uint32_t dmask = (1u << b) - 1u;
        const uint32_t NW = b ? (1u << (b - 1)) : 1u;  // packed counter words
        auto digit_of = [&](const Rec<WORDS>& x) { return uint32_t(key_image<WORDS, FLOATK>(x, xorv) >> shift) & dmask; };

        // ---- count ----
#pragma unroll
        for (int j = 0; j < ITEMS; ++j) {
            if (j < mine) {
                const uint32_t d = digit_of(r[j]);
                red_add_shared(cnt_base + (d >> 1) * 4, 1u << ((d & 1u) * 16));
            }
        }
        __syncthreads();  // (1)
        // ---- scan: counters := exclusive bases ----
        const bool vec = NW >= uint32_t(WARPS) * 128u;  // block-uniform
        if (vec) {
            // every warp owns a contiguous run of words and walks it 128 words (one uint4 per lane) at a time
            const uint32_t per_warp = NW / WARPS;
            uint4* base4 = reinterpret_cast<uint4*>(s_cnt + uint32_t(warp) * per_warp);
            const uint32_t rounds = per_warp / 128;
            uint32_t acc = 0, mx = 0;  // packed halves: per-lane sums stay below 2^16 (a segment has <= 16 384 records)
            for (uint32_t q = 0; q < rounds; ++q) {
                const uint4 v = base4[q * 32 + lane];
                acc += v.x + v.y + v.z + v.w;
                mx = __vmaxu2(mx, __vmaxu2(__vmaxu2(v.x, v.y), __vmaxu2(v.z, v.w)));
            }
            uint32_t tot = (acc & 0xffffu) + (acc >> 16);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
            if (lane == 0) s_wtot[warp] = tot;
            if (max(mx & 0xffffu, mx >> 16) > uint32_t(LC_MAXRUN)) s_seg[2] = 1;
            __syncthreads();  // (2)
            uint32_t run = 0;  // records before this warp's words
#pragma unroll
            for (int w = 0; w < WARPS; ++w)
                if (w < warp) run += s_wtot[w];
            for (uint32_t q = 0; q < rounds; ++q) {
                uint4 v = base4[q * 32 + lane];
                const uint32_t packed = v.x + v.y + v.z + v.w;
                const uint32_t mysum = (packed & 0xffffu) + (packed >> 16);
                uint32_t incl = mysum;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
                    if (lane >= o) incl += y;
                }
                uint32_t at = run + incl - mysum;
                run += __shfl_sync(0xffffffffu, incl, 31);
                // word {c1:c0} -> {at + c0 : at}; at += c0 + c1
                uint32_t t;
                t = v.x, v.x = at * 0x10001u + (t << 16), at += (t & 0xffffu) + (t >> 16);
                t = v.y, v.y = at * 0x10001u + (t << 16), at += (t & 0xffffu) + (t >> 16);
                t = v.z, v.z = at * 0x10001u + (t << 16), at += (t & 0xffffu) + (t >> 16);
                t = v.w, v.w = at * 0x10001u + (t << 16);
                base4[q * 32 + lane] = v;
            }
        } else {
            // small tables: thread t owns words [t * W, (t + 1) * W), walked from word (t mod W) on and around
            const uint32_t W = NW >= uint32_t(THREADS) ? NW / THREADS : 1u;
            const bool scanner = uint32_t(tid) < NW;
            const uint32_t w0 = uint32_t(tid) * W, rot = uint32_t(tid) & (W - 1);
            uint32_t total = 0, wrapped = 0, mx = 0;  // wrapped = sum of the words before `rot`
            if (scanner) {
                for (uint32_t k = 0; k < W; ++k) {
                    const uint32_t idx = (rot + k) & (W - 1);
                    const uint32_t v = s_cnt[w0 + idx];
                    const uint32_t lo = v & 0xffffu, hi = v >> 16;
                    mx = max(mx, max(lo, hi));
                    total += lo + hi;
                    if (idx < rot) wrapped += lo + hi;
                }
                if (mx > uint32_t(LC_MAXRUN)) s_seg[2] = 1;
            }
            uint32_t incl = total;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += y;
            }
            if (lane == 31) s_wtot[warp] = incl;
            __syncthreads();  // (2)
            if (scanner) {
                uint32_t excl = incl - total;
#pragma unroll
                for (int w = 0; w < WARPS; ++w)
                    if (w < warp) excl += s_wtot[w];
                uint32_t run = excl + wrapped;
                for (uint32_t k = 0; k < W; ++k) {
                    const uint32_t idx = (rot + k) & (W - 1);
                    if (k && idx == 0) run = excl;  // wrapped around: back at the first word of my run
                    const uint32_t v = s_cnt[w0 + idx];
                    const uint32_t lo = v & 0xffffu, hi = v >> 16;
                    s_cnt[w0 + idx] = run | ((run + lo) << 16);
                    run += lo + hi;
                }
            }
        }
        __syncthreads();  // (3)
        // ---- scatter: the atomic hands out positions; afterwards counter d = end of digit d's run ----
        // (in batches of 8: eight atomics in flight per thread before the first dependent store)
#pragma unroll
        for (int j0 = 0; j0 < ITEMS; j0 += 8) {
            uint32_t pos[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int j = j0 + u;
                pos[u] = 0;
                if (j < mine) {
                    const uint32_t d = digit_of(r[j]);
                    const uint32_t sh = (d & 1u) * 16;
                    pos[u] = (atom_add_shared(cnt_base + (d >> 1) * 4, 1u << sh) >> sh) & 0xffffu;
                }
            }
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int j = j0 + u;
                if (j < mine) store_rec<WORDS>(s_out, pos[u], r[j]);
            }
        }
        __syncthreads();  // (4)
        const uint16_t* c16 = reinterpret_cast<const uint16_t*>(s_cnt);
        if (shift != 0 && s_seg[2]) {
            // a run longer than LC_MAXRUN: fine if all its keys are equal (nothing to order), else the
            // stable segment sort takes this segment
            const uint32_t NC = dmask + 1u;
            for (uint32_t d0 = uint32_t(warp) * 32; d0 < NC; d0 += WARPS * 32) {
                const uint32_t d = d0 + lane;
                uint32_t start = 0, end = 0;
                if (d < NC) {
                    end = c16[d];
                    start = d ? c16[d - 1] : 0u;
                }
                uint32_t m = __ballot_sync(0xffffffffu, end - start > uint32_t(LC_MAXRUN));
                while (m) {
                    const int src = __ffs(m) - 1;
                    m &= m - 1;
                    const uint32_t s = __shfl_sync(0xffffffffu, start, src), e = __shfl_sync(0xffffffffu, end, src);
                    const KT k0 = key_image<WORDS, FLOATK>(load_rec<WORDS>(s_out, s), xorv);
                    bool same = true;
                    for (uint32_t i = s + 1 + lane; i < e; i += 32)
                        same = same && key_image<WORDS, FLOATK>(load_rec<WORDS>(s_out, i), xorv) == k0;
                    if (!__all_sync(0xffffffffu, same) && lane == 0) s_seg[3] = 1;
                }
            }
            __syncthreads();
        }
        const bool slow = s_seg[3] != 0;
        // ---- prefetch the next segment, then fix-up + write-out ----
        seg = next_seg;
        fetch(seg);
        if (slow) {
            if (tid == 0) {
                const uint32_t at = atomicAdd(p.slow_list, 1u);
                p.slow_list[1 + at] = cur_seg;
            }
        } else if (shift == 0) {
            // the digit was everything that was left of the key: the staging buffer is sorted
            for (uint32_t i = tid; i < cur_cnt; i += THREADS) store_rec<WORDS>(cur_g, uint64_t(i), load_rec<WORDS>(s_out, i));
        } else {
            // four records per thread and trip, loads first: the run bounds and the neighbours' keys are
            // chains of dependent shared-memory loads
            constexpr int FB = 4;
            for (uint32_t i0 = tid; i0 < cur_cnt; i0 += FB * THREADS) {
                Rec<WORDS> x[FB];
                KT kx[FB];
                uint32_t start[FB], end[FB];
#pragma unroll
                for (int u = 0; u < FB; ++u) {
                    const uint32_t i = i0 + u * THREADS;
                    if (i < cur_cnt) x[u] = load_rec<WORDS>(s_out, i);
                }
#pragma unroll
                for (int u = 0; u < FB; ++u) {
                    const uint32_t i = i0 + u * THREADS;
                    start[u] = end[u] = 0;
                    kx[u] = 0;
                    if (i < cur_cnt) {
                        kx[u] = key_image<WORDS, FLOATK>(x[u], xorv);
                        const uint32_t d = uint32_t(kx[u] >> shift) & dmask;
                        end[u] = c16[d];
                        start[u] = d ? c16[d - 1] : 0u;
                    }
                }
#pragma unroll
                for (int u = 0; u < FB; ++u) {
                    const uint32_t i = i0 + u * THREADS;
                    if (i < cur_cnt) {
                        uint32_t pos = i;
                        const uint32_t c = end[u] - start[u];
                        if (c > 1 && c <= uint32_t(LC_MAXRUN)) {
                            pos = start[u];
                            for (uint32_t j = start[u]; j < end[u]; ++j) {
                                const KT ky = key_image<WORDS, FLOATK>(load_rec<WORDS>(s_out, j), xorv);
                                pos += (ky < kx[u] || (ky == kx[u] && j < i)) ? 1u : 0u;
                            }
                        }
                        store_rec<WORDS>(cur_g, uint64_t(pos), x[u]);
                    }
                }
            }
        }
        __syncthreads();  // (5) counters no longer read
        for (uint32_t i = tid; i < NW / 4; i += THREADS) reinterpret_cast<uint4*>(s_cnt)[i] = make_uint4(0, 0, 0, 0);
        if (NW < 4 && uint32_t(tid) < NW) s_cnt[tid] = 0;
    }
}

}  // namespace b200
