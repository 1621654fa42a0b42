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
//   fix-up   the scan also lists every counter that holds >= 2 records (a "run" of the staging
//            buffer); one thread per run orders it by the whole key image — a compare-and-swap for
//            the common run of two, an insertion sort up to LC_MAXRUN records;
//   write-out  the staging buffer goes back where the segment came from, coalesced.
//
// A wide digit leaves runs of expected length < 0.5, so the radix part is one level instead of the
// stable engine's two (local_sort.cuh) and nothing has to be stable: records with equal keys are
// identical.  Runs longer than LC_MAXRUN are legal only if all their keys are equal (heavy
// duplicates: checked by a warp, nothing to do then); otherwise the segment is appended to
// `slow_list` and the stable segment sort (local_sort_kernel in list mode) handles it afterwards.
#pragma once
#include <type_traits>

#include "pass_atomic.cuh"

namespace b200 {

constexpr int LC_MAXRUN = 16;

struct LocalCountParams {
    uint32_t* data;           // where the sorted segments go (sorted in place when src is null)
    const uint32_t* src;      // where the segments are read from; nullptr = data
    const uint64_t* seg_off;  // [nseg + 1] record offsets of the segments
    const uint2* seg_pairs;   // alternative to seg_off: [nseg] {first record, records} of arbitrary, non-adjacent segments
    uint32_t seg_first;       // this launch sorts segments [seg_first, nseg)
    uint32_t nseg;
    uint32_t* ticket;         // zeroed before the launch
    SplitParams key;          // 64-bit order-preserving image of the whole key
    uint32_t top_bit;         // image bits [0, top_bit) are unsorted inside a segment (>= 1)
    uint32_t bmax;            // cap on the digit width (<= CBITS of the instantiation)
    uint32_t* slow_list;      // [0] = count, [1 ..] = ids of the segments left to the stable segment sort
    unsigned long long* prof; // optional [8] cycle counters per phase (thread 0 of every CTA), nullptr = off
};
// AUTOTOP instantiations (b200_sort_segments_device: segments of unknown origin): every segment finds its own top bit,
// the highest image bit in which ITS keys differ — the buckets of an MSD split share their leading bytes, and a digit
// cut from constant bits would put a whole segment into one counter.

template <int WORDS, int THREADS, int ITEMS, int CBITS>
struct LocalCountSmem {
    static constexpr int CAP = THREADS * ITEMS;
    static constexpr size_t OUT_BYTES = (size_t(CAP + 2) * WORDS * 4 + 127) / 128 * 128;  // + the shift that aligns record pairs
    static constexpr size_t CNT_BYTES = (size_t(1) << CBITS) * 2;
    static constexpr size_t RUN_BYTES = size_t(CAP / 2) * 4;  // a segment of c records has at most c / 2 runs of >= 2
    static constexpr size_t MISC_BYTES = 512;  // flags | warp totals (<= 32 warps) | auto_top: per-warp vary masks
    static constexpr size_t TOTAL = OUT_BYTES + CNT_BYTES + RUN_BYTES + MISC_BYTES;
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

template <int WORDS, bool FLOATK, int THREADS, int ITEMS, int CBITS, int MINB, bool AUTOTOP = false>
__global__ void __launch_bounds__(THREADS, MINB) local_count_kernel(const LocalCountParams p) {
    using SM = LocalCountSmem<WORDS, THREADS, ITEMS, CBITS>;
    using KT = typename KeyImage<WORDS>::T;
    constexpr int WARPS = THREADS / 32;
    static_assert(ITEMS % 8 == 0, "scatter batches of eight");
    extern __shared__ __align__(128) uint8_t smem_raw[];
    uint32_t* s_out = reinterpret_cast<uint32_t*>(smem_raw);
    uint32_t* s_cnt = reinterpret_cast<uint32_t*>(smem_raw + SM::OUT_BYTES);
    uint32_t* s_runs = reinterpret_cast<uint32_t*>(smem_raw + SM::OUT_BYTES + SM::CNT_BYTES);  // {count:16 | first position:16}
    uint8_t* misc = smem_raw + SM::OUT_BYTES + SM::CNT_BYTES + SM::RUN_BYTES;
    // [0] first id, [1] next id; per trip parity p: [4p + 2] long run seen, [4p + 3] slow
    // (a trip's flags are cleared two trips later, so no thread can still be reading them)
    volatile uint32_t* s_seg = reinterpret_cast<volatile uint32_t*>(misc);
    uint32_t par = 0;
    uint32_t* s_wtot = reinterpret_cast<uint32_t*>(misc + 64);  // WARPS warp totals
    unsigned long long* s_vary = reinterpret_cast<unsigned long long*>(misc + 256);  // auto_top: per-warp OR of (key ^ first key)
    static_assert(WARPS <= 32, "s_wtot / s_vary hold 32 warps");

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t cnt_base = smem_u32(s_cnt);
    const KT xorv = KT(p.key.xorv);
    for (int i = tid; i < int(SM::CNT_BYTES / 16); i += THREADS) reinterpret_cast<uint4*>(s_cnt)[i] = make_uint4(0, 0, 0, 0);

    // Global memory is touched two records at a time (16 bytes for u64 keys): half the load / store
    // instructions, which is what the memory phases of this kernel wait on (lg_throttle).  A segment starts
    // at any record, so: `head` records (0 or 1) up to pair alignment, npairs aligned pairs, maybe one tail
    // record.  The order of records inside a segment is free, so thread t simply takes pairs t, t + THREADS, ...;
    // head and tail go to the last slots of the last thread (free whenever they exist, as cnt <= CAP).
    uint32_t cnt = 0, head = 0, vmask = 0;  // vmask: bit j = r[j] holds a record
    const uint32_t* g = nullptr;
    uint64_t gbase = 0;
    Rec<WORDS> r[ITEMS];
    using Vec2 = typename std::conditional<WORDS == 1, uint2, uint4>::type;
    auto fetch = [&](uint32_t sg) {
        cnt = 0;
        vmask = 0;
        if (sg >= p.nseg) return;
        uint64_t base;
        if (p.seg_pairs) {
            const uint2 pr = p.seg_pairs[sg];
            base = pr.x, cnt = pr.y;
        } else {
            base = p.seg_off[sg];
            cnt = uint32_t(p.seg_off[sg + 1] - base);
        }
        if (cnt > uint32_t(SM::CAP)) cnt = 0;  // oversized segments are somebody else's (heavy.cuh): skip
        gbase = base;
        g = (p.src ? p.src : p.data) + base * WORDS;
        head = uint32_t(base) & 1u;
        if (cnt <= 1) {
            // a single record still has to travel when the segments are read from another buffer
            if (cnt == 1 && p.src && tid == 0) store_rec<WORDS>(p.data, base, load_rec<WORDS>(g, 0));
            return;
        }
        const uint32_t npairs = (cnt - head) >> 1;
        const Vec2* gp = reinterpret_cast<const Vec2*>(g + head * WORDS);
#pragma unroll
        for (int k0 = 0; k0 < ITEMS / 2; k0 += 4) {
            if (uint32_t(k0 * THREADS) < npairs) {  // block-uniform: small segments skip the empty batches
#pragma unroll
                for (int k = k0; k < k0 + 4; ++k) {
                    const uint32_t v = uint32_t(k * THREADS + tid);
                    if (v < npairs) {
                        const Vec2 x = gp[v];
                        if constexpr (WORDS == 1) {
                            r[2 * k].w[0] = x.x, r[2 * k + 1].w[0] = x.y;
                        } else {
                            r[2 * k].w[0] = x.x, r[2 * k].w[1] = x.y, r[2 * k + 1].w[0] = x.z, r[2 * k + 1].w[1] = x.w;
                        }
                        vmask |= 3u << (2 * k);
                    }
                }
            }
        }
        if (tid == THREADS - 1) {
            if (head) r[ITEMS - 2] = load_rec<WORDS>(g, 0), vmask |= 1u << (ITEMS - 2);
            if ((cnt - head) & 1u) r[ITEMS - 1] = load_rec<WORDS>(g, cnt - 1), vmask |= 1u << (ITEMS - 1);
        }
    };

    long long t_last = p.prof ? clock64() : 0;
    auto tick = [&](int k) {
        if (p.prof && tid == 0) {
            const long long t = clock64();
            atomicAdd(p.prof + k, (unsigned long long)(t - t_last));
            t_last = t;
        }
    };
    if (tid == 0) s_seg[0] = p.seg_first + atomicAdd(p.ticket, 1u);
    __syncthreads();
    uint32_t seg = s_seg[0];
    fetch(seg);
    while (seg < p.nseg) {
        par ^= 1u;
        volatile uint32_t* flag = s_seg + 4 * par;  // [2] long run seen, [3] slow
        if (AUTOTOP && cnt > 1) {
            // bits in which this segment's keys differ: OR of (key ^ the segment's first key), per warp; combined after
            // barrier (0) (the slots are rewritten one trip later, behind barriers (1)..(5))
            const KT kref = key_image<WORDS, FLOATK>(load_rec<WORDS>(g, 0), xorv);
            KT v = 0;
#pragma unroll
            for (int j = 0; j < ITEMS; ++j)
                if (vmask >> j & 1u) v |= key_image<WORDS, FLOATK>(r[j], xorv) ^ kref;
            uint32_t lo = __reduce_or_sync(0xffffffffu, uint32_t(v)), hi = 0;
            if constexpr (WORDS == 2) hi = __reduce_or_sync(0xffffffffu, uint32_t(uint64_t(v) >> 32));
            if (lane == 0) s_vary[warp] = (unsigned long long)lo | ((unsigned long long)hi << 32);
        }
        if (tid == 0) {
            s_seg[1] = p.seg_first + atomicAdd(p.ticket, 1u);
            flag[2] = 0;
            flag[3] = 0;
        }
        __syncthreads();  // (0) previous segment done, counters clean, next id visible
        tick(0);
        const uint32_t next_seg = s_seg[1];
        if (cnt <= 1) {
            seg = next_seg;
            fetch(seg);
            __syncthreads();  // everyone has read the next id before thread 0 claims another one
            continue;
        }
        const uint32_t cur_cnt = cnt, cur_seg = seg, cur_head = head, mine = vmask;
        uint32_t* const cur_g = p.data + gbase * WORDS;
        // staging positions are shifted by `head`, so that the aligned pairs of the write-out are aligned here too
        uint32_t* const st = s_out + cur_head * WORDS;
        auto key_at = [&](uint32_t i) { return key_image<WORDS, FLOATK>(load_rec<WORDS>(st, i), xorv); };
        // digit width of this segment: 2..4 counters per record, within [1, min(bmax, top_bit)]
        uint32_t b = 33u - uint32_t(__clz(int(cur_cnt - 1)));  // ceil(log2 cnt) + 1 for cnt >= 2
        b = b < p.bmax ? b : p.bmax;
        uint32_t top_bit = p.top_bit;
        if constexpr (AUTOTOP) {
            unsigned long long vary = 0;
#pragma unroll
            for (int w = 0; w < WARPS; ++w) vary |= s_vary[w];
            const uint32_t bits = 64u - uint32_t(__clzll((long long)vary));  // 0 when every key is the same
            top_bit = min(top_bit, max(bits, 1u));
        }
        b = b < top_bit ? b : top_bit;
        const uint32_t shift = top_bit - b;
        const bool exact = shift == 0;  // the digit is everything that is left of the key: no fix-up
        const uint32_t dmask = (1u << b) - 1u;
        const uint32_t NW = b ? (1u << (b - 1)) : 1u;  // packed counter words
        auto digit_of = [&](const Rec<WORDS>& x) { return uint32_t(key_image<WORDS, FLOATK>(x, xorv) >> shift) & dmask; };

        // ---- count ----
        // (batches of 8 records = 4 pair slots; a batch no thread has records in is skipped: block-uniform test)
        const uint32_t cur_pairs = (cur_cnt - cur_head) >> 1;
        const bool leftovers = cur_head || ((cur_cnt - cur_head) & 1u);
        auto batch_live = [&](int j0) { return uint32_t(j0 / 2 * THREADS) < cur_pairs || (j0 == ITEMS - 8 && leftovers); };
#pragma unroll
        for (int j0 = 0; j0 < ITEMS; j0 += 8) {
            if (batch_live(j0)) {
#pragma unroll
                for (int j = j0; j < j0 + 8; ++j) {
                    if (mine >> j & 1u) {
                        const uint32_t d = digit_of(r[j]);
                        red_add_shared(cnt_base + (d >> 1) * 4, 1u << ((d & 1u) * 16));
                    }
                }
            }
        }
        __syncthreads();  // (1)
        tick(1);
        // ---- scan: counters := exclusive bases; counters holding >= 2 records go to the run list ----
        // The run list is private to the warp: a warp whose counters hold `t` records after `r` records of the
        // warps before it lists at most t / 2 runs, in entries [r / 2, (r + t) / 2) — no shared counter.
        // warp-collective: reserve `nr` consecutive entries for this lane, returns the first
        uint32_t my_runs = 0, runs_at = 0;  // warp-uniform: entries written so far, first entry of my region
        auto reserve = [&](uint32_t nr) -> uint32_t {
            uint32_t incl = nr;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += y;
            }
            const uint32_t first = runs_at + my_runs + incl - nr;
            my_runs += __shfl_sync(0xffffffffu, incl, 31);
            return first;
        };
        const bool vec = NW >= uint32_t(WARPS) * 128u;  // block-uniform
        if (vec) {
            // every warp owns a contiguous run of words and walks it 128 words (one uint4 per lane) at a time
            const uint32_t per_warp = NW / WARPS;
            uint4* base4 = reinterpret_cast<uint4*>(s_cnt + uint32_t(warp) * per_warp);
            const uint32_t rounds = per_warp / 128;
            uint32_t acc = 0;  // packed halves: per-lane sums stay below 2^16 (a segment has <= 16 384 records)
            for (uint32_t q = 0; q < rounds; ++q) {
                const uint4 v = base4[q * 32 + lane];
                acc += v.x + v.y + v.z + v.w;
            }
            uint32_t tot = (acc & 0xffffu) + (acc >> 16);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
            if (lane == 0) s_wtot[warp] = tot;
            __syncthreads();  // (2)
            uint32_t run = (lane < warp && lane < WARPS) ? s_wtot[lane] : 0u;  // records before this warp's words
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) run += __shfl_xor_sync(0xffffffffu, run, o);
            runs_at = run >> 1;
            for (uint32_t q = 0; q < rounds; ++q) {
                uint4 v = base4[q * 32 + lane];
                // word {c1:c0} -> {at + c0 : at}; counters with c >= 2 -> run list entries (static indices: registers)
                uint32_t c[8], a8[9];
                c[0] = v.x & 0xffffu, c[1] = v.x >> 16, c[2] = v.y & 0xffffu, c[3] = v.y >> 16;
                c[4] = v.z & 0xffffu, c[5] = v.z >> 16, c[6] = v.w & 0xffffu, c[7] = v.w >> 16;
                uint32_t nr = 0;
                if (!exact) {
#pragma unroll
                    for (int k = 0; k < 8; ++k) nr += c[k] >= 2 ? 1u : 0u;
                }
                // one warp scan for both: {runs of this lane : records of this lane}
                const uint32_t packed = v.x + v.y + v.z + v.w;
                const uint32_t both = ((packed & 0xffffu) + (packed >> 16)) | (nr << 16);
                uint32_t incl = both;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
                    if (lane >= o) incl += y;
                }
                const uint32_t excl = incl - both, tot = __shfl_sync(0xffffffffu, incl, 31);
                a8[0] = run + (excl & 0xffffu);
#pragma unroll
                for (int k = 0; k < 8; ++k) a8[k + 1] = a8[k] + c[k];
                v.x = a8[0] | (a8[1] << 16), v.y = a8[2] | (a8[3] << 16), v.z = a8[4] | (a8[5] << 16), v.w = a8[6] | (a8[7] << 16);
                base4[q * 32 + lane] = v;
                if (!exact) {
                    uint32_t slot = runs_at + my_runs + (excl >> 16);
#pragma unroll
                    for (int k = 0; k < 8; ++k)
                        if (c[k] >= 2) s_runs[slot++] = a8[k] | (c[k] << 16);
                    my_runs += tot >> 16;
                }
                run += tot & 0xffffu;
            }
        } else {
            // small tables: thread t owns words [t * W, (t + 1) * W), walked from word (t mod W) on and around
            const uint32_t W = NW >= uint32_t(THREADS) ? NW / THREADS : 1u;
            const bool scanner = uint32_t(tid) < NW;
            const uint32_t w0 = uint32_t(tid) * W, rot = uint32_t(tid) & (W - 1);
            uint32_t total = 0, wrapped = 0;  // wrapped = sum of the words before `rot`
            if (scanner) {
                for (uint32_t k = 0; k < W; ++k) {
                    const uint32_t idx = (rot + k) & (W - 1);
                    const uint32_t v = s_cnt[w0 + idx];
                    const uint32_t lo = v & 0xffffu, hi = v >> 16;
                    total += lo + hi;
                    if (idx < rot) wrapped += lo + hi;
                }
            }
            uint32_t incl = total;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += y;
            }
            if (lane == 31) s_wtot[warp] = incl;
            __syncthreads();  // (2)
            uint32_t excl = incl - total, wbase = 0;
#pragma unroll
            for (int w = 0; w < WARPS; ++w)
                if (w < warp) wbase += s_wtot[w];
            excl += wbase;
            runs_at = wbase >> 1;
            uint32_t run = excl + wrapped;
            for (uint32_t k = 0; k < W; ++k) {  // W is block-uniform: the appends below are warp-collective
                const uint32_t idx = (rot + k) & (W - 1);
                if (k && idx == 0) run = excl;  // wrapped around: back at the first word of my run
                uint32_t lo = 0, hi = 0, at = 0;
                if (scanner) {
                    const uint32_t v = s_cnt[w0 + idx];
                    lo = v & 0xffffu, hi = v >> 16, at = run;
                    s_cnt[w0 + idx] = run | ((run + lo) << 16);
                    run += lo + hi;
                }
                if (!exact) {
                    uint32_t slot = reserve((lo >= 2 ? 1u : 0u) + (hi >= 2 ? 1u : 0u));
                    if (lo >= 2) s_runs[slot++] = at | (lo << 16);
                    if (hi >= 2) s_runs[slot] = (at + lo) | (hi << 16);
                }
            }
        }
        __syncthreads();  // (3)
        tick(2);
        // ---- scatter: the atomic hands out positions (in batches of 8: eight atomics in flight per thread
        // before the first dependent store) ----
#pragma unroll
        for (int j0 = 0; j0 < ITEMS; j0 += 8) {
            if (!batch_live(j0)) continue;
            uint32_t pos[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int j = j0 + u;
                pos[u] = 0;
                if (mine >> j & 1u) {
                    const uint32_t d = digit_of(r[j]);
                    const uint32_t sh = (d & 1u) * 16;
                    pos[u] = (atom_add_shared(cnt_base + (d >> 1) * 4, 1u << sh) >> sh) & 0xffffu;
                }
            }
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int j = j0 + u;
                if (mine >> j & 1u) store_rec<WORDS>(st, pos[u], r[j]);
            }
        }
        __syncthreads();  // (4) the staging buffer is sorted by the digit; the run list is complete
        tick(3);
        // ---- prefetch the next segment ----
        seg = next_seg;
        fetch(seg);
        tick(4);
        // ---- fix-up: one thread per run of records that share a digit, ordered by the whole key ----
        // (every warp works through its own list)
        const uint32_t nruns = exact ? 0u : my_runs;
        for (uint32_t q = lane; q < nruns; q += 32) {
            const uint32_t e = s_runs[runs_at + q], at = e & 0xffffu, c = e >> 16;
            if (c == 2) {
                const Rec<WORDS> x = load_rec<WORDS>(st, at), y = load_rec<WORDS>(st, at + 1);
                if (key_image<WORDS, FLOATK>(y, xorv) < key_image<WORDS, FLOATK>(x, xorv)) {
                    store_rec<WORDS>(st, at, y);
                    store_rec<WORDS>(st, at + 1, x);
                }
            } else if (c <= uint32_t(LC_MAXRUN)) {
                for (uint32_t a = at + 1; a < at + c; ++a) {  // insertion sort
                    const Rec<WORDS> x = load_rec<WORDS>(st, a);
                    const KT kx = key_image<WORDS, FLOATK>(x, xorv);
                    uint32_t j = a;
                    while (j > at) {
                        const Rec<WORDS> y = load_rec<WORDS>(st, j - 1);
                        if (key_image<WORDS, FLOATK>(y, xorv) <= kx) break;
                        store_rec<WORDS>(st, j, y);
                        --j;
                    }
                    if (j != a) store_rec<WORDS>(st, j, x);
                }
            } else {
                flag[2] = 1;  // a long run: checked below
            }
        }
        __syncthreads();  // (5)
        tick(5);
        if (flag[2]) {
            // Runs longer than LC_MAXRUN are heavy duplicates: one key many times over, plus at most a few other keys
            // that share its digit.  A warp reads the run once, keeps the records that differ from the dominant key
            // (the one in the middle of the run) in its lanes — up to 32 of them —, ranks those among themselves and
            // rewrites the run as [smaller keys, sorted][the dominant key, copied][larger keys, sorted]: equal keys
            // are identical records, so copies are as good as the originals.  More than 32 other keys in one run:
            // the stable segment sort takes the segment.
            for (uint32_t q0 = 0; q0 < nruns; q0 += 32) {
                const uint32_t e = q0 + lane < nruns ? s_runs[runs_at + q0 + lane] : 0u;
                uint32_t lm = __ballot_sync(0xffffffffu, (e >> 16) > uint32_t(LC_MAXRUN));
                while (lm) {
                    const int srcl = __ffs(lm) - 1;
                    lm &= lm - 1;
                    const uint32_t ee = __shfl_sync(0xffffffffu, e, srcl), at = ee & 0xffffu, c = ee >> 16;
                    const Rec<WORDS> pv = load_rec<WORDS>(st, at + c / 2);
                    const KT kpv = key_image<WORDS, FLOATK>(pv, xorv);
                    uint32_t m = 0;  // other keys found so far (warp-uniform); lane l < m holds the l-th of them
                    Rec<WORDS> other;
#pragma unroll
                    for (int k = 0; k < WORDS; ++k) other.w[k] = 0;
                    bool overflow = false;
                    for (uint32_t i0 = at; i0 < at + c && !overflow; i0 += 32) {
                        const uint32_t i = i0 + lane;
                        const bool have = i < at + c;
                        Rec<WORDS> x = pv;
                        if (have) x = load_rec<WORDS>(st, i);
                        const uint32_t nm = __ballot_sync(0xffffffffu, have && key_image<WORDS, FLOATK>(x, xorv) != kpv);
                        const uint32_t found = __popc(nm);
                        if (m + found > 32) {
                            overflow = true;
                        } else if (found) {
                            // lane l in [m, m + found) takes the (l - m + 1)-th record flagged in nm
                            const uint32_t from = __fns(nm, 0, int(uint32_t(lane) - m) + 1);
                            Rec<WORDS> y;
#pragma unroll
                            for (int k = 0; k < WORDS; ++k) y.w[k] = __shfl_sync(0xffffffffu, x.w[k], from & 31u);
                            if (uint32_t(lane) >= m && uint32_t(lane) < m + found) other = y;
                            m += found;
                        }
                    }
                    if (overflow) {
                        if (lane == 0) flag[3] = 1;
                    } else if (m) {
                        const KT ko = key_image<WORDS, FLOATK>(other, xorv);
                        const bool mine_ok = uint32_t(lane) < m;
                        uint32_t rank = 0;  // among the other keys
                        for (uint32_t i = 0; i < m; ++i) {
                            const KT ky = KT(__shfl_sync(0xffffffffu, ko, int(i)));
                            rank += (ky < ko || (ky == ko && i < uint32_t(lane))) ? 1u : 0u;
                        }
                        const uint32_t nlt = __popc(__ballot_sync(0xffffffffu, mine_ok && ko < kpv)), neq = c - m;
                        __syncwarp();
                        for (uint32_t i = at + nlt + lane; i < at + nlt + neq; i += 32) store_rec<WORDS>(st, i, pv);
                        if (mine_ok) store_rec<WORDS>(st, at + rank + (ko > kpv ? neq : 0u), other);
                    }
                }
            }
            __syncthreads();
        }
        // ---- write-out ----
        // a segment for the slow list is written out all the same (as it stands: a permutation of its records), so
        // that the stable segment sort finds it in `data` whatever buffer it was read from
        if (flag[3] && tid == 0) {
            const uint32_t at = atomicAdd(p.slow_list, 1u);
            p.slow_list[1 + at] = cur_seg;
        }
        {
            // aligned pairs: one vector load from the staging buffer, one vector store
            const uint32_t npairs = (cur_cnt - cur_head) >> 1;
            Vec2* gp = reinterpret_cast<Vec2*>(cur_g + cur_head * WORDS);
            const Vec2* sp = reinterpret_cast<const Vec2*>(s_out + 2 * cur_head * WORDS);  // = st + head: record `head`
            for (uint32_t v = tid; v < npairs; v += THREADS) gp[v] = sp[v];
            if (tid == 0 && cur_head) store_rec<WORDS>(cur_g, 0, load_rec<WORDS>(st, 0));
            if (tid == 32 % THREADS && ((cur_cnt - cur_head) & 1u)) store_rec<WORDS>(cur_g, uint64_t(cur_cnt - 1), load_rec<WORDS>(st, cur_cnt - 1));
        }
        tick(6);
        // the counters are dead since barrier (4); barrier (0) of the next trip orders this against the next count
        for (uint32_t i = tid; i < NW / 4; i += THREADS) reinterpret_cast<uint4*>(s_cnt)[i] = make_uint4(0, 0, 0, 0);
        if (NW < 4 && uint32_t(tid) < NW) s_cnt[tid] = 0;
        tick(7);
    }
}

}  // namespace b200
