// local_sort.cuh — K3, the shared-memory segment sort of the hybrid path.
//
// rdst's default route is MSB-first: split on the top byte, recurse into the 256 buckets, and
// finish small buckets with an LSB sort that runs in cache (reference src/sorter.rs:24-171,
// src/tuners/standard_tuner.rs:13-63, src/sorts/lsb_sort.rs:38-128).  The GPU analogue: after
// two global MSD passes (K2 on the two most significant active levels) the array is 65 536
// segments that each fit in one SM's shared memory (<= 16 384 records); one CTA then sorts a
// segment completely on chip — read once, written once — instead of six more passes through HBM.
//
// Per segment: records live in registers (32 per thread, 512 threads); every local level is one
// stable counting-sort step exactly like K2's tile step (half-warp OR-matching ranks, digit scan,
// scatter into the staging buffer) followed by a read-back.  Fast path: only the top `nfast` local
// levels are radix-sorted; what is left are short runs of records that agree on those levels, and
// the thread at the head of each run insertion-sorts it by the whole key (stable).  Runs longer
// than MAXRUN (duplicate-heavy or adversarial data) switch the segment to the slow path: a full
// LSD over all local levels.  Both paths give the unique stable order, so the result is
// bit-identical to the pure LSD engine.
#pragma once
#include "pass_or.cuh"

namespace b200 {

constexpr int LS_THREADS = 512, LS_ITEMS = 32, LS_CAP = LS_THREADS * LS_ITEMS;  // 16 384 records
constexpr int LS_GROUPS = LS_THREADS / 16, LS_MAXRUN = 24, LS_MAXLEV = 8;

struct LocalSortParams {
    uint32_t* data;            // sorted in place, segment by segment
    const uint64_t* seg_off;   // [nseg + 1] record offsets of the segments
    uint32_t nseg;
    uint32_t* ticket;          // zeroed before the launch
    uint32_t nlev;             // local radix levels, least significant first
    uint32_t nfast;            // fast path sorts only the top nfast of them (0 or nlev: always all)
    uint32_t fix_shift;        // bits of the key image below the lowest fast level
    DigitParams lv[LS_MAXLEV];
    SplitParams key;           // 64-bit order-preserving image of the whole key (LEVELS <= 8)
    uint32_t* slow_segments;   // telemetry: segments that fell back to the full LSD
    unsigned long long* prof;  // optional [8] cycle counters per phase (thread 0 of every CTA), nullptr = off
    const uint2* seg_pairs;    // alternative to seg_off: [nseg] {first record, records} (segments need not be adjacent)
    const uint32_t* list;      // list mode (nullptr = off): sort segments list[1 .. list[0]] only — the ones K3-U
                               // (local_count.cuh) left behind; the count is read on the device, no host round trip
};

template <int WORDS>
struct LocalSmem {
    static constexpr size_t OUT_BYTES = size_t(LS_CAP) * WORDS * 4;
    static constexpr size_t SLOT_BYTES = size_t(LS_GROUPS) * RADIX * 4;
    static constexpr size_t PART_BYTES = size_t(LS_THREADS) * 4;
    static constexpr size_t MISC_BYTES = 128;
    static constexpr size_t TOTAL = OUT_BYTES + SLOT_BYTES + PART_BYTES + MISC_BYTES;
};

template <int WORDS, bool FLOATK>
__global__ void __launch_bounds__(LS_THREADS, 1) local_sort_kernel(const LocalSortParams p) {
    using SM = LocalSmem<WORDS>;
    constexpr int ITEMS = LS_ITEMS, THREADS = LS_THREADS, GROUPS = LS_GROUPS;
    constexpr int GPH = GROUPS / 2;  // two threads per digit in the digit phase
    extern __shared__ __align__(128) uint8_t smem_raw[];
    uint32_t* s_out = reinterpret_cast<uint32_t*>(smem_raw);
    uint8_t* s_slot_raw = smem_raw + SM::OUT_BYTES;
    uint32_t* s_part = reinterpret_cast<uint32_t*>(smem_raw + SM::OUT_BYTES + SM::SLOT_BYTES);
    uint8_t* misc = smem_raw + SM::OUT_BYTES + SM::SLOT_BYTES + SM::PART_BYTES;
    volatile uint32_t* s_seg = reinterpret_cast<volatile uint32_t*>(misc);       // [0] segment id, [1] slow-path flag, [2] run heads found, [3] next segment id
    uint32_t* s_wtot = reinterpret_cast<uint32_t*>(misc + 32);                   // 8 digit-warp totals

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t lane_bit = 1u << (lane & 15);
    const uint32_t lt_mask = lane_bit - 1u;
    const uint32_t slot_base = smem_u32(s_slot_raw);
    const uint32_t my_slots = slot_base + uint32_t(tid >> 4) * RADIX * 4;
    const uint32_t dg = tid & (RADIX - 1), h = tid / RADIX;

    {
        uint4* z = reinterpret_cast<uint4*>(s_slot_raw);
        for (int i = tid; i < int(SM::SLOT_BYTES / 16); i += THREADS) z[i] = make_uint4(0, 0, 0, 0);
    }

    // Segments are claimed one ahead, and the next segment's records are fetched into the registers as
    // soon as the current one sits sorted in the staging buffer, so the loads fly during the write-out.
    uint32_t cnt = 0, IT = 0, e0 = 0;
    uint32_t* g = nullptr;
    Rec<WORDS> r[ITEMS];
    const uint32_t nseg = p.list ? min(__ldg(p.list), p.nseg) : p.nseg;
    auto fetch = [&](uint32_t sg) {  // header + records of segment sg (nothing for cnt <= 1)
        cnt = 0;
        if (sg >= nseg) return;
        if (p.list) sg = __ldg(p.list + 1 + sg);
        uint64_t base;
        if (p.seg_pairs) {
            const uint2 pr = p.seg_pairs[sg];
            base = pr.x, cnt = pr.y;
        } else {
            base = p.seg_off[sg];
            cnt = uint32_t(p.seg_off[sg + 1] - base);
        }
        g = p.data + base * WORDS;
        // records per thread actually used: element (group, j, lane16) = group*16*IT + j*16 + lane16,
        // so that every half-warp owns a contiguous run and (group, j, lane) order is memory order
        IT = (cnt + THREADS - 1) / THREADS;
        e0 = uint32_t(tid >> 4) * 16u * IT + (lane & 15);
        if (cnt <= 1) return;
#pragma unroll
        for (int j = 0; j < ITEMS; ++j) {
            const uint32_t e = e0 + j * 16;
            if (uint32_t(j) < IT && e < cnt) {
                r[j] = load_rec<WORDS>(g, e);
            } else {
#pragma unroll
                for (int k = 0; k < WORDS; ++k) r[j].w[k] = 0;
            }
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
    if (tid == 0) s_seg[0] = atomicAdd(p.ticket, 1u);
    __syncthreads();
    uint32_t seg = s_seg[0];
    fetch(seg);
    while (seg < nseg) {
        if (tid == 0) {
            s_seg[3] = atomicAdd(p.ticket, 1u);  // the segment after this one
            s_seg[1] = 0;
            s_seg[2] = 0;
        }
        __syncthreads();  // previous segment's write-out done; slots clean; next id visible
        tick(0);
        const uint32_t next_seg = s_seg[3];
        if (cnt <= 1) {  // nothing to sort
            seg = next_seg;
            fetch(seg);
            __syncthreads();  // everyone has read the next id before thread 0 claims another one
            continue;
        }
        const uint32_t cur_cnt = cnt;
        uint32_t* const cur_g = g;

        bool fast = p.nfast > 0 && p.nfast < p.nlev;
        uint32_t li = fast ? p.nlev - p.nfast : 0;
        while (true) {  // one trip on the fast path; a second, over all levels, if a run was too long
            for (; li < p.nlev; ++li) {
                const DigitParams dp = p.lv[li];
                // ---- rank inside the half-warp by OR-matching (as K2) ----
                uint32_t rk[ITEMS / 2];
#pragma unroll
                for (int q = 0; q < ITEMS / 2; ++q) rk[q] = 0;
#pragma unroll
                for (int j = 0; j < ITEMS; ++j) {
                    if (uint32_t(j) < IT) {
                        const uint32_t e = e0 + j * 16;
                        uint32_t d = 255u;  // padding sorts behind every record
                        if (e < cnt) d = rec_digit_w<WORDS, FLOATK, true>(dp, r[j]);
                        const uint32_t sa = my_slots + d * 4;
                        slot_or(sa, lane_bit);
                        __syncwarp();
                        const uint32_t v = lds_u32(sa);
                        __syncwarp();
                        const uint32_t m = v & 0xffffu, c = v >> 16;
                        const uint32_t lt = m & lt_mask;
                        if (lt == 0) sts_u32(sa, (c + __popc(m)) << 16);
                        __syncwarp();
                        rk[j / 2] |= (c + __popc(lt)) << ((j & 1) * 16);
                    }
                }
                __syncthreads();
                tick(1);
                // ---- two threads per digit: counts of my 16 groups -> positions in the sorted segment ----
                uint32_t gc[GPH], part = 0, sum = 0, before = 0, incl = 0;
                {
                    const uint32_t ca = slot_base + (h * GPH * RADIX + dg) * 4;
#pragma unroll
                    for (int w = 0; w < GPH; ++w) {
                        gc[w] = lds_u32(ca + w * RADIX * 4) >> 16;
                        part += gc[w];
                    }
                    s_part[tid] = part;
                }
                __syncthreads();
                {
                    const uint32_t x0 = s_part[dg], x1 = s_part[RADIX + dg];
                    sum = x0 + x1;
                    before = h ? x0 : 0u;
                    incl = sum;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) {
                        const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
                        if (lane >= o) incl += y;
                    }
                    if (h == 0 && lane == 31) s_wtot[warp] = incl;
                }
                __syncthreads();
                {
                    uint32_t wbase = 0;
#pragma unroll
                    for (int w = 0; w < RADIX / 32; ++w)
                        if (uint32_t(w) < dg / 32) wbase += s_wtot[w];
                    const uint32_t ca = slot_base + (h * GPH * RADIX + dg) * 4;
                    uint32_t run = wbase + incl - sum + before;
#pragma unroll
                    for (int w = 0; w < GPH; ++w) {
                        sts_u32(ca + w * RADIX * 4, run << 16);
                        run += gc[w];
                    }
                }
                __syncthreads();
                tick(2);
                // ---- scatter into the staging buffer in sorted order ----
#pragma unroll
                for (int j = 0; j < ITEMS; ++j) {
                    if (uint32_t(j) < IT) {
                        const uint32_t e = e0 + j * 16;
                        uint32_t d = 255u;
                        if (e < cnt) d = rec_digit_w<WORDS, FLOATK, true>(dp, r[j]);
                        const uint32_t pos = (lds_u32(my_slots + d * 4) >> 16) + ((rk[j / 2] >> ((j & 1) * 16)) & 0xffffu);
                        store_rec<WORDS>(s_out, pos, r[j]);
                    }
                }
                __syncwarp();
                {
                    uint4* z = reinterpret_cast<uint4*>(s_slot_raw + size_t(warp) * 2 * RADIX * 4);
#pragma unroll
                    for (int i = lane; i < 2 * RADIX * 4 / 16; i += 32) z[i] = make_uint4(0, 0, 0, 0);
                }
                __syncthreads();
                tick(3);
                if (li + 1 < p.nlev) {
                    // read back in element order for the next level
#pragma unroll
                    for (int j = 0; j < ITEMS; ++j) {
                        const uint32_t e = e0 + j * 16;
                        if (uint32_t(j) < IT && e < cnt) r[j] = load_rec<WORDS>(s_out, e);
                    }
                    __syncthreads();  // everyone has its records before the next scatter overwrites them
                    tick(4);
                }
            }
            if (!fast) break;
            // ---- fix-up: runs of records equal on the fast levels, ordered by the rest of the key ----
            // Phase 1: find the run heads (record i differs from i-1 and equals i+1 on the fast levels).
            // Every warp scans one contiguous chunk, so eq(i) = "record i equals record i + 1 on the fast
            // levels" comes from two conflict-free loads and eq(i - 1) from the warp's ballot (bit 31 of
            // the previous round carries over).  The heads go into the warp's own list (it borrows the slot
            // tables, idle and zero here; no shared counter: same-address shared atomics with a return
            // value serialise at ~36 cycles each and made this phase a third of the kernel).
            // Phase 2: the same warp, one run per thread: find its end, insertion-sort it by the whole key
            // image (stable).  A run may reach into the next warp's chunk while that warp is still
            // scanning: its records all agree on the fast levels, so the scan reads the same eq() either way.
            // "equal on the fast levels" on the raw record words: the key image is the raw key xor a
            // constant (or, for floats, a constant chosen by the sign bit, which is above the fast levels),
            // so two records agree on image bits >= fix_shift iff their raw words agree under these masks
            uint32_t eqm_w[WORDS];
            {
                const uint64_t m = (p.key.mask >> p.fix_shift) << p.fix_shift;
#pragma unroll
                for (int k = 0; k < WORDS; ++k)
                    eqm_w[k] = uint32_t(k) == p.key.lo_word ? uint32_t(m) : (uint32_t(k) == p.key.hi_word ? uint32_t(m >> 32) : 0u);
            }
            auto same_fast = [&](const Rec<WORDS>& a, const Rec<WORDS>& b) {
                uint32_t diff = 0;
#pragma unroll
                for (int k = 0; k < WORDS; ++k) diff |= (a.w[k] ^ b.w[k]) & eqm_w[k];
                return diff == 0;
            };
            constexpr uint32_t HEADS_PER_WARP = LS_CAP / (THREADS / 32) / 2;  // a chunk of c records has <= c / 2 heads
            uint16_t* my_heads = reinterpret_cast<uint16_t*>(s_slot_raw) + uint32_t(warp) * HEADS_PER_WARP;
            uint32_t nheads = 0;  // warp-uniform
            {
                auto same_at = [&](uint32_t i) { return same_fast(load_rec<WORDS>(s_out, i), load_rec<WORDS>(s_out, i + 1)); };
                const uint32_t per = ((cnt + THREADS - 1) / THREADS) * 32;
                const uint32_t wlo = uint32_t(warp) * per, whi = wlo + per < cnt ? wlo + per : cnt;
                uint32_t carry = 0;
                if (wlo > 0 && wlo < cnt) carry = same_at(wlo - 1) ? 1u : 0u;
                for (uint32_t i0 = wlo; i0 < whi; i0 += 128) {
                    // four rounds at a time, loads first: the list stores below order against shared loads
                    uint32_t eqm[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const uint32_t i = i0 + u * 32 + lane;
                        bool eq = false;
                        if (i + 1 < cnt && i < whi) eq = same_at(i);
                        eqm[u] = __ballot_sync(0xffffffffu, eq);
                    }
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const uint32_t headm = eqm[u] & ~((eqm[u] << 1) | carry);
                        carry = eqm[u] >> 31;
                        if ((headm >> lane) & 1u)
                            my_heads[nheads + __popc(headm & lanemask_lt())] = uint16_t(i0 + u * 32 + lane);
                        nheads += __popc(headm);
                    }
                }
            }
            __syncwarp();
            tick(5);
            // nine runs in ten are two records long: three independent loads and at most one swap, all lanes
            // together; the heads of longer runs are compacted to the front of the list for a second sweep
            uint32_t nlong = 0;  // warp-uniform
            for (uint32_t hbase = 0; hbase < nheads; hbase += 32) {
                const bool act = hbase + lane < nheads;
                const uint32_t i = act ? my_heads[hbase + lane] : 0u;
                bool more = false;
                if (act) {
                    const Rec<WORDS> ri = load_rec<WORDS>(s_out, i), r1 = load_rec<WORDS>(s_out, i + 1);
                    more = i + 2 < cnt && same_fast(ri, load_rec<WORDS>(s_out, i + 2));
                    if (!more && rec_prefix<WORDS, FLOATK>(p.key, r1) < rec_prefix<WORDS, FLOATK>(p.key, ri)) {
                        store_rec<WORDS>(s_out, i, r1);
                        store_rec<WORDS>(s_out, i + 1, ri);
                    }
                }
                const uint32_t lm = __ballot_sync(0xffffffffu, more);  // (every lane has read its entry by now)
                if (more) my_heads[nlong + __popc(lm & lanemask_lt())] = uint16_t(i);
                nlong += __popc(lm);
            }
            __syncwarp();
            for (uint32_t hidx = lane; hidx < nlong; hidx += 32) {
                const uint32_t i = my_heads[hidx];
                const Rec<WORDS> ri = load_rec<WORDS>(s_out, i);
                uint32_t q = i + 2;
                while (q < cnt && q - i <= uint32_t(LS_MAXRUN) && same_fast(ri, load_rec<WORDS>(s_out, q))) ++q;
                if (q - i > uint32_t(LS_MAXRUN)) {
                    s_seg[1] = 1;  // too long for an insertion sort: the whole segment takes the slow path
                    continue;
                }
                for (uint32_t a = i + 1; a < q; ++a) {
                    const Rec<WORDS> x = load_rec<WORDS>(s_out, a);
                    const uint64_t kx = rec_prefix<WORDS, FLOATK>(p.key, x);
                    uint32_t b = a;
                    while (b > i) {
                        const Rec<WORDS> y = load_rec<WORDS>(s_out, b - 1);
                        if (rec_prefix<WORDS, FLOATK>(p.key, y) <= kx) break;
                        store_rec<WORDS>(s_out, b, y);
                        --b;
                    }
                    if (b != a) store_rec<WORDS>(s_out, b, x);
                }
            }
            __syncwarp();
            // give the slot tables back clean
            for (uint32_t w = lane; w < (nheads + 1) / 2; w += 32) reinterpret_cast<uint32_t*>(my_heads)[w] = 0;
            __syncthreads();
            tick(6);
            if (s_seg[1] == 0) break;
            // slow path: the staging buffer holds a stable rearrangement of the segment; start over with
            // a full LSD over every local level
            if (tid == 0 && p.slow_segments) atomicAdd(p.slow_segments, 1u);
#pragma unroll
            for (int j = 0; j < ITEMS; ++j) {
                const uint32_t e = e0 + j * 16;
                if (uint32_t(j) < IT && e < cnt) r[j] = load_rec<WORDS>(s_out, e);
            }
            __syncthreads();
            fast = false;
            li = 0;
        }
        // ---- prefetch the next segment, then write-out: the sorted segment goes back where it came from ----
        seg = next_seg;
        fetch(seg);
        for (uint32_t i = tid; i < cur_cnt; i += THREADS)
            store_rec<WORDS>(cur_g, uint64_t(i), load_rec<WORDS>(s_out, i));
        tick(7);
    }
}

}  // namespace b200

namespace b200 {

// Joint histogram of two levels for an array already partitioned (stably) on the first of them:
// joint[d1][d2] = records with digit d1 on level A1 and d2 on level A2, i.e. the sizes of the
// 65 536 segments the second MSD pass produces.  A CTA works on contiguous chunks; a chunk of an
// array sorted by d1 spans few d1 values, so counts go to a small shared table first.
constexpr int H2_THREADS = 256, H2_CHUNK = 65536, H2_SPAN = 4;
struct Hist2Params {
    const uint32_t* src;
    uint64_t n;
    DigitParams l1, l2;
    unsigned long long* joint;  // [256][256], zeroed before the launch
};

// Histograms of two bit windows over the whole array in one read (the two MSD digits when the key's
// varying bits do not end on a byte boundary, so K1's per-level histograms do not describe them).
// Per-warp private tables; a skewed window serialises on its hot bins (correct, slower).
struct HistWinParams {
    const uint32_t* src;
    uint64_t n;
    DigitParams l1, l2;
    unsigned long long* hist;  // [2][256], zeroed before the launch
};

template <int WORDS, bool FLOATK>
__global__ void __launch_bounds__(H2_THREADS) hist_win_kernel(const HistWinParams p) {
    constexpr int WARPS = H2_THREADS / 32;
    __shared__ uint32_t sh[2 * WARPS * RADIX];
    for (int i = threadIdx.x; i < 2 * WARPS * RADIX; i += H2_THREADS) sh[i] = 0;
    __syncthreads();
    uint32_t* mine1 = sh + (threadIdx.x / 32) * RADIX;
    uint32_t* mine2 = mine1 + WARPS * RADIX;
    const uint64_t stride = uint64_t(gridDim.x) * H2_THREADS;
    uint64_t i = uint64_t(blockIdx.x) * H2_THREADS + threadIdx.x;
    for (; i + 3 * stride < p.n; i += 4 * stride) {
        Rec<WORDS> r[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) r[k] = load_rec<WORDS>(p.src + (i + k * stride) * WORDS, 0);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            atomicAdd(&mine1[rec_digit_w<WORDS, FLOATK, true>(p.l1, r[k])], 1u);
            atomicAdd(&mine2[rec_digit_w<WORDS, FLOATK, true>(p.l2, r[k])], 1u);
        }
    }
    for (; i < p.n; i += stride) {
        const Rec<WORDS> r = load_rec<WORDS>(p.src + i * WORDS, 0);
        atomicAdd(&mine1[rec_digit_w<WORDS, FLOATK, true>(p.l1, r)], 1u);
        atomicAdd(&mine2[rec_digit_w<WORDS, FLOATK, true>(p.l2, r)], 1u);
    }
    __syncthreads();
    for (int d = threadIdx.x; d < 2 * RADIX; d += H2_THREADS) {
        const uint32_t* t = sh + (d / RADIX) * WARPS * RADIX + (d % RADIX);
        unsigned long long c = 0;
#pragma unroll
        for (int w = 0; w < WARPS; ++w) c += t[w * RADIX];
        if (c) atomicAdd(p.hist + d, c);
    }
}

template <int WORDS, bool FLOATK>
__global__ void __launch_bounds__(H2_THREADS) hist2_kernel(const Hist2Params p) {
    __shared__ uint32_t sh[H2_SPAN * RADIX];
    const uint64_t chunks = (p.n + H2_CHUNK - 1) / H2_CHUNK;
    for (uint64_t ch = blockIdx.x; ch < chunks; ch += gridDim.x) {
        for (int i = threadIdx.x; i < H2_SPAN * RADIX; i += H2_THREADS) sh[i] = 0;
        const uint64_t lo = ch * H2_CHUNK;
        const uint64_t hi = lo + H2_CHUNK < p.n ? lo + H2_CHUNK : p.n;
        const uint32_t dfirst = rec_digit_w<WORDS, FLOATK, true>(p.l1, load_rec<WORDS>(p.src + lo * WORDS, 0));
        __syncthreads();
        auto count = [&](const Rec<WORDS>& r) {
            const uint32_t d1 = rec_digit_w<WORDS, FLOATK, true>(p.l1, r), d2 = rec_digit_w<WORDS, FLOATK, true>(p.l2, r);
            const uint32_t rel = d1 - dfirst;
            if (rel < uint32_t(H2_SPAN))
                atomicAdd(&sh[rel * RADIX + d2], 1u);
            else
                atomicAdd(p.joint + size_t(d1) * RADIX + d2, 1ull);
        };
        uint64_t i = lo + threadIdx.x;
        for (; i + 3 * H2_THREADS < hi; i += 4 * H2_THREADS) {  // four loads in flight per thread
            Rec<WORDS> r[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) r[k] = load_rec<WORDS>(p.src + (i + k * H2_THREADS) * WORDS, 0);
#pragma unroll
            for (int k = 0; k < 4; ++k) count(r[k]);
        }
        for (; i < hi; i += H2_THREADS) count(load_rec<WORDS>(p.src + i * WORDS, 0));
        __syncthreads();
        for (int i = threadIdx.x; i < H2_SPAN * RADIX; i += H2_THREADS) {
            const uint32_t c = sh[i];
            const uint32_t d1 = dfirst + i / RADIX;
            if (c && d1 < uint32_t(RADIX)) atomicAdd(p.joint + size_t(d1) * RADIX + (i % RADIX), (unsigned long long)c);
        }
        __syncthreads();
    }
}

}  // namespace b200
