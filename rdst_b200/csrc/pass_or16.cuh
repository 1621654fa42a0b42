// pass_or16.cuh — K2 variant "or16": as pass_or.cuh, but the matching group is a HALF-warp.
// A half-warp's slot is one 32-bit word {count:16 | mask:16}, so the three shared-memory
// operations per record (RED.OR, LDS, leader STS) are 32-bit accesses spread over all 32
// banks instead of 64-bit accesses on 16 bank pairs: ncu showed the warp-wide variant
// bound by the L1 data pipe (93 %), 58 % of its shared wavefronts being bank conflicts
// (profiles/r1c_*).  Each half-warp owns a contiguous run of the tile so that
// (half-warp, round, lane) order is memory order and the pass stays stable.
#pragma once
#include "pass_or.cuh"

namespace b200 {

template <int WORDS, int THREADS, int ITEMS, int SUB = 1>
struct Or16Smem {
    static constexpr int WARPS = THREADS / 32;
    static constexpr int TILE = THREADS * ITEMS;
    static constexpr size_t OUT_BYTES = size_t(TILE) * WORDS * 4;
    static constexpr int GROUPS = THREADS / 16 * SUB;  // SUB slot tables per half-warp (rounds split in SUB runs)
    static constexpr size_t SLOT_BYTES = size_t(GROUPS) * RADIX * 4;
    static constexpr size_t GOFS_BYTES = RADIX * 8 * 2 + RADIX * 4;  // u64 offsets, REMAP pointers, u32 offsets
    static constexpr size_t MISC_BYTES = 128;
    static constexpr size_t PART_BYTES = THREADS >= 2 * RADIX ? size_t(THREADS) * 4 : 0;
    static constexpr size_t TOTAL = OUT_BYTES + SLOT_BYTES + GOFS_BYTES + MISC_BYTES + PART_BYTES;
};

template <int WORDS, int THREADS, int ITEMS, bool FLOATK, bool REMAP, int MINB, int SUB = 1, bool HOT = false, bool SEG = false>
__global__ void __launch_bounds__(THREADS, MINB) lsd_pass_or16_kernel(const PassParams p) {
    static_assert(SUB == 1, "one slot table per half-warp");
    constexpr int JH = ITEMS / SUB;  // rounds per sub-table
    static_assert(THREADS >= RADIX && THREADS % 32 == 0, "one thread per digit is needed for the look-back");
    using SM = Or16Smem<WORDS, THREADS, ITEMS, SUB>;
    constexpr int GROUPS = SM::GROUPS;
    static_assert(SM::TILE <= 65536, "16-bit positions");
    constexpr int TILE = SM::TILE;
    constexpr int LB_BATCH = 4;
    constexpr int DSPLIT = (THREADS % RADIX == 0) ? THREADS / RADIX : 1;  // threads per digit in the digit phase
    constexpr int GPH = GROUPS / DSPLIT;
    constexpr int RBITS = (16 * JH <= 256) ? 8 : 16, RPR = 32 / RBITS;
    constexpr uint32_t RMASK = (1u << RBITS) - 1u;

    extern __shared__ __align__(128) uint8_t smem_raw[];
    uint32_t* s_out = reinterpret_cast<uint32_t*>(smem_raw);
    uint8_t* s_slot_raw = smem_raw + SM::OUT_BYTES;
    uint64_t* s_gofs = reinterpret_cast<uint64_t*>(smem_raw + SM::OUT_BYTES + SM::SLOT_BYTES);
    uint64_t* s_gptr = s_gofs + RADIX;  // REMAP only: per-part destination base pointers
    uint32_t* s_gofs32 = reinterpret_cast<uint32_t*>(s_gptr + RADIX);  // offsets mod 2^32 (n < 2^32)
    uint8_t* misc = smem_raw + SM::OUT_BYTES + SM::SLOT_BYTES + SM::GOFS_BYTES;
    volatile uint32_t* s_tile = reinterpret_cast<volatile uint32_t*>(misc);  // [0] first, [1] next tile id
    uint32_t* s_wtot = reinterpret_cast<uint32_t*>(misc + 32);  // 8 digit-warp totals
    uint32_t* s_part = reinterpret_cast<uint32_t*>(misc + SM::MISC_BYTES);  // DSPLIT > 1: per-part digit counts

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t lane_bit = 1u << (lane & 15);
    const uint32_t lt_mask = lane_bit - 1u;                                  // lanes below me in my half-warp
    const uint32_t slot_base = smem_u32(s_slot_raw);                         // [GROUPS][256] x 4 B
    const uint32_t my_slots = slot_base + uint32_t(tid >> 4) * SUB * RADIX * 4;  // this half-warp's SUB x 256 slots
    const DigitParams dp = p.dp;

    // all slots start as {mask 0, count 0}
    {
        uint4* z = reinterpret_cast<uint4*>(s_slot_raw);
        for (int i = tid; i < int(SM::SLOT_BYTES / 16); i += THREADS) z[i] = make_uint4(0, 0, 0, 0);
    }

    // segmented mode: (bucket, first tile of the bucket, first record, records) of a tile
    constexpr bool seg = SEG && !REMAP;  // a separate instantiation: the plain pass stays lean
    auto seg_find = [&](uint32_t t) -> uint32_t {  // bucket b with seg_tile_start[b] <= t < seg_tile_start[b + 1]
        uint32_t lo = 0, hi = RADIX;               // invariant: start[lo] <= t < start[hi]
        while (hi - lo > 1) {
            const uint32_t mid = (lo + hi) >> 1;
            if (__ldg(p.seg_tile_start + mid) <= t) lo = mid; else hi = mid;
        }
        return lo;
    };
    uint32_t cur_b = 0, cur_t0 = 0;  // bucket of the current tile and its first tile
    auto tile_valid_at = [&](uint32_t t, uint32_t b, uint32_t t0) -> uint32_t {
        const uint64_t nrec = seg ? __ldg(p.seg_rec_start + b + 1) - __ldg(p.seg_rec_start + b) : p.n;
        const uint64_t rem = nrec - uint64_t(t - t0) * TILE;
        return rem < uint64_t(TILE) ? uint32_t(rem) : uint32_t(TILE);
    };
    // element (half-warp h, j, lane16) = h*16*ITEMS + j*16 + lane16: every half-warp owns a
    // contiguous run of the tile, so (h, j, lane16) order is memory order
    const uint32_t e0 = uint32_t(tid >> 4) * 16u * ITEMS + (lane & 15);
    Rec<WORDS> r[ITEMS];
    auto load_tile = [&](uint32_t t, uint32_t b, uint32_t t0) {
        if (t >= p.num_tiles) return;
        const uint32_t v = tile_valid_at(t, b, t0);
        const bool f = v == uint32_t(TILE);
        const uint64_t first = (seg ? __ldg(p.seg_rec_start + b) : 0ull) + uint64_t(t - t0) * TILE;
        const uint32_t* g = p.src + first * WORDS;
#pragma unroll
        for (int j = 0; j < ITEMS; ++j) {
            const uint32_t e = e0 + j * 16;
            if (f || e < v) {
                r[j] = load_rec<WORDS>(g, e);
            } else {
#pragma unroll
                for (int k = 0; k < WORDS; ++k) r[j].w[k] = 0;
            }
        }
    };

    // Tiles are claimed in ticket order, one ahead: while a tile is processed the next one
    // is already claimed, and its records are fetched as soon as the registers are free
    // (right after the scatter), so the loads fly during the look-back and the write-out.
    if (tid == 0) {
        const uint32_t t = atomicAdd(p.ticket, 1u);
        s_tile[0] = t;
        s_tile[2] = (seg && t < p.num_tiles) ? seg_find(t) : 0u;
    }
    __syncthreads();  // first tile id visible; slots are clean
    uint32_t tile = s_tile[0];
    cur_b = s_tile[2];
    cur_t0 = seg ? __ldg(p.seg_tile_start + cur_b) : 0u;
    load_tile(tile, cur_b, cur_t0);

    while (tile < p.num_tiles) {
        if (tid == 0) {  // read by everyone after barrier (B)
            const uint32_t t = atomicAdd(p.ticket, 1u);
            s_tile[1] = t;
            s_tile[3] = (seg && t < p.num_tiles) ? seg_find(t) : 0u;
        }
        const uint32_t valid = tile_valid_at(tile, cur_b, cur_t0);
        const bool first_tile = tile == cur_t0;  // head of its look-back chain
        const bool full = valid == uint32_t(TILE);

        // ---- rank inside the warp by OR-matching; stable: lanes in order, rounds in order ----
        // ranks inside the half-warp group are < 16 * ITEMS: packed 4 (or 2) per register
        uint32_t rk[(ITEMS + RPR - 1) / RPR];
#pragma unroll
        for (int q = 0; q < (ITEMS + RPR - 1) / RPR; ++q) rk[q] = 0;
        uint32_t dcache[REMAP ? (ITEMS + 3) / 4 : 1];
#pragma unroll
        for (int q = 0; q < (REMAP ? (ITEMS + 3) / 4 : 1); ++q) dcache[q] = 0;
        // Three ways to find the lanes of my half-warp that hold my bucket:
        //  * OR-matching through the slot's mask field (default: 1.6 cycles per warp-op on random
        //    digits, but same-address RED.OR serialises: 6.6 .. 12 cycles when lanes share a digit,
        //    profiles/r1r_microbench_skew.txt);
        //  * hot buckets (up to two per pass, chosen by the host from the K1 histogram when a bucket
        //    holds more than ~3 % of the records): matched by one ballot each, counted in a register,
        //    never touching shared memory in this loop;
        //  * range partitions into few parts (REMAP, match_bits = ceil(log2 parts) <= 4): every lane
        //    collides, so the bucket id is matched bit by bit with ballots.
        const uint32_t hshift = lane & 16u;
        const bool use_hot = HOT && p.hot[0] != 0xffffffffu;  // HOT = false instantiations ignore p.hot
        const uint32_t mbits = REMAP ? p.match_bits : 0u;
        uint32_t hot_cnt0 = 0, hot_cnt1 = 0;  // hot records of this half-warp so far (uniform over its lanes)
#pragma unroll
        for (int j = 0; j < ITEMS; ++j) {
            const uint32_t e = e0 + j * 16;
            uint32_t d = 255u;  // padding sorts to the very end of the tile
            if (full || e < valid) {
                d = seg ? rec_digit_w<WORDS, FLOATK, seg>(dp, r[j]) : pass_bucket<WORDS, FLOATK, REMAP>(p, dp, r[j]);
            }
            if (REMAP) dcache[j / 4] |= d << ((j % 4) * 8);  // the part id is costly to recompute
            const uint32_t sa = my_slots + d * 4;
            uint32_t rank;
            if (REMAP && mbits) {
                uint32_t m = 0xffffu;
                if constexpr (REMAP) {
#pragma unroll 1
                    for (uint32_t bit = 0; bit < mbits; ++bit) {
                        const bool on = (d >> bit) & 1u;
                        const uint32_t bal = (__ballot_sync(0xffffffffu, on) >> hshift) & 0xffffu;
                        m &= on ? bal : ~bal;
                    }
                }
                const uint32_t c = lds_u32(sa) >> 16;
                __syncwarp();
                const uint32_t lt = m & lt_mask;
                if (lt == 0) sts_u32(sa, (c + __popc(m)) << 16);
                __syncwarp();
                rank = c + __popc(lt);
            } else {
                bool cold = true;
                uint32_t hrank = 0;
                if (HOT && use_hot) {
                    const bool h0 = d == p.hot[0], h1 = d == p.hot[1];
                    const uint32_t b0 = (__ballot_sync(0xffffffffu, h0) >> hshift) & 0xffffu;
                    const uint32_t b1 = (__ballot_sync(0xffffffffu, h1) >> hshift) & 0xffffu;
                    cold = !(h0 || h1);
                    hrank = h0 ? hot_cnt0 + __popc(b0 & lt_mask) : hot_cnt1 + __popc(b1 & lt_mask);
                    hot_cnt0 += __popc(b0);
                    hot_cnt1 += __popc(b1);
                }
                if (cold) slot_or(sa, lane_bit);
                __syncwarp();
                uint32_t v = 0;
                if (cold) v = lds_u32(sa);  // [15:0] lanes with my digit, [31:16] records of my digit in earlier rounds
                __syncwarp();
                const uint32_t m = v & 0xffffu, c = v >> 16;
                const uint32_t lt = m & lt_mask;
                if (cold && lt == 0) sts_u32(sa, (c + __popc(m)) << 16);
                __syncwarp();
                rank = cold ? c + __popc(lt) : hrank;
            }
            rk[j / RPR] |= rank << ((j % RPR) * RBITS);
        }
        if (HOT && use_hot && (lane & 15) == 0) {
            // the hot buckets' counts join the slot table for the digit phase (nobody else wrote these slots)
            if (hot_cnt0) sts_u32(my_slots + p.hot[0] * 4, hot_cnt0 << 16);
            if (hot_cnt1) sts_u32(my_slots + p.hot[1] * 4, hot_cnt1 << 16);
        }
        __syncthreads();  // (B) every warp's counts final, masks back to zero

        // ---- DSPLIT threads per digit: tile count, publish, local bases ----
        // thread (digit dg, part h) owns the GPH half-warp groups h*GPH .. of its digit
        uint32_t sum = 0, incl = 0, cnt = 0, part = 0, before = 0;
        uint32_t gc[GPH];  // this digit's count in each of my groups (read once)
        const uint32_t dg = tid & (RADIX - 1), h = tid / RADIX;
        const bool dthread = tid < RADIX * DSPLIT;
        if (dthread) {
            const uint32_t ca = slot_base + (h * GPH * RADIX + dg) * 4;
#pragma unroll
            for (int w = 0; w < GPH; ++w) {
                gc[w] = lds_u32(ca + w * RADIX * 4) >> 16;
                part += gc[w];
            }
            if (DSPLIT > 1) s_part[tid] = part;
        }
        if (DSPLIT > 1) __syncthreads();  // (B1)
        if (dthread) {
            sum = part;
            if (DSPLIT > 1) {
                sum = 0;
#pragma unroll
                for (int q = 0; q < DSPLIT; ++q) {
                    const uint32_t x = s_part[q * RADIX + dg];
                    sum += x;
                    if (uint32_t(q) < h) before += x;
                }
            }
            cnt = sum;
            if (dg == RADIX - 1) cnt -= uint32_t(TILE) - valid;
            if (h == 0)
                lb_store(&p.lookback[uint64_t(tile) * RADIX + dg], lb_pack(first_tile ? LB_PREFIX : LB_AGG, p.tag, cnt));
            incl = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += y;
            }
            if (h == 0 && lane == 31) s_wtot[warp] = incl;
        }
        __syncthreads();  // (B2)
        uint32_t local_off = 0;
        if (dthread) {
            uint32_t wbase = 0;
#pragma unroll
            for (int w = 0; w < RADIX / 32; ++w)
                if (uint32_t(w) < dg / 32) wbase += s_wtot[w];
            local_off = wbase + incl - sum;
            if (REMAP && h == 0) s_gofs32[dg] = local_off;  // REMAP: where every part's run starts in the staged tile
            // slot count := position of the (group, digit) run inside the locally sorted tile
            const uint32_t ca = slot_base + (h * GPH * RADIX + dg) * 4;
            uint32_t run = local_off + before;
#pragma unroll
            for (int w = 0; w < GPH; ++w) {
                sts_u32(ca + w * RADIX * 4, run << 16);
                run += gc[w];
            }
        }
        __syncthreads();  // (C)

        // ---- scatter into the staging buffer in locally sorted order ----
#pragma unroll
        for (int j = 0; j < ITEMS; ++j) {
            const uint32_t e = e0 + j * 16;
            uint32_t d = 255u;
            if (REMAP) {
                d = (dcache[j / 4] >> ((j % 4) * 8)) & 0xffu;
            } else if (full || e < valid) {
                d = rec_digit_w<WORDS, FLOATK, seg>(dp, r[j]);
            }
            const uint32_t pos = (lds_u32(my_slots + (j / JH) * RADIX * 4 + d * 4) >> 16) + ((rk[j / RPR] >> ((j % RPR) * RBITS)) & RMASK);
            store_rec<WORDS>(s_out, pos, r[j]);
        }
        const uint32_t next_tile = s_tile[1];
        const uint32_t next_b = s_tile[3];
        const uint32_t next_t0 = seg ? __ldg(p.seg_tile_start + next_b) : 0u;
        load_tile(next_tile, next_b, next_t0);  // registers are free again: fetch the next tile now
        __syncwarp();
        // this warp's slots are only read by this warp: reset them for the next tile now
        {
            uint4* z = reinterpret_cast<uint4*>(s_slot_raw + size_t(warp) * 2 * SUB * RADIX * 4);
#pragma unroll
            for (int i = lane; i < 2 * SUB * RADIX * 4 / 16; i += 32) z[i] = make_uint4(0, 0, 0, 0);
        }

        // ---- decoupled look-back, one thread per digit, LB_BATCH predecessors per round trip ----
        if (tid < RADIX) {
            uint64_t excl = 0;
            if (!first_tile) {
                int64_t t = int64_t(tile) - 1;
                bool done = false;
                while (!done) {
                    const int64_t t_before = t;
                    uint64_t v[LB_BATCH];
#pragma unroll
                    for (int b = 0; b < LB_BATCH; ++b) {
                        const int64_t tb = t - b;
                        v[b] = tb >= 0 ? lb_load(&p.lookback[uint64_t(tb) * RADIX + tid]) : 0ull;
                    }
#pragma unroll
                    for (int b = 0; b < LB_BATCH; ++b) {
                        if (done) break;
                        const uint64_t st = v[b] >> 62;
                        if (st == 0 || uint32_t((v[b] >> 54) & 0xffu) != (p.tag & 0xffu)) break;  // not published yet
                        excl += v[b] & LB_VALUE_MASK;
                        --t;
                        if (st == LB_PREFIX) done = true;
                    }
                    if (!done && t == t_before) __nanosleep(64);  // predecessor not published yet: back off
                }
                lb_store(&p.lookback[uint64_t(tile) * RADIX + tid], lb_pack(LB_PREFIX, p.tag, excl + cnt));
            }
            if (REMAP && p.dst_tab) {
                // range partition with per-part destinations (possibly peer GPUs over NVLink):
                // keep a record index relative to the part's own base pointer
                s_gofs[tid] = p.bucket_base[tid] + excl - local_off;
                s_gptr[tid] = p.dst_tab[tid];
            } else {
                const uint64_t bb = seg ? __ldg(p.seg_bases + size_t(cur_b) * RADIX + tid) : p.bucket_base[tid];
                const uint64_t go = bb + excl - local_off;
                s_gofs[tid] = go;
                s_gofs32[tid] = uint32_t(go);
                if (REMAP) s_gptr[tid] = reinterpret_cast<uint64_t>(p.dst);
            }
        }
        __syncthreads();  // (D)

        // ---- write-out: runs of equal digits are contiguous in s_out and in dst ----
        if constexpr (WORDS == 1 || WORDS == 2 || WORDS == 4) {
            if (REMAP && p.dst_tab && p.nparts != 0 && p.nparts <= 16) {
                // few parts, each with its own destination (peer GPUs over NVLink): part by part, with the warps lined
                // up on the DESTINATION's 128-byte lines.  A run starts wherever the previous tile's run ended; stores
                // issued in staging order straddle three lines per warp, two of them partially, which costs the link
                // an eighth of its rate (profiles/r12_exchange_experiments.txt: 628 against 715 GB/s).
                constexpr uint32_t LINE = 128 / (4 * WORDS);  // records per line
                for (uint32_t q = 0; q < p.nparts; ++q) {
                    const uint32_t a = s_gofs32[q], b = q + 1 < uint32_t(RADIX) ? s_gofs32[q + 1] : valid;
                    if (a == b) continue;
                    uint32_t* const dstq = reinterpret_cast<uint32_t*>(s_gptr[q]);
                    const uint64_t go = s_gofs[q];
                    const uint32_t mis = uint32_t((s_gptr[q] / (4 * WORDS) + go + a) % LINE);
                    for (int64_t i = int64_t(a) - mis + tid; i < int64_t(b); i += THREADS)
                        if (i >= int64_t(a)) store_rec<WORDS>(dstq, go + uint64_t(i), load_rec<WORDS>(s_out, uint32_t(i)));
                }
            } else if (!REMAP && full && p.n <= 0xffffffffull) {
                // record indices fit 32 bits: 32-bit offset table (half the shared-memory wavefronts)
#pragma unroll 4
                for (int k = 0; k < ITEMS; ++k) {
                    const uint32_t i = tid + k * THREADS;
                    const Rec<WORDS> x = load_rec<WORDS>(s_out, i);
                    const uint32_t d = rec_digit_w<WORDS, FLOATK, seg>(dp, x);
                    store_rec<WORDS>(p.dst, uint64_t(s_gofs32[d] + i), x);
                }
            } else if (full) {
#pragma unroll 4
                for (int k = 0; k < ITEMS; ++k) {
                    const uint32_t i = tid + k * THREADS;
                    const Rec<WORDS> x = load_rec<WORDS>(s_out, i);
                    const uint32_t d = seg ? rec_digit_w<WORDS, FLOATK, seg>(dp, x) : pass_bucket<WORDS, FLOATK, REMAP>(p, dp, x);
                    store_rec<WORDS>(REMAP ? reinterpret_cast<uint32_t*>(s_gptr[d]) : p.dst, s_gofs[d] + i, x);
                }
            } else {
                for (uint32_t i = tid; i < valid; i += THREADS) {
                    const Rec<WORDS> x = load_rec<WORDS>(s_out, i);
                    const uint32_t d = seg ? rec_digit_w<WORDS, FLOATK, seg>(dp, x) : pass_bucket<WORDS, FLOATK, REMAP>(p, dp, x);
                    store_rec<WORDS>(REMAP ? reinterpret_cast<uint32_t*>(s_gptr[d]) : p.dst, s_gofs[d] + i, x);
                }
            }
        } else {
            for (uint32_t wi = tid; wi < valid * WORDS; wi += THREADS) {
                const uint32_t pos = wi / WORDS;
                const uint32_t k = wi - pos * WORDS;
                const uint32_t kw = s_out[pos * WORDS + dp.word];
                const uint32_t sw = FLOATK ? s_out[pos * WORDS + dp.sign_word] : 0u;
                uint32_t d;
                if (REMAP && p.sp.nsplit) {
                    d = prefix_part(p.sp, rec_prefix<WORDS, FLOATK>(p.sp, load_rec<WORDS>(s_out, pos)));
                } else {
                    d = digit_from_words<FLOATK>(dp, kw, sw);
                    if (REMAP) d = p.remap[d];
                }
                (REMAP ? reinterpret_cast<uint32_t*>(s_gptr[d]) : p.dst)[(s_gofs[d] + pos) * WORDS + k] = s_out[wi];
            }
        }
        // barriers (B)..(C) of the next iteration order this write-out before the next scatter
        tile = next_tile;
        cur_b = next_b;
        cur_t0 = next_t0;
    }
}

}  // namespace b200
