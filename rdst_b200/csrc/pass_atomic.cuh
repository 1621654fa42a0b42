// pass_atomic.cuh — K2-U, the UNSTABLE one-digit MSD pass of the hybrid route.
//
// rdst's default entry point is radix_sort_UNSTABLE (reference src/radix_sort.rs:21-45): its MSB-first
// algorithms (ska, regions, recombinating, scanning: reference src/sorts/*.rs) permute a bucket's
// records freely, because a later level sorts each bucket anyway.  The GPU form of that freedom: when
// a record is nothing but its key (every byte is a level), records that tie on an MSD digit may land
// in any order inside the digit's bucket — the segment sort that follows (K3-U, local_count.cuh)
// orders them by the rest of the key, and equal keys are indistinguishable, so the final array is
// bit-identical to the stable engine's.  Giving up stability inside a tile removes the half-warp
// OR-matching (RED.OR + LDS + leader STS per record, three dependent shared-memory round trips):
//
//   count    one RED.ADD per record into ONE 256-counter table per CTA (no return value, no ordering)
//   scan     256 threads: tile count -> look-back word, exclusive scan -> the counters become bases
//   scatter  one ATOMS.ADD per record returns its position in the locally sorted tile -> staging buffer
//   look-back + coalesced write-out exactly as K2 (pass_or16.cuh)
//
// Two shared-memory atomics and one scatter per record instead of five random-address operations;
// no per-group slot tables and no per-group digit phase.  The digit is a bit window (hybrid_plan),
// the array is 1..256 independently partitioned buckets (PassParams::seg_*), as in K2's SEG mode.
// Same-address atomics serialise, so the planner keeps skewed digits on the stable K2 (its ballot
// path for hot buckets) — see pick_unstable() in b200sort.cu.
#pragma once
#include <type_traits>

#include "pass_or.cuh"

namespace b200 {

template <int WORDS, int THREADS, int ITEMS>
struct AtomSmem {
    static constexpr int TILE = THREADS * ITEMS;
    static constexpr size_t OUT_BYTES = size_t(TILE) * WORDS * 4;
    static constexpr size_t CNT_BYTES = 2 * RADIX * 4;          // two counter tables, alternating per tile
    static constexpr size_t GOFS_BYTES = RADIX * 8 + RADIX * 4;  // u64 and u32 global offsets
    static constexpr size_t MISC_BYTES = 128;
    static constexpr size_t TOTAL = OUT_BYTES + CNT_BYTES + GOFS_BYTES + MISC_BYTES;
};

__device__ __forceinline__ void red_add_shared(uint32_t saddr, uint32_t v) {
    asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(saddr), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t atom_add_shared(uint32_t saddr, uint32_t v) {
    uint32_t old;
    asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(old) : "r"(saddr), "r"(v) : "memory");
    return old;
}

template <int WORDS, int THREADS, int ITEMS, bool FLOATK, int MINB, bool VEC = false>
__global__ void __launch_bounds__(THREADS, MINB) msd_pass_atomic_kernel(const PassParams p) {
    static_assert(!VEC || (ITEMS % 2 == 0 && ITEMS <= 32), "vector loads take records in pairs");
    static_assert(THREADS >= RADIX && THREADS % 32 == 0, "one thread per digit is needed for the look-back");
    using SM = AtomSmem<WORDS, THREADS, ITEMS>;
    constexpr int TILE = SM::TILE;
    // predecessors read per look-back round trip (8 / 16 / 32 measured slower: 4.62 / 5.20 / 9.1 ms per pass against
    // 4.58 ms at 1 B u64 keys, profiles/r6_k2u_experiments.txt)
    constexpr int LB_BATCH = 4;

    extern __shared__ __align__(128) uint8_t smem_raw[];
    uint32_t* s_out = reinterpret_cast<uint32_t*>(smem_raw);
    uint32_t* s_cnt = reinterpret_cast<uint32_t*>(smem_raw + SM::OUT_BYTES);  // [2][256]
    uint64_t* s_gofs = reinterpret_cast<uint64_t*>(smem_raw + SM::OUT_BYTES + SM::CNT_BYTES);
    uint32_t* s_gofs32 = reinterpret_cast<uint32_t*>(s_gofs + RADIX);
    uint8_t* misc = smem_raw + SM::OUT_BYTES + SM::CNT_BYTES + SM::GOFS_BYTES;
    volatile uint32_t* s_tile = reinterpret_cast<volatile uint32_t*>(misc);  // [0] first, [1] next tile id, [2],[3] their buckets
    uint32_t* s_wtot = reinterpret_cast<uint32_t*>(misc + 32);               // 8 digit-warp totals

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const DigitParams dp = p.dp;
    const uint32_t cnt_base = smem_u32(s_cnt);

    for (int i = tid; i < 2 * RADIX; i += THREADS) s_cnt[i] = 0;

    auto seg_find = [&](uint32_t t) -> uint32_t {  // bucket b with seg_tile_start[b] <= t < seg_tile_start[b + 1]
        uint32_t lo = 0, hi = RADIX;
        while (hi - lo > 1) {
            const uint32_t mid = (lo + hi) >> 1;
            if (__ldg(p.seg_tile_start + mid) <= t) lo = mid; else hi = mid;
        }
        return lo;
    };
    auto tile_valid_at = [&](uint32_t t, uint32_t b, uint32_t t0) -> uint32_t {
        const uint64_t nrec = __ldg(p.seg_rec_start + b + 1) - __ldg(p.seg_rec_start + b);
        const uint64_t rem = nrec - uint64_t(t - t0) * TILE;
        return rem < uint64_t(TILE) ? uint32_t(rem) : uint32_t(TILE);
    };
    // the order of records inside a bucket is free: plain coalesced element order e = j * THREADS + tid, or
    // (VEC) records in aligned pairs — one 16-byte load per two u64 keys, half the load instructions
    Rec<WORDS> r[ITEMS];
    uint32_t vmask = 0;  // VEC: bit j = r[j] holds a record of the tile
    auto load_tile = [&](uint32_t t, uint32_t b, uint32_t t0) {
        vmask = 0;
        if (t >= p.num_tiles) return;
        const uint32_t v = tile_valid_at(t, b, t0);
        const uint64_t first = __ldg(p.seg_rec_start + b) + uint64_t(t - t0) * TILE;
        const uint32_t* g = p.src + first * WORDS;
        if constexpr (VEC) {
            // a tile starts at any record: `head` records (0 or 1) up to pair alignment, npairs aligned pairs, maybe
            // one tail record; head and tail go to the last slots of the last thread (free whenever they exist)
            using Vec2 = typename std::conditional<WORDS == 1, uint2, uint4>::type;
            const uint32_t head = uint32_t(first) & 1u, npairs = (v - head) >> 1;
            const Vec2* gp = reinterpret_cast<const Vec2*>(g + head * WORDS);
#pragma unroll
            for (int k = 0; k < ITEMS / 2; ++k) {
                const uint32_t q = uint32_t(k * THREADS + tid);
                if (q < npairs) {
                    const Vec2 x = gp[q];
                    if constexpr (WORDS == 1) {
                        r[2 * k].w[0] = x.x, r[2 * k + 1].w[0] = x.y;
                    } else {
                        r[2 * k].w[0] = x.x, r[2 * k].w[1] = x.y, r[2 * k + 1].w[0] = x.z, r[2 * k + 1].w[1] = x.w;
                    }
                    vmask |= 3u << (2 * k);
                }
            }
            if (tid == THREADS - 1) {
                if (head) r[ITEMS - 2] = load_rec<WORDS>(g, 0), vmask |= 1u << (ITEMS - 2);
                if ((v - head) & 1u) r[ITEMS - 1] = load_rec<WORDS>(g, v - 1), vmask |= 1u << (ITEMS - 1);
            }
        } else if (v == uint32_t(TILE)) {
#pragma unroll
            for (int j = 0; j < ITEMS; ++j) r[j] = load_rec<WORDS>(g, uint32_t(j * THREADS + tid));
        } else {
#pragma unroll
            for (int j = 0; j < ITEMS; ++j) {
                const uint32_t e = uint32_t(j * THREADS + tid);
                if (e < v) {
                    r[j] = load_rec<WORDS>(g, e);
                } else {
#pragma unroll
                    for (int k = 0; k < WORDS; ++k) r[j].w[k] = 0;
                }
            }
        }
    };

    if (tid == 0) {
        const uint32_t t = atomicAdd(p.ticket, 1u);
        s_tile[0] = t;
        s_tile[2] = t < p.num_tiles ? seg_find(t) : 0u;
    }
    __syncthreads();  // first tile id visible; counters are clean
    uint32_t tile = s_tile[0];
    uint32_t cur_b = s_tile[2];
    uint32_t cur_t0 = __ldg(p.seg_tile_start + cur_b);
    load_tile(tile, cur_b, cur_t0);
    uint32_t par = 0;  // which counter table this tile uses

    while (tile < p.num_tiles) {
        // (claiming the next tile later in the trip — after the scan, or after the scatter — so that it publishes its
        // counts sooner after the claim changed nothing: 4.75 / 4.65 against 4.55 ms, profiles/r6_k2u_experiments.txt)
        if (tid == 0) {  // read by everyone after barrier (B)
            const uint32_t t = atomicAdd(p.ticket, 1u);
            s_tile[1] = t;
            s_tile[3] = t < p.num_tiles ? seg_find(t) : 0u;
        }
        const uint32_t valid = tile_valid_at(tile, cur_b, cur_t0);
        const bool first_tile = tile == cur_t0;  // head of its bucket's look-back chain
        const bool full = valid == uint32_t(TILE);
        const uint32_t my_cnt = cnt_base + par * RADIX * 4;

        // ---- count (the digits are kept, four per register, for the scatter) ----
        const uint32_t mine = vmask;
        auto have = [&](int j) { return VEC ? bool(mine >> j & 1u) : (full || uint32_t(j * THREADS + tid) < valid); };
        uint32_t dcache[(ITEMS + 3) / 4];
#pragma unroll
        for (int q = 0; q < (ITEMS + 3) / 4; ++q) dcache[q] = 0;
#pragma unroll
        for (int j = 0; j < ITEMS; ++j) {
            if (have(j)) {
                const uint32_t d = rec_digit_w<WORDS, FLOATK, true>(dp, r[j]);
                dcache[j / 4] |= d << ((j % 4) * 8);
                red_add_shared(my_cnt + d * 4, 1u);
            }
        }
        __syncthreads();  // (B)

        // ---- one thread per digit: publish the tile count, exclusive scan, counters := local bases ----
        uint32_t cnt = 0, local_off = 0;
        uint64_t pre0 = 0, pre1 = 0;
        if (tid < RADIX) {
            cnt = s_cnt[par * RADIX + tid];
            lb_store(&p.lookback[uint64_t(tile) * RADIX + tid], lb_pack(first_tile ? LB_PREFIX : LB_AGG, p.tag, cnt));
            // first look-back reads, issued now: they fly during the scan and the scatter
            if (!first_tile) {
                pre0 = lb_load(&p.lookback[uint64_t(tile - 1) * RADIX + tid]);
                pre1 = tile - 1 > cur_t0 ? lb_load(&p.lookback[uint64_t(tile - 2) * RADIX + tid]) : 0ull;
            }
            uint32_t incl = cnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += y;
            }
            if (lane == 31) s_wtot[warp] = incl;
            local_off = incl - cnt;
            s_cnt[(par ^ 1) * RADIX + tid] = 0;  // the other table: last touched before barrier (D) of the previous tile
        }
        __syncthreads();  // (B2)
        if (tid < RADIX) {
#pragma unroll
            for (int w = 0; w < RADIX / 32; ++w)
                if (w < warp) local_off += s_wtot[w];
            s_cnt[par * RADIX + tid] = local_off;
        }
        __syncthreads();  // (C)

        // ---- scatter into the staging buffer: the atomic hands out the positions of a digit's run ----
        // (in batches: SB atomics in flight per thread before the first dependent store)
        constexpr int SB = ITEMS % 8 == 0 ? 8 : (ITEMS % 6 == 0 ? 6 : 4);
        static_assert(ITEMS % SB == 0, "scatter batches");
#pragma unroll
        for (int j0 = 0; j0 < ITEMS; j0 += SB) {
            uint32_t pos[SB];
#pragma unroll
            for (int u = 0; u < SB; ++u) {
                const int j = j0 + u;
                pos[u] = 0;
                if (have(j)) pos[u] = atom_add_shared(my_cnt + ((dcache[j / 4] >> ((j % 4) * 8)) & 0xffu) * 4, 1u);
            }
#pragma unroll
            for (int u = 0; u < SB; ++u) {
                const int j = j0 + u;
                if (have(j)) store_rec<WORDS>(s_out, pos[u], r[j]);
            }
        }
        const uint32_t next_tile = s_tile[1];
        const uint32_t next_b = s_tile[3];
        const uint32_t next_t0 = __ldg(p.seg_tile_start + next_b);
        load_tile(next_tile, next_b, next_t0);  // registers are free again: fetch the next tile now

        // ---- decoupled look-back, one thread per digit ----
        if (tid < RADIX) {
            uint64_t excl = 0;
            if (!first_tile) {
                int64_t t = int64_t(tile) - 1;
                bool done = false;
                {   // the early reads first (an entry, once published, stays valid for the whole pass)
                    const uint64_t pre[2] = {pre0, pre1};
#pragma unroll
                    for (int b = 0; b < 2; ++b) {
                        if (done) break;
                        const uint64_t st = pre[b] >> 62;
                        if (st == 0 || uint32_t((pre[b] >> 54) & 0xffu) != (p.tag & 0xffu)) break;
                        excl += pre[b] & LB_VALUE_MASK;
                        --t;
                        if (st == LB_PREFIX) done = true;
                    }
                }
                while (!done) {
                    const int64_t t_before = t;
                    uint64_t v[LB_BATCH];
#pragma unroll
                    for (int b = 0; b < LB_BATCH; ++b) {
                        const int64_t tb = t - b;
                        v[b] = tb >= int64_t(cur_t0) ? lb_load(&p.lookback[uint64_t(tb) * RADIX + tid]) : 0ull;
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
                    if (!done && t == t_before) __nanosleep(64);
                }
                lb_store(&p.lookback[uint64_t(tile) * RADIX + tid], lb_pack(LB_PREFIX, p.tag, excl + cnt));
            }
            const uint64_t go = __ldg(p.seg_bases + size_t(cur_b) * RADIX + tid) + excl - local_off;
            s_gofs[tid] = go;
            s_gofs32[tid] = uint32_t(go);
        }
        __syncthreads();  // (D) staging complete, offsets visible

        // ---- write-out: runs of equal digits are contiguous in s_out and in dst ----
        if (full && p.n <= 0xffffffffull) {
#pragma unroll 4
            for (int k = 0; k < ITEMS; ++k) {
                const uint32_t i = tid + k * THREADS;
                const Rec<WORDS> x = load_rec<WORDS>(s_out, i);
                const uint32_t d = rec_digit_w<WORDS, FLOATK, true>(dp, x);
                store_rec<WORDS>(p.dst, uint64_t(s_gofs32[d] + i), x);
            }
        } else {
            for (uint32_t i = tid; i < valid; i += THREADS) {
                const Rec<WORDS> x = load_rec<WORDS>(s_out, i);
                const uint32_t d = rec_digit_w<WORDS, FLOATK, true>(dp, x);
                store_rec<WORDS>(p.dst, s_gofs[d] + i, x);
            }
        }
        // barriers (B)..(C) of the next iteration order this write-out before the next scatter
        tile = next_tile;
        cur_b = next_b;
        cur_t0 = next_t0;
        par ^= 1;
    }
}

}  // namespace b200
