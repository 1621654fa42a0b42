// pass_or.cuh — K2 variant "or": the stable one-level LSD pass with shared-memory
// OR-matching instead of ballot matching.
//
// Same contract as onesweep_pass_kernel (onesweep.cuh): one launch = one stable 256-way
// counting-sort scatter of every record for one 8-bit level, the device counterpart of
//   out_of_place_sort / lr_out_of_place_sort   (reference src/sorts/out_of_place_sort.rs:51-108,201-275)
//   the per-level body of mt_lsb_sort          (reference src/sorts/mt_lsb_sort.rs:40-133)
// What differs is how a warp finds, for each of its 32 records, the lanes holding the same
// digit.  Measured on B200 (profiles/r1a_microbench.txt): eight ballots cost 18.3
// SM-cycles per warp-step and MATCH.ANY 47.8 on random digits, while a shared-memory
// atomic costs 1.8.  So each warp owns 256 8-byte slots {mask, count}: every lane ORs its
// lane bit into its digit's mask (RED.OR, order-free so the result is deterministic),
// reads {mask, count} back, and the lowest lane of each digit group stores
// {0, count + popc(mask)}.  rank = count + popc(mask & lanes_below): stable by lane order.
//
// Tiles are small (THREADS*ITEMS records) and several CTAs share an SM, so the phases of
// different tiles overlap; records are loaded straight into registers (coalesced LDG),
// reordered once through shared memory and written as per-digit runs.
#pragma once
#include "onesweep.cuh"

namespace b200 {

template <int WORDS, int THREADS, int ITEMS>
struct OrSmem {
    static constexpr int WARPS = THREADS / 32;
    static constexpr int TILE = THREADS * ITEMS;
    static constexpr size_t OUT_BYTES = size_t(TILE) * WORDS * 4;
    static constexpr size_t SLOT_BYTES = size_t(WARPS) * RADIX * 8;
    static constexpr size_t GOFS_BYTES = RADIX * 8;
    static constexpr size_t MISC_BYTES = 128;
    static constexpr size_t TOTAL = OUT_BYTES + SLOT_BYTES + GOFS_BYTES + MISC_BYTES;
};

__device__ __forceinline__ void slot_or(uint32_t saddr, uint32_t bit) {
    asm volatile("red.shared.or.b32 [%0], %1;" ::"r"(saddr), "r"(bit) : "memory");
}
__device__ __forceinline__ uint2 slot_ld(uint32_t saddr) {
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(saddr) : "memory");
    return v;
}
__device__ __forceinline__ void slot_st(uint32_t saddr, uint32_t a, uint32_t b) {
    asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(saddr), "r"(a), "r"(b) : "memory");
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t saddr) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(saddr) : "memory");
    return v;
}
__device__ __forceinline__ void sts_u32(uint32_t saddr, uint32_t v) {
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(saddr), "r"(v) : "memory");
}

template <int WORDS, int THREADS, int ITEMS, bool FLOATK, bool REMAP, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) lsd_pass_or_kernel(const PassParams p) {
    static_assert(THREADS >= RADIX && THREADS % 32 == 0, "one thread per digit is needed for the look-back");
    using SM = OrSmem<WORDS, THREADS, ITEMS>;
    constexpr int WARPS = SM::WARPS;
    constexpr int TILE = SM::TILE;
    constexpr int LB_BATCH = 4;

    extern __shared__ __align__(128) uint8_t smem_raw[];
    uint32_t* s_out = reinterpret_cast<uint32_t*>(smem_raw);
    uint8_t* s_slot_raw = smem_raw + SM::OUT_BYTES;
    uint64_t* s_gofs = reinterpret_cast<uint64_t*>(smem_raw + SM::OUT_BYTES + SM::SLOT_BYTES);
    uint8_t* misc = smem_raw + SM::OUT_BYTES + SM::SLOT_BYTES + SM::GOFS_BYTES;
    volatile uint32_t* s_tile = reinterpret_cast<volatile uint32_t*>(misc);  // [0] first, [1] next tile id
    uint32_t* s_wtot = reinterpret_cast<uint32_t*>(misc + 32);  // 8 digit-warp totals

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t lt_mask = lanemask_lt();
    const uint32_t lane_bit = 1u << lane;
    const uint32_t slot_base = smem_u32(s_slot_raw);                   // [WARPS][256] x 8 B
    const uint32_t my_slots = slot_base + uint32_t(warp) * RADIX * 8;  // this warp's 256 slots
    const DigitParams dp = p.dp;

    // all slots start as {mask 0, count 0}
    {
        uint4* z = reinterpret_cast<uint4*>(s_slot_raw);
        for (int i = tid; i < int(SM::SLOT_BYTES / 16); i += THREADS) z[i] = make_uint4(0, 0, 0, 0);
    }

    auto tile_valid = [&](uint32_t t) -> uint32_t {
        const uint64_t rem = p.n - uint64_t(t) * TILE;
        return rem < uint64_t(TILE) ? uint32_t(rem) : uint32_t(TILE);
    };
    // element (warp, j, lane) = warp*32*ITEMS + j*32 + lane, i.e. memory order
    const uint32_t e0 = uint32_t(warp) * 32u * ITEMS + lane;
    Rec<WORDS> r[ITEMS];
    auto load_tile = [&](uint32_t t) {
        if (t >= p.num_tiles) return;
        const uint32_t v = tile_valid(t);
        const bool f = v == uint32_t(TILE);
        const uint32_t* g = p.src + uint64_t(t) * TILE * WORDS;
#pragma unroll
        for (int j = 0; j < ITEMS; ++j) {
            const uint32_t e = e0 + j * 32;
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
    if (tid == 0) s_tile[0] = atomicAdd(p.ticket, 1u);
    __syncthreads();  // first tile id visible; slots are clean
    uint32_t tile = s_tile[0];
    load_tile(tile);

    while (tile < p.num_tiles) {
        if (tid == 0) s_tile[1] = atomicAdd(p.ticket, 1u);  // read by everyone after barrier (B)
        const uint32_t valid = tile_valid(tile);
        const bool full = valid == uint32_t(TILE);

        // ---- rank inside the warp by OR-matching; stable: lanes in order, rounds in order ----
        uint32_t rank[ITEMS];
#pragma unroll
        for (int j = 0; j < ITEMS; ++j) {
            const uint32_t e = e0 + j * 32;
            uint32_t d = 255u;  // padding sorts to the very end of the tile
            if (full || e < valid) {
                d = rec_digit<WORDS, FLOATK>(dp, r[j]);
                if (REMAP) d = p.remap[d];
            }
            const uint32_t sa = my_slots + d * 8;
            slot_or(sa, lane_bit);
            __syncwarp();
            const uint2 v = slot_ld(sa);  // {lanes with my digit, records of my digit in earlier rounds}
            __syncwarp();
            const uint32_t lt = v.x & lt_mask;
            if (lt == 0) slot_st(sa, 0u, v.y + __popc(v.x));
            __syncwarp();
            rank[j] = v.y + __popc(lt);
        }
        __syncthreads();  // (B) every warp's counts final, masks back to zero

        // ---- one thread per digit: tile count, publish, local bases ----
        uint32_t sum = 0, incl = 0, cnt = 0;
        if (tid < RADIX) {
            const uint32_t ca = slot_base + uint32_t(tid) * 8 + 4;
#pragma unroll
            for (int w = 0; w < WARPS; ++w) sum += lds_u32(ca + w * RADIX * 8);
            cnt = sum;
            if (tid == RADIX - 1) cnt -= uint32_t(TILE) - valid;
            lb_store(&p.lookback[uint64_t(tile) * RADIX + tid], lb_pack(tile == 0 ? LB_PREFIX : LB_AGG, p.tag, cnt));
            incl = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += y;
            }
            if (lane == 31) s_wtot[warp] = incl;
        }
        __syncthreads();  // (B2)
        uint32_t local_off = 0;
        if (tid < RADIX) {
            uint32_t wbase = 0;
#pragma unroll
            for (int w = 0; w < RADIX / 32; ++w)
                if (w < warp) wbase += s_wtot[w];
            local_off = wbase + incl - sum;
            // slot count := position of the (warp, digit) group inside the locally sorted tile
            const uint32_t ca = slot_base + uint32_t(tid) * 8 + 4;
            uint32_t run = local_off;
#pragma unroll
            for (int w = 0; w < WARPS; ++w) {
                const uint32_t c = lds_u32(ca + w * RADIX * 8);
                sts_u32(ca + w * RADIX * 8, run);
                run += c;
            }
        }
        __syncthreads();  // (C)

        // ---- scatter into the staging buffer in locally sorted order ----
#pragma unroll
        for (int j = 0; j < ITEMS; ++j) {
            const uint32_t e = e0 + j * 32;
            uint32_t d = 255u;
            if (full || e < valid) {
                d = rec_digit<WORDS, FLOATK>(dp, r[j]);
                if (REMAP) d = p.remap[d];
            }
            const uint32_t pos = lds_u32(my_slots + d * 8 + 4) + rank[j];
            store_rec<WORDS>(s_out, pos, r[j]);
        }
        const uint32_t next_tile = s_tile[1];
        load_tile(next_tile);  // registers are free again: fetch the next tile now
        __syncwarp();
        // this warp's slots are only read by this warp: reset them for the next tile now
        {
            uint4* z = reinterpret_cast<uint4*>(s_slot_raw + size_t(warp) * RADIX * 8);
#pragma unroll
            for (int i = lane; i < RADIX * 8 / 16; i += 32) z[i] = make_uint4(0, 0, 0, 0);
        }

        // ---- decoupled look-back, one thread per digit, LB_BATCH predecessors per round trip ----
        if (tid < RADIX) {
            uint64_t excl = 0;
            if (tile > 0) {
                int64_t t = int64_t(tile) - 1;
                bool done = false;
                while (!done) {
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
                }
                lb_store(&p.lookback[uint64_t(tile) * RADIX + tid], lb_pack(LB_PREFIX, p.tag, excl + cnt));
            }
            s_gofs[tid] = p.bucket_base[tid] + excl - local_off;
        }
        __syncthreads();  // (D)

        // ---- write-out: runs of equal digits are contiguous in s_out and in dst ----
        if constexpr (WORDS == 1 || WORDS == 2 || WORDS == 4) {
            if (full) {
#pragma unroll 4
                for (int k = 0; k < ITEMS; ++k) {
                    const uint32_t i = tid + k * THREADS;
                    const Rec<WORDS> x = load_rec<WORDS>(s_out, i);
                    uint32_t d = rec_digit<WORDS, FLOATK>(dp, x);
                    if (REMAP) d = p.remap[d];
                    store_rec<WORDS>(p.dst, s_gofs[d] + i, x);
                }
            } else {
                for (uint32_t i = tid; i < valid; i += THREADS) {
                    const Rec<WORDS> x = load_rec<WORDS>(s_out, i);
                    uint32_t d = rec_digit<WORDS, FLOATK>(dp, x);
                    if (REMAP) d = p.remap[d];
                    store_rec<WORDS>(p.dst, s_gofs[d] + i, x);
                }
            }
        } else {
            for (uint32_t wi = tid; wi < valid * WORDS; wi += THREADS) {
                const uint32_t pos = wi / WORDS;
                const uint32_t k = wi - pos * WORDS;
                const uint32_t kw = s_out[pos * WORDS + dp.word];
                const uint32_t sw = FLOATK ? s_out[pos * WORDS + dp.sign_word] : 0u;
                uint32_t d = digit_from_words<FLOATK>(dp, kw, sw);
                if (REMAP) d = p.remap[d];
                p.dst[(s_gofs[d] + pos) * WORDS + k] = s_out[wi];
            }
        }
        // barriers (B)..(C) of the next iteration order this write-out before the next scatter
        tile = next_tile;
    }
}

}  // namespace b200
