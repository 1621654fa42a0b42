// onesweep.cuh — K2, the stable one-level LSD pass of the B200 radix sort engine.
//
// One launch moves every record from `src` to its final place for one 8-bit
// level in `dst`: a stable 256-way counting-sort scatter.  This is the device
// counterpart of rdst's per-level scatter routines
//   out_of_place_sort / lr_out_of_place_sort   (reference src/sorts/out_of_place_sort.rs:51-108,201-275)
//   the per-level body of mt_lsb_sort          (reference src/sorts/mt_lsb_sort.rs:40-133)
// where rdst's "(radix, tile)-ordered pre-split of dst" (mt_lsb_sort.rs:51-63)
// becomes: global bucket base (from the K1 histogram) + exclusive prefix over
// earlier tiles, the latter obtained by decoupled look-back instead of a
// counting pre-pass, so each pass reads and writes every record exactly once.
//
// Layout: a record is WORDS little-endian 32-bit words; a tile is THREADS*ITEMS
// consecutive records.  Persistent CTAs claim tiles in ticket order (so every
// predecessor of a claimed tile is owned by a running CTA and look-back cannot
// deadlock).  With TMA the next tile is fetched by a 1-D bulk asynchronous copy
// (UBLKCP) into shared memory while the current one is ranked and written out.
#pragma once
#include "common.cuh"

namespace b200 {

// Range partition on full-resolution splitters (sample-sort style):
// the record's 64-bit key prefix = its top min(LEVELS, 8) levels as an order-preserving
// integer, and part = #{g : prefix > split[g] or (prefix == split[g] and incl bit g)}.
// The incl bits differ per source rank so that a key value heavier than one part is split
// between adjacent parts by source rank (which keeps the exchange stable).
constexpr int MAX_SPLITTERS = 15;
struct SplitParams {
    uint32_t nsplit;      // 0 = partition on one digit through `remap`
    uint32_t incl_mask;
    // partition kernel: part = part0 + #{g < ncmp : prefix > split[g]} with the tie rule folded into
    // the thresholds by the host (split[g] = s_g - incl_g; ncmp = 1, 3, 7 or 15, padded with ~0)
    uint32_t part0, ncmp;
    uint32_t lo_word, hi_word;  // record words holding prefix bits [31:0] / [63:32] (hi_word >= WORDS: none)
    uint32_t sign_word, sign_shift;
    uint64_t mask, xorv;
    uint64_t split[MAX_SPLITTERS];
};

struct PassParams {
    const uint32_t* src;
    uint32_t* dst;
    uint64_t n;                   // records
    DigitParams dp;
    const uint64_t* bucket_base;  // [256] exclusive prefix of this level's global histogram
    uint64_t* lookback;           // [num_tiles][256], zeroed once per sort
    uint32_t* ticket;             // zeroed before the launch
    uint32_t tag;                 // pass tag, 1..255
    uint32_t num_tiles;
    const uint8_t* remap;         // optional 256 -> part table (range partition); nullptr = identity
    const uint64_t* dst_tab;      // optional, with remap: [256] destination base pointer per part
                                  // (peer GPUs' buffers for the fused partition + exchange)
    SplitParams sp;               // remap kernels only: splitter mode when sp.nsplit != 0
    uint32_t hot[2];              // half-warp kernel: buckets matched by ballot (0xffffffff = none)
    uint32_t match_bits;          // half-warp remap kernel: match the part id by this many ballots (0 = OR-matching)
    uint32_t nparts;              // half-warp remap kernel with dst_tab: parts in use (0 = unknown); <= 16 parts are
                                  // written out part by part, the warps lined up on the destination's 128-byte lines
    // half-warp kernel, segmented mode (second MSD pass of the hybrid path): the array is 256 buckets, every
    // bucket is partitioned on its own (own tiles, own look-back chain, own bucket bases)
    const uint32_t* seg_tile_start;  // [257] first global tile id of every bucket; nullptr = one array
    const uint64_t* seg_rec_start;   // [257] first record of every bucket
    const uint64_t* seg_bases;       // [256][256] destination base of (bucket, digit)
};

template <int WORDS>
struct Rec {
    uint32_t w[WORDS];
};

template <int WORDS>
__device__ __forceinline__ Rec<WORDS> load_rec(const uint32_t* base, uint32_t idx) {
    Rec<WORDS> r;
    if constexpr (WORDS == 1) {
        r.w[0] = base[idx];
    } else if constexpr (WORDS == 2) {
        uint2 v = reinterpret_cast<const uint2*>(base)[idx];
        r.w[0] = v.x, r.w[1] = v.y;
    } else if constexpr (WORDS == 4) {
        uint4 v = reinterpret_cast<const uint4*>(base)[idx];
        r.w[0] = v.x, r.w[1] = v.y, r.w[2] = v.z, r.w[3] = v.w;
    } else {
#pragma unroll
        for (int k = 0; k < WORDS; ++k) r.w[k] = base[idx * WORDS + k];
    }
    return r;
}
template <int WORDS>
__device__ __forceinline__ void store_rec(uint32_t* base, uint64_t idx, const Rec<WORDS>& r) {
    if constexpr (WORDS == 1) {
        base[idx] = r.w[0];
    } else if constexpr (WORDS == 2) {
        reinterpret_cast<uint2*>(base)[idx] = make_uint2(r.w[0], r.w[1]);
    } else if constexpr (WORDS == 4) {
        reinterpret_cast<uint4*>(base)[idx] = make_uint4(r.w[0], r.w[1], r.w[2], r.w[3]);
    } else {
#pragma unroll
        for (int k = 0; k < WORDS; ++k) base[idx * WORDS + k] = r.w[k];
    }
}

template <int WORDS>
__device__ __forceinline__ uint32_t pick_word(const Rec<WORDS>& r, uint32_t word) {
    uint32_t v = r.w[0];
#pragma unroll
    for (int k = 1; k < WORDS; ++k)
        if (word == uint32_t(k)) v = r.w[k];
    return v;
}

template <int WORDS, bool FLOATK>
__device__ __forceinline__ uint32_t rec_digit(const DigitParams& dp, const Rec<WORDS>& r) {
    const uint32_t kw = pick_word<WORDS>(r, dp.word);
    const uint32_t sw = FLOATK ? pick_word<WORDS>(r, dp.sign_word) : 0u;
    return digit_from_words<FLOATK>(dp, kw, sw);
}

// WIN = true: bit-window digit (hybrid path), false: byte-aligned level
template <int WORDS, bool FLOATK, bool WIN>
__device__ __forceinline__ uint32_t rec_digit_w(const DigitParams& dp, const Rec<WORDS>& r) {
    if constexpr (WIN) {
        // a window that crosses a word boundary has its upper bits in the record's last word (the host
        // checks dp.word2 == WORDS - 1); one that does not never looks at the second operand
        const uint32_t kw = pick_word<WORDS>(r, dp.word);
        const uint32_t kw2 = r.w[WORDS - 1];
        const uint32_t sw = FLOATK ? pick_word<WORDS>(r, dp.sign_word) : 0u;
        return window_from_words<FLOATK>(dp, kw, kw2, sw);
    } else {
        return rec_digit<WORDS, FLOATK>(dp, r);
    }
}

template <int WORDS, bool FLOATK>
__device__ __forceinline__ uint64_t rec_prefix(const SplitParams& sp, const Rec<WORDS>& r) {
    const uint32_t lo = pick_word<WORDS>(r, sp.lo_word);
    const uint32_t hi = sp.hi_word < uint32_t(WORDS) ? pick_word<WORDS>(r, sp.hi_word) : 0u;
    uint64_t k = (uint64_t(hi) << 32) | lo;
    if (FLOATK) {
        const uint32_t neg = (pick_word<WORDS>(r, sp.sign_word) >> sp.sign_shift) & 1u;
        k ^= neg ? ~0ull : sp.xorv;
    } else {
        k ^= sp.xorv;
    }
    return k & sp.mask;
}

template <int N>
__device__ __forceinline__ uint32_t count_above(const SplitParams& sp, uint64_t k) {
    uint32_t c = 0;
#pragma unroll
    for (int g = 0; g < N; ++g) c += k > sp.split[g] ? 1u : 0u;
    return c;
}
__device__ __forceinline__ uint32_t prefix_part(const SplitParams& sp, uint64_t k) {
    if (sp.ncmp <= 1) return sp.part0 + count_above<1>(sp, k);
    if (sp.ncmp <= 3) return sp.part0 + count_above<3>(sp, k);
    if (sp.ncmp <= 7) return sp.part0 + count_above<7>(sp, k);
    return sp.part0 + count_above<MAX_SPLITTERS>(sp, k);
}

// The bucket a record goes to in this pass: its digit, the digit's part (range partition on
// one level) or the part its key prefix falls in (range partition on splitters).
template <int WORDS, bool FLOATK, bool REMAP>
__device__ __forceinline__ uint32_t pass_bucket(const PassParams& p, const DigitParams& dp, const Rec<WORDS>& r) {
    if constexpr (REMAP) {
        if (p.sp.nsplit) return prefix_part(p.sp, rec_prefix<WORDS, FLOATK>(p.sp, r));
    }
    uint32_t d = rec_digit<WORDS, FLOATK>(dp, r);
    if (REMAP) d = p.remap[d];
    return d;
}

template <int WORDS, int THREADS, int ITEMS, bool TMA>
struct PassSmem {
    static constexpr int WARPS = THREADS / 32;
    static constexpr int TILE = THREADS * ITEMS;
    static constexpr size_t OUT_BYTES = size_t(TILE) * WORDS * 4;
    static constexpr size_t IN_BYTES = TMA ? OUT_BYTES : 0;
    static constexpr size_t HIST_BYTES = size_t(WARPS) * RADIX * 4;
    static constexpr size_t GOFS_BYTES = RADIX * 8;
    static constexpr size_t MISC_BYTES = 128;
    static constexpr size_t TOTAL = OUT_BYTES + IN_BYTES + HIST_BYTES + GOFS_BYTES + MISC_BYTES;
};

template <int WORDS, int THREADS, int ITEMS, bool FLOATK, bool TMA, bool MATCH_HW, bool REMAP, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) onesweep_pass_kernel(const PassParams p) {
    static_assert(THREADS >= RADIX && THREADS % 32 == 0, "one thread per digit is needed for the look-back");
    using SM = PassSmem<WORDS, THREADS, ITEMS, TMA>;
    constexpr int WARPS = SM::WARPS;
    constexpr int TILE = SM::TILE;

    extern __shared__ __align__(128) uint8_t smem_raw[];
    uint32_t* s_out = reinterpret_cast<uint32_t*>(smem_raw);
    uint32_t* s_in = reinterpret_cast<uint32_t*>(smem_raw + SM::OUT_BYTES);
    uint32_t* s_hist = reinterpret_cast<uint32_t*>(smem_raw + SM::OUT_BYTES + SM::IN_BYTES);
    uint64_t* s_gofs = reinterpret_cast<uint64_t*>(smem_raw + SM::OUT_BYTES + SM::IN_BYTES + SM::HIST_BYTES);
    uint8_t* misc = smem_raw + SM::OUT_BYTES + SM::IN_BYTES + SM::HIST_BYTES + SM::GOFS_BYTES;
    uint64_t* mbar = reinterpret_cast<uint64_t*>(misc);                    // 8 B
    volatile uint32_t* s_next = reinterpret_cast<volatile uint32_t*>(misc + 16);
    uint32_t* s_wtot = reinterpret_cast<uint32_t*>(misc + 32);            // 8 warp totals

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint32_t* my_hist = s_hist + warp * RADIX;
    const uint32_t lt_mask = lanemask_lt();
    const DigitParams dp = p.dp;

    auto tile_valid = [&](uint32_t t) -> uint32_t {
        const uint64_t rem = p.n - uint64_t(t) * TILE;
        return rem < uint64_t(TILE) ? uint32_t(rem) : uint32_t(TILE);
    };
    auto claim_and_prefetch = [&]() {
        const uint32_t t = atomicAdd(p.ticket, 1u);
        *s_next = t;
        if (TMA && t < p.num_tiles && tile_valid(t) == uint32_t(TILE)) {
            mbar_expect_tx(mbar, uint32_t(SM::OUT_BYTES));
            bulk_g2s(s_in, p.src + uint64_t(t) * TILE * WORDS, uint32_t(SM::OUT_BYTES), mbar);
        }
    };

    if (tid == 0) {
        if (TMA) {
            mbar_init(mbar, 1);
            mbar_fence_init();
        }
        claim_and_prefetch();
    }
    __syncthreads();
    uint32_t phase = 0;

    while (true) {
        const uint32_t tile = *s_next;
        if (tile >= p.num_tiles) break;
        const uint64_t tile_base = uint64_t(tile) * TILE;
        const uint32_t valid = tile_valid(tile);
        const bool full = valid == uint32_t(TILE);

#pragma unroll
        for (int i = lane; i < RADIX; i += 32) my_hist[i] = 0;

        // ---- load: element (warp, j, lane) = warp*32*ITEMS + j*32 + lane (tile order) ----
        Rec<WORDS> r[ITEMS];
        uint32_t dig[ITEMS];
        const uint32_t e0 = uint32_t(warp) * 32u * ITEMS + lane;
        if (TMA) {
            if (full) {
                mbar_wait(mbar, phase);
                phase ^= 1u;
            } else {
                const uint32_t* g = p.src + tile_base * WORDS;
                for (uint32_t i = tid; i < valid * WORDS; i += THREADS) s_in[i] = g[i];
                __syncthreads();
            }
#pragma unroll
            for (int j = 0; j < ITEMS; ++j) {
                const uint32_t e = e0 + j * 32;
                if (full || e < valid) {
                    r[j] = load_rec<WORDS>(s_in, e);
                    dig[j] = rec_digit<WORDS, FLOATK>(dp, r[j]);
                    if (REMAP) dig[j] = p.remap[dig[j]];
                } else {
#pragma unroll
                    for (int k = 0; k < WORDS; ++k) r[j].w[k] = 0;
                    dig[j] = 255u;
                }
            }
        } else {
            const uint32_t* g = p.src + tile_base * WORDS;
#pragma unroll
            for (int j = 0; j < ITEMS; ++j) {
                const uint32_t e = e0 + j * 32;
                if (full || e < valid) {
                    r[j] = load_rec<WORDS>(g, e);
                } else {
#pragma unroll
                    for (int k = 0; k < WORDS; ++k) r[j].w[k] = 0;
                }
            }
#pragma unroll
            for (int j = 0; j < ITEMS; ++j) {
                const uint32_t e = e0 + j * 32;
                if (full || e < valid) {
                    dig[j] = rec_digit<WORDS, FLOATK>(dp, r[j]);
                    if (REMAP) dig[j] = p.remap[dig[j]];
                } else {
                    dig[j] = 255u;
                }
            }
        }
        __syncthreads();  // (A) s_in consumed, previous tile's write-out finished everywhere

        if (tid == 0) claim_and_prefetch();

        // ---- rank inside the warp: stable, lanes in order, rounds in order ----
        uint32_t rank[ITEMS];
#pragma unroll
        for (int j = 0; j < ITEMS; ++j) {
            const uint32_t d = dig[j];
            const uint32_t m = match_digit<MATCH_HW>(d);
            const uint32_t lt = m & lt_mask;
            const uint32_t before = my_hist[d];
            __syncwarp();
            if (lt == 0) my_hist[d] = before + __popc(m);
            __syncwarp();
            rank[j] = before + __popc(lt);
        }
        __syncthreads();  // (B) all warp histograms final

        // ---- digits: exclusive scan over warps, publish aggregate, scan over digits ----
        uint32_t cnt = 0, incl = 0, total = 0;
        if (tid < RADIX) {
            uint32_t sum = 0;
#pragma unroll
            for (int w = 0; w < WARPS; ++w) {
                const uint32_t c = s_hist[w * RADIX + tid];
                s_hist[w * RADIX + tid] = sum;
                sum += c;
            }
            total = sum;
            cnt = sum;
            if (tid == RADIX - 1) cnt -= uint32_t(TILE) - valid;  // padding sits at the end of bucket 255
            lb_store(&p.lookback[uint64_t(tile) * RADIX + tid],
                     lb_pack(tile == 0 ? LB_PREFIX : LB_AGG, p.tag, cnt));
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
            local_off = wbase + incl - total;
#pragma unroll
            for (int w = 0; w < WARPS; ++w) s_hist[w * RADIX + tid] += local_off;
        }
        __syncthreads();  // (C) per-(warp, digit) local bases final

        // ---- scatter into the staging buffer in locally sorted order ----
#pragma unroll
        for (int j = 0; j < ITEMS; ++j) {
            const uint32_t pos = my_hist[dig[j]] + rank[j];
            store_rec<WORDS>(s_out, pos, r[j]);
        }

        // ---- decoupled look-back, one thread per digit ----
        if (tid < RADIX) {
            uint64_t excl = 0;
            if (tile > 0) {
                int64_t t = int64_t(tile) - 1;
                while (true) {
                    const uint64_t v = lb_load(&p.lookback[uint64_t(t) * RADIX + tid]);
                    const uint64_t st = v >> 62;
                    if (st == 0 || uint32_t((v >> 54) & 0xffu) != (p.tag & 0xffu)) continue;
                    excl += v & LB_VALUE_MASK;
                    if (st == LB_PREFIX) break;
                    --t;
                }
                lb_store(&p.lookback[uint64_t(tile) * RADIX + tid], lb_pack(LB_PREFIX, p.tag, excl + cnt));
            }
            s_gofs[tid] = p.bucket_base[tid] + excl - local_off;
        }
        __syncthreads();  // (D)

        // ---- write-out: runs of equal digits are contiguous in s_out and in dst ----
        if constexpr (WORDS == 1 || WORDS == 2 || WORDS == 4) {
            if (full) {
#pragma unroll
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
            // word-granular so that odd record sizes still store fully coalesced
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
        // no barrier here: barrier (A) of the next tile orders this write-out
        // before the next scatter into s_out / s_gofs update.
    }
}

// ---------------------------------------------------------------------------
// Byte-granular pass for record sizes that are not a multiple of 4 (u8, u16,
// [u8; N], packed structs).  Same ranking and look-back; records are copied
// byte by byte straight to dst.  Correctness path, not a tuned one.
// ---------------------------------------------------------------------------
struct BytePassParams {
    const uint8_t* src;
    uint8_t* dst;
    uint64_t n;
    uint32_t elem_size;
    uint32_t byte_offset, xor_mask;
    uint32_t is_float, sign_byte;
    const uint64_t* bucket_base;
    uint64_t* lookback;
    uint32_t* ticket;
    uint32_t tag;
    uint32_t num_tiles;
    const uint8_t* remap;
};

__device__ __forceinline__ uint32_t byte_digit(const uint8_t* rec, uint32_t byte_offset, uint32_t xor_mask,
                                               uint32_t is_float, uint32_t sign_byte) {
    uint32_t b = rec[byte_offset];
    if (is_float && (rec[sign_byte] & 0x80u)) return b ^ 0xffu;
    return b ^ xor_mask;
}

template <int THREADS, int ITEMS>
__global__ void __launch_bounds__(THREADS) onesweep_pass_bytes_kernel(const BytePassParams p) {
    constexpr int WARPS = THREADS / 32;
    constexpr int TILE = THREADS * ITEMS;
    static_assert(THREADS == RADIX, "byte path uses one thread per digit");
    __shared__ uint32_t s_hist[WARPS * RADIX];
    __shared__ uint64_t s_gbase[RADIX];
    __shared__ uint32_t s_wtot[WARPS];
    __shared__ uint32_t s_tile;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint32_t* my_hist = s_hist + warp * RADIX;
    const uint32_t lt_mask = lanemask_lt();

    while (true) {
        __syncthreads();
        if (tid == 0) s_tile = atomicAdd(p.ticket, 1u);
        __syncthreads();
        const uint32_t tile = s_tile;
        if (tile >= p.num_tiles) break;
        const uint64_t tile_base = uint64_t(tile) * TILE;
        const uint64_t rem = p.n - tile_base;
        const uint32_t valid = rem < uint64_t(TILE) ? uint32_t(rem) : uint32_t(TILE);
        for (int i = lane; i < RADIX; i += 32) my_hist[i] = 0;
        __syncwarp();

        uint32_t dig[ITEMS], rank[ITEMS];
        const uint32_t e0 = uint32_t(warp) * 32u * ITEMS + lane;
#pragma unroll
        for (int j = 0; j < ITEMS; ++j) {
            const uint32_t e = e0 + j * 32;
            uint32_t d = 255u;
            if (e < valid) {
                d = byte_digit(p.src + (tile_base + e) * p.elem_size, p.byte_offset, p.xor_mask, p.is_float,
                               p.sign_byte);
                if (p.remap) d = p.remap[d];
            }
            dig[j] = d;
            const uint32_t m = match_digit<false>(d);
            const uint32_t lt = m & lt_mask;
            const uint32_t before = my_hist[d];
            __syncwarp();
            if (lt == 0) my_hist[d] = before + __popc(m);
            __syncwarp();
            rank[j] = before + __popc(lt);
        }
        __syncthreads();
        // exclusive scan over warps for my digit; counts per (warp, digit) become
        // "records of this digit in earlier warps"
        uint32_t sum = 0;
#pragma unroll
        for (int w = 0; w < WARPS; ++w) {
            const uint32_t c = s_hist[w * RADIX + tid];
            s_hist[w * RADIX + tid] = sum;
            sum += c;
        }
        uint32_t cnt = sum;
        if (tid == RADIX - 1) cnt -= uint32_t(TILE) - valid;
        lb_store(&p.lookback[uint64_t(tile) * RADIX + tid], lb_pack(tile == 0 ? LB_PREFIX : LB_AGG, p.tag, cnt));
        uint64_t excl = 0;
        if (tile > 0) {
            int64_t t = int64_t(tile) - 1;
            while (true) {
                const uint64_t v = lb_load(&p.lookback[uint64_t(t) * RADIX + tid]);
                const uint64_t st = v >> 62;
                if (st == 0 || uint32_t((v >> 54) & 0xffu) != (p.tag & 0xffu)) continue;
                excl += v & LB_VALUE_MASK;
                if (st == LB_PREFIX) break;
                --t;
            }
            lb_store(&p.lookback[uint64_t(tile) * RADIX + tid], lb_pack(LB_PREFIX, p.tag, excl + cnt));
        }
        s_gbase[tid] = p.bucket_base[tid] + excl;
        (void)s_wtot;
        __syncthreads();
#pragma unroll
        for (int j = 0; j < ITEMS; ++j) {
            const uint32_t e = e0 + j * 32;
            if (e < valid) {
                const uint64_t g = s_gbase[dig[j]] + my_hist[dig[j]] + rank[j];
                const uint8_t* s = p.src + (tile_base + e) * p.elem_size;
                uint8_t* d = p.dst + g * p.elem_size;
                for (uint32_t b = 0; b < p.elem_size; ++b) d[b] = s[b];
            }
        }
    }
}

}  // namespace b200
