// hist.cuh — K1, the one-read multi-level histogram, and the tiny bucket-base scan.
//
// K1 replaces every counting call rdst makes before and between its passes
//   get_counts / get_counts_with_ends / par_get_counts_with_ends
//   get_tile_counts + aggregate_tile_counts      (reference src/sort_utils.rs:33-249)
// with ONE read of the array that fills all LEVELS 256-bin histograms
// (u64 counts, as rdst's usize counts: src/radix_array.rs:4).
// The scan kernel is get_prefix_sums (reference src/sort_utils.rs:9-20) for
// every level at once.
#pragma once
#include "common.cuh"

namespace b200 {

constexpr int HIST_THREADS = 512;

struct HistParams {
    const uint32_t* src;
    uint64_t n;              // records
    uint32_t kw0;            // first key word inside the record
    uint32_t sign_shift;     // FLOAT: bit of the sign inside the last key word
    uint32_t xorw[4];        // per-level xor masks packed as key words (byte l of the key ^= mask[l])
    unsigned long long* ghist;  // [L][256], zeroed before the launch
    const uint32_t* gate;       // nullable: K0's out-of-order flag; 0 there = input already sorted, nothing to count
};

// Bank-rotated counting.  The CTA keeps 32 private 256-bin tables, table b living entirely
// in shared-memory bank b (counter (d, b) at word d*32 + b).  Table b = level * NCOPY + copy
// with NCOPY = 32 / LP copies of each level (LP = LEVELS rounded up to a power of two).
// In step s lane i counts level (i + s) % LP of ITS OWN record into copy i / LP, so the 32
// lanes of every shared atomic hit 32 different banks whatever the digits are: one
// wavefront per instruction on uniform, skewed and constant levels alike.  (A table indexed
// by the digit alone costs 3.3 wavefronts per instruction on random digits and serialises
// on constant ones: profiles/r1g_ncu_summary.txt.)
template <int L>
struct HistRot {
    static constexpr int LP = L <= 1 ? 1 : L <= 2 ? 2 : L <= 4 ? 4 : L <= 8 ? 8 : 16;
    static constexpr int NCOPY = 32 / LP;
};

// Per-lane constants of the rotation.  Levels are walked word by word so that one PRMT
// with a zero second operand yields the zero-extended digit: in step s the lane counts
// level  4 * ((wsel + s / 4) % NKWP) + (lvl_lo + s) % 4  (for LEVELS < 4: (lvl_lo + s) % LP),
// which for a fixed s is a bijection from the LP lanes of a copy group onto the levels.
template <int L>
struct HistLane {
    static constexpr int LP = HistRot<L>::LP, NCOPY = HistRot<L>::NCOPY;
    static constexpr int LOWN = LP < 4 ? LP : 4;   // levels walked inside one word
    static constexpr int NKWP = LP < 4 ? 1 : LP / 4;  // key words (rounded up to a power of two)
    uint32_t sel[LOWN];   // PRMT selectors: digit byte -> bits [7:0], zeros above
    uint32_t addr[LP];    // shared address of (digit 0, level(s), my copy)
    uint32_t lvl[LP];     // level counted in step s (only read when LP != L)
    uint32_t wsel;
    __device__ __forceinline__ void init(uint32_t smem_base, uint32_t lane) {
        const uint32_t lvl_lo = lane & (LOWN - 1);
        wsel = (lane / LOWN) & (NKWP - 1);
        const uint32_t copy = lane / LP;
#pragma unroll
        for (int s = 0; s < LOWN; ++s) sel[s] = ((lvl_lo + s) & (LOWN - 1)) | 0x4440u;
#pragma unroll
        for (int s = 0; s < LP; ++s) {
            const uint32_t level = 4 * ((wsel + s / 4) & (NKWP - 1)) + ((lvl_lo + s) & (LOWN - 1));
            lvl[s] = level;
            addr[s] = smem_base + (level * NCOPY + copy) * 4;
        }
    }
};

__device__ __forceinline__ void red_shared_inc(uint32_t saddr, bool pred) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.u32 p, %1, 0;\n"
        "@p red.shared.add.u32 [%0], 1;\n"
        "}\n" ::"r"(saddr),
        "r"(uint32_t(pred))
        : "memory");
}

__device__ __forceinline__ void red_shared_inc_all(uint32_t saddr) {
    asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(saddr) : "memory");
}

// CHECK = false: every lane's record is valid and LEVELS is a power of two, so the shared
// atomics carry no predicate (ptxas turns a predicated one into a branch around it).
template <int L, bool FLOATK, bool CHECK>
__device__ __forceinline__ void hist_record(const HistLane<L>& hl, const HistParams& hp, const uint32_t* kraw,
                                            bool active) {
    constexpr int NKW = (L + 3) / 4;
    constexpr int LP = HistRot<L>::LP;
    constexpr int LOWN = HistLane<L>::LOWN, NKWP = HistLane<L>::NKWP;
    uint32_t k[NKWP];
    uint32_t neg = 0;
    if (FLOATK) neg = (kraw[NKW - 1] >> hp.sign_shift) & 1u;
#pragma unroll
    for (int c = 0; c < NKWP; ++c) k[c] = c < NKW ? (kraw[c] ^ ((FLOATK && neg) ? 0xffffffffu : hp.xorw[c])) : 0u;
#pragma unroll
    for (int g = 0; g < NKWP; ++g) {
        // the word this lane walks in steps 4g .. 4g+3
        uint32_t w;
        if constexpr (NKWP == 1) {
            w = k[0];
        } else if constexpr (NKWP == 2) {
            w = ((hl.wsel + g) & 1u) ? k[1] : k[0];
        } else {
            const uint32_t wi = (hl.wsel + g) & 3u;
            const uint32_t a = (wi & 1u) ? k[1] : k[0], b = (wi & 1u) ? k[3] : k[2];
            w = (wi & 2u) ? b : a;
        }
#pragma unroll
        for (int q = 0; q < LOWN; ++q) {
            const int s = g * LOWN + q;
            const uint32_t d = __byte_perm(w, 0u, hl.sel[q]);
            if constexpr (!CHECK && LP == L) {
                red_shared_inc_all(hl.addr[s] + d * 128u);
            } else {
                const bool on = active && (LP == L || hl.lvl[s] < uint32_t(L));
                red_shared_inc(hl.addr[s] + d * 128u, on);
            }
        }
    }
}

// WORDS = record size in words.  PACKED = the record is exactly the key words
// (key-only arrays): 128-bit loads, several records per load.
template <int WORDS, int L, bool FLOATK>
__global__ void __launch_bounds__(HIST_THREADS) hist_fast_kernel(const HistParams hp) {
    constexpr int NKW = (L + 3) / 4;
    constexpr int NCOPY = HistRot<L>::NCOPY;
    constexpr bool PACKED = (WORDS == NKW) && (WORDS == 1 || WORDS == 2 || WORDS == 4);
    __shared__ uint32_t sh[32 * RADIX];
    if (hp.gate && *hp.gate == 0) return;  // K0 ran in front on the same stream and found the input in order
    for (int i = threadIdx.x; i < 32 * RADIX; i += HIST_THREADS) sh[i] = 0;
    __syncthreads();

    const uint32_t lane = threadIdx.x & 31;
    HistLane<L> hl;
    hl.init(smem_u32(sh), lane);
    const uint64_t gtid = uint64_t(blockIdx.x) * HIST_THREADS + threadIdx.x;
    const uint64_t gstride = uint64_t(gridDim.x) * HIST_THREADS;

    if constexpr (PACKED) {
        constexpr int RPV = 4 / WORDS;  // records per 16-byte vector
        const uint64_t nvec = hp.n / RPV;
        const uint4* v4 = reinterpret_cast<const uint4*>(hp.src);
        constexpr int UNROLL = 4;
        uint64_t base = gtid - lane;
        // every vector of the warp's UNROLL loads in range: no predicates at all
        for (; base + 31 + (UNROLL - 1) * gstride < nvec; base += gstride * UNROLL) {
            uint4 q[UNROLL];
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) q[u] = ldg_stream(v4 + base + lane + u * gstride);
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) {
                const uint32_t w[4] = {q[u].x, q[u].y, q[u].z, q[u].w};
#pragma unroll
                for (int rr = 0; rr < RPV; ++rr) hist_record<L, FLOATK, false>(hl, hp, w + rr * WORDS, true);
            }
        }
        for (; base < nvec; base += gstride) {
            const uint64_t v = base + lane;
            const bool ok = v < nvec;
            const uint4 q = ok ? ldg_stream(v4 + v) : make_uint4(0, 0, 0, 0);
            const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
            for (int rr = 0; rr < RPV; ++rr) hist_record<L, FLOATK, true>(hl, hp, w + rr * WORDS, ok);
        }
        // tail records (n % RPV), first warp of the grid
        if (blockIdx.x == 0 && threadIdx.x < 32) {
            const uint64_t rec = nvec * RPV + threadIdx.x;
            const bool ok = rec < hp.n;
            uint32_t k[NKW];
#pragma unroll
            for (int c = 0; c < NKW; ++c) k[c] = ok ? hp.src[rec * WORDS + c] : 0u;
            hist_record<L, FLOATK, true>(hl, hp, k, ok);
        }
    } else {
        constexpr int UNROLL = 4;
        uint64_t base = gtid - lane;
        for (; base + 31 + (UNROLL - 1) * gstride < hp.n; base += gstride * UNROLL) {
            uint32_t k[UNROLL][NKW];
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) {
                const uint64_t rec = base + lane + u * gstride;
#pragma unroll
                for (int c = 0; c < NKW; ++c) k[u][c] = __ldg(hp.src + rec * WORDS + hp.kw0 + c);
            }
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) hist_record<L, FLOATK, false>(hl, hp, k[u], true);
        }
        for (; base < hp.n; base += gstride) {
            const uint64_t rec = base + lane;
            const bool ok = rec < hp.n;
            uint32_t k[NKW];
#pragma unroll
            for (int c = 0; c < NKW; ++c) k[c] = ok ? __ldg(hp.src + rec * WORDS + hp.kw0 + c) : 0u;
            hist_record<L, FLOATK, true>(hl, hp, k, ok);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < L * RADIX; i += HIST_THREADS) {
        const int l = i / RADIX, d = i % RADIX;
        uint32_t c = 0;
#pragma unroll
        for (int q = 0; q < NCOPY; ++q) c += sh[d * 32 + l * NCOPY + q];
        if (c) atomicAdd(hp.ghist + i, (unsigned long long)c);
    }
}

// Any record size, any byte-offset table (BYTES kind, partial float keys,
// unaligned keys): one byte load per level.
struct HistGenericParams {
    const uint8_t* src;
    uint64_t n;
    uint32_t elem_size, levels;
    uint32_t is_float, sign_byte;
    unsigned long long* ghist;
    const uint32_t* gate;  // as HistParams::gate
    LevelLowering lv[MAX_LEVELS];
};

static __global__ void __launch_bounds__(HIST_THREADS) hist_generic_kernel(const HistGenericParams hp) {
    extern __shared__ uint32_t shg[];
    if (hp.gate && *hp.gate == 0) return;
    for (uint32_t i = threadIdx.x; i < hp.levels * RADIX; i += HIST_THREADS) shg[i] = 0;
    __syncthreads();
    const uint64_t gstride = uint64_t(gridDim.x) * HIST_THREADS;
    for (uint64_t rec = uint64_t(blockIdx.x) * HIST_THREADS + threadIdx.x; rec < hp.n; rec += gstride) {
        const uint8_t* r = hp.src + rec * hp.elem_size;
        const bool neg = hp.is_float && (r[hp.sign_byte] & 0x80u);
        for (uint32_t l = 0; l < hp.levels; ++l) {
            uint32_t d = r[hp.lv[l].byte_offset];
            d ^= neg ? 0xffu : hp.lv[l].xor_mask;
            atomicAdd(&shg[l * RADIX + d], 1u);
        }
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < hp.levels * RADIX; i += HIST_THREADS) {
        const uint32_t c = shg[i];
        if (c) atomicAdd(hp.ghist + i, (unsigned long long)c);
    }
}

// One block per level: exclusive prefix sums of the 256 counts.
static __global__ void __launch_bounds__(RADIX) scan_hist_kernel(const unsigned long long* ghist,
                                                         unsigned long long* gbase) {
    __shared__ unsigned long long wtot[RADIX / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned long long c = ghist[blockIdx.x * RADIX + tid];
    unsigned long long incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long y = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += y;
    }
    if (lane == 31) wtot[warp] = incl;
    __syncthreads();
    unsigned long long wbase = 0;
    for (int w = 0; w < warp; ++w) wbase += wtot[w];
    gbase[blockIdx.x * RADIX + tid] = wbase + incl - c;
}

}  // namespace b200
