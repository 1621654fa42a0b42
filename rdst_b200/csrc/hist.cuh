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
    uint32_t xor_mask[16];   // per level
    unsigned long long* ghist;  // [L][256], zeroed before the launch
};

// Adds one digit to a CTA-private shared histogram.  With AGG, a warp whose 32
// lanes all carry the same digit (constant or heavily skewed levels) sends one
// add of 32 instead of a 32-way same-address conflict.
template <bool AGG>
__device__ __forceinline__ void hist_add(uint32_t* sh_level, uint32_t d, bool active) {
    if (AGG) {
        const uint32_t act = __ballot_sync(0xffffffffu, active);
        const uint32_t d0 = __shfl_sync(0xffffffffu, d, __ffs(act) - 1);
        const bool same = __all_sync(0xffffffffu, !active || d == d0);
        if (same) {
            if (active && (threadIdx.x & 31) == __ffs(act) - 1) atomicAdd(&sh_level[d], __popc(act));
            return;
        }
    }
    if (active) atomicAdd(&sh_level[d], 1u);
}

// Digits of one record whose key is NKW consecutive words k[0..NKW), level l =
// byte l of the key (UINT / SINT / FLOAT kinds, src/radix_key_impl.rs:3-185).
template <int L, bool FLOATK, bool AGG>
__device__ __forceinline__ void hist_record(uint32_t* sh, const HistParams& hp, const uint32_t* k, bool active) {
    constexpr int NKW = (L + 3) / 4;
    uint32_t neg = 0;
    if (FLOATK) neg = (k[NKW - 1] >> hp.sign_shift) & 1u;
#pragma unroll
    for (int l = 0; l < L; ++l) {
        uint32_t d = (k[l >> 2] >> (8 * (l & 3))) & 0xffu;
        if (FLOATK)
            d ^= neg ? 0xffu : hp.xor_mask[l];
        else
            d ^= hp.xor_mask[l];
        hist_add<AGG>(sh + l * RADIX, d, active);
    }
}

// WORDS = record size in words (0 = runtime hp-provided stride is not supported
// here; see hist_generic_kernel).  PACKED = the record is exactly the key words
// (key-only arrays): 128-bit loads, several records per load.
template <int WORDS, int L, bool FLOATK, bool AGG>
__global__ void __launch_bounds__(HIST_THREADS) hist_fast_kernel(const HistParams hp) {
    constexpr int NKW = (L + 3) / 4;
    constexpr bool PACKED = (WORDS == NKW) && (WORDS == 1 || WORDS == 2 || WORDS == 4);
    __shared__ uint32_t sh[L * RADIX];
    for (int i = threadIdx.x; i < L * RADIX; i += HIST_THREADS) sh[i] = 0;
    __syncthreads();

    const uint64_t gtid = uint64_t(blockIdx.x) * HIST_THREADS + threadIdx.x;
    const uint64_t gstride = uint64_t(gridDim.x) * HIST_THREADS;

    if constexpr (PACKED) {
        constexpr int RPV = 4 / WORDS;  // records per 16-byte vector
        const uint64_t nvec = hp.n / RPV;
        const uint4* v4 = reinterpret_cast<const uint4*>(hp.src);
        constexpr int UNROLL = 4;
        // whole warps stay converged: the loop bound is rounded up per warp
        for (uint64_t base = gtid - (threadIdx.x & 31); base < nvec; base += gstride * UNROLL) {
            uint4 q[UNROLL];
            bool ok[UNROLL];
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) {
                const uint64_t v = base + (threadIdx.x & 31) + u * gstride;
                ok[u] = v < nvec;
                q[u] = ok[u] ? ldg_stream(v4 + v) : make_uint4(0, 0, 0, 0);
            }
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) {
                const uint32_t w[4] = {q[u].x, q[u].y, q[u].z, q[u].w};
#pragma unroll
                for (int rr = 0; rr < RPV; ++rr) hist_record<L, FLOATK, AGG>(sh, hp, w + rr * WORDS, ok[u]);
            }
        }
        // tail records (n % RPV), first warp of the grid
        if (blockIdx.x == 0 && threadIdx.x < 32) {
            const uint64_t rec = nvec * RPV + threadIdx.x;
            const bool ok = rec < hp.n;
            uint32_t k[NKW];
#pragma unroll
            for (int c = 0; c < NKW; ++c) k[c] = ok ? hp.src[rec * WORDS + c] : 0u;
            hist_record<L, FLOATK, AGG>(sh, hp, k, ok);
        }
    } else {
        constexpr int UNROLL = 4;
        for (uint64_t base = gtid - (threadIdx.x & 31); base < hp.n; base += gstride * UNROLL) {
            uint32_t k[UNROLL][NKW];
            bool ok[UNROLL];
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) {
                const uint64_t rec = base + (threadIdx.x & 31) + u * gstride;
                ok[u] = rec < hp.n;
#pragma unroll
                for (int c = 0; c < NKW; ++c) k[u][c] = ok[u] ? __ldg(hp.src + rec * WORDS + hp.kw0 + c) : 0u;
            }
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) hist_record<L, FLOATK, AGG>(sh, hp, k[u], ok[u]);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < L * RADIX; i += HIST_THREADS) {
        const uint32_t c = sh[i];
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
    LevelLowering lv[MAX_LEVELS];
};

static __global__ void __launch_bounds__(HIST_THREADS) hist_generic_kernel(const HistGenericParams hp) {
    extern __shared__ uint32_t shg[];
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
