// common.cuh — device-side building blocks shared by the sm_100a kernels of
// the B200 radix sort engine: digit extraction (the lowered form of rdst's
// RadixKey::get_level, reference src/radix_key.rs:1-5 and
// src/radix_key_impl.rs), look-back word packing, mbarrier / bulk-copy (TMA)
// wrappers and warp digit matching.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace b200 {

constexpr int RADIX = 256;
constexpr int MAX_LEVELS = 64;

// How one level's digit is pulled out of a record that is viewed as an array
// of little-endian 32-bit words:  digit = ((rec[word] >> shift) & 0xff) ^ x,
// x = xor_mask, or 0xff when the record's float key is negative
// (src/radix_key_impl.rs:162-185: negative floats flip every bit, others
// only the sign bit, i.e. only the top level's byte ^ 0x80).
struct LevelLowering {
    uint32_t byte_offset;  // byte inside the record
    uint32_t xor_mask;     // applied after the load
};

struct DigitParams {
    uint32_t word;       // 32-bit word index inside the record
    uint32_t shift;      // 0, 8, 16, 24
    uint32_t xor_mask;   // per level
    uint32_t sign_word;  // FLOAT keys: word holding the sign bit
    uint32_t sign_shift; // FLOAT keys: bit position of the sign inside that word
    // bit windows (hybrid path only; the byte-aligned kernels ignore these): the digit is
    // bits [shift, shift + width) of the 64-bit pair (record word `word2` : record word `word`)
    uint32_t word2;      // word holding the bits above bit 31 of the pair (== word when the window does not cross)
    uint32_t mask;       // (1 << width) - 1; xor_mask lies inside it
};

template <bool FLOATK>
__device__ __forceinline__ uint32_t digit_from_words(const DigitParams& dp, uint32_t key_word,
                                                     uint32_t sign_word_val) {
    uint32_t b = (key_word >> dp.shift) & 0xffu;
    if (FLOATK) {
        uint32_t neg = (sign_word_val >> dp.sign_shift) & 1u;
        return b ^ (neg ? 0xffu : dp.xor_mask);
    }
    return b ^ dp.xor_mask;
}

// Window digit: up to 8 bits at any bit position of the key image, possibly across two record words.
template <bool FLOATK>
__device__ __forceinline__ uint32_t window_from_words(const DigitParams& dp, uint32_t key_word, uint32_t key_word2,
                                                      uint32_t sign_word_val) {
    const uint32_t b = __funnelshift_r(key_word, key_word2, dp.shift) & dp.mask;
    if (FLOATK) {
        const uint32_t neg = (sign_word_val >> dp.sign_shift) & 1u;
        return b ^ (neg ? dp.mask : dp.xor_mask);
    }
    return b ^ dp.xor_mask;
}

// ---------------------------------------------------------------------------
// Decoupled look-back words.  One 64-bit word per (tile, digit):
//   [63:62] state (0 = not written, 1 = tile aggregate, 2 = inclusive prefix)
//   [61:54] pass tag (the array is zeroed once per sort and shared by all passes)
//   [53:0]  count
// A word is written and read with single 64-bit relaxed accesses, so state and
// count always travel together and no fence is needed.
// ---------------------------------------------------------------------------
constexpr uint64_t LB_AGG = 1ull, LB_PREFIX = 2ull;
constexpr uint64_t LB_VALUE_MASK = (1ull << 54) - 1;

__device__ __forceinline__ uint64_t lb_pack(uint64_t state, uint32_t tag, uint64_t value) {
    return (state << 62) | (uint64_t(tag & 0xffu) << 54) | value;
}
__device__ __forceinline__ uint64_t lb_load(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void lb_store(uint64_t* p, uint64_t v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// ---------------------------------------------------------------------------
// mbarrier + 1-D bulk asynchronous copy (TMA engine, SASS UBLKCP).
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {
    }
}
// global -> shared bulk copy; bytes and both addresses must be multiples of 16.
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(smem_dst)),
        "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

__device__ __forceinline__ uint32_t lanemask_lt() {
    uint32_t m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

// Lanes of the warp whose 8-bit digit equals this lane's digit.
// MATCH_HW = the MATCH.ANY instruction; otherwise eight ballots.
template <bool MATCH_HW>
__device__ __forceinline__ uint32_t match_digit(uint32_t d) {
    if (MATCH_HW) {
        return __match_any_sync(0xffffffffu, d);
    } else {
        uint32_t mask = 0xffffffffu;
#pragma unroll
        for (int b = 0; b < 8; ++b) {
            const bool bit = (d >> b) & 1u;
            const uint32_t bal = __ballot_sync(0xffffffffu, bit);
            mask &= bit ? bal : ~bal;
        }
        return mask;
    }
}

__device__ __forceinline__ uint4 ldg_stream(const uint4* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}

}  // namespace b200
