// sorted_check.cuh — K0, the already-sorted early-out.
//
// rdst scans every chunk while its digits are still non-decreasing before it counts them
// (reference src/sort_utils.rs:125-136) and skips the sort of a level whose digits already arrive
// in order (src/sorter.rs:59-65); a fully presorted input therefore costs it counting passes only
// (the `struct_sort` bench's "fully presorted" case, benches/struct_sort.rs:43-74).  The GPU
// counterpart checks the whole key once: one read of the array (n*S bytes, HBM bound), and a
// stable sort of an input that is already non-decreasing by its key is the identity, so the
// engine returns without running K1/K2/K3.
//
// Every thread compares adjacent records i and i+1 by the order-preserving image of their keys.
// The first violation raises a global flag; CTAs read the flag once per chunk and leave as soon as
// it is up, and all of them probe the head of the array first, so an unsorted input costs a few
// microseconds, not a pass.
#pragma once
#include "common.cuh"

namespace b200 {

constexpr int SC_THREADS = 256, SC_ITEMS = 4, SC_VEC_ITEMS = 4;

struct SortedCheckParams {
    const uint8_t* data;
    uint64_t n_pairs;  // n - 1: pair i is (record i, record i + 1)
    uint32_t* flag;    // zeroed before the launch; 1 = some pair is out of order
    uint32_t elem_size;
    uint32_t levels;
    // word path: the key is an aligned little-endian word of W bytes at key_offset
    uint32_t key_offset;
    uint32_t is_float;
    uint64_t flip;  // xor applied to the word (SINT: the sign bit; FLOAT non-negative: the sign bit)
    uint64_t keep;  // the low `levels` bytes
    // byte path: one digit per level, most significant level compared first
    uint32_t sign_byte;
    LevelLowering lv[MAX_LEVELS];
};

template <int W>
__device__ __forceinline__ uint64_t sc_load_word(const uint8_t* p) {
    if (W == 8) return *reinterpret_cast<const uint64_t*>(p);
    if (W == 4) return *reinterpret_cast<const uint32_t*>(p);
    if (W == 2) return *reinterpret_cast<const uint16_t*>(p);
    return *p;
}

// Order-preserving unsigned image of the low `levels` bytes of the key word
// (reference src/radix_key_impl.rs:3-46 unsigned, :87-160 signed, :162-185 floats).
template <int W>
__device__ __forceinline__ uint64_t sc_image(const SortedCheckParams& p, uint64_t i) {
    uint64_t v = sc_load_word<W>(p.data + i * p.elem_size + p.key_offset);
    if (p.is_float && ((v >> (8 * W - 1)) & 1ull))
        v = ~v;
    else
        v ^= p.flip;
    return v & p.keep;
}

template <int W>
__device__ __forceinline__ uint64_t sc_xform(const SortedCheckParams& p, uint64_t v) {
    if (p.is_float && ((v >> (8 * W - 1)) & 1ull))
        v = ~v;
    else
        v ^= p.flip;
    return v & p.keep;
}

// key j of a 16-byte vector of W-byte keys
template <int W>
__device__ __forceinline__ uint64_t sc_unpack(const uint4& v, int j) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    if (W == 8) return uint64_t(w[2 * j]) | (uint64_t(w[2 * j + 1]) << 32);
    if (W == 4) return w[j];
    if (W == 2) return (w[j / 2] >> ((j & 1) * 16)) & 0xffffu;
    return (w[j / 4] >> ((j & 3) * 8)) & 0xffu;
}

// Flag protocol shared by both kernels.  Every CTA first checks the first chunk of the ARRAY (the
// same few KB for all of them, served by L2): unsorted data of any ordinary kind fails right there,
// every CTA sees it on its own and leaves, and CTA 0 alone writes the flag — no traffic on the flag's
// cache line at all (a thousand CTAs storing to and polling one line cost 0.2 ms on the two-die L2,
// measured; profiles/r3_k0.txt).  Past the probe a CTA reads the flag once per chunk (thread 0) and
// writes it at most once.
__device__ __forceinline__ bool sc_leave(bool bad, uint32_t raised, uint32_t* flag) {
    if (!__syncthreads_or(bad || raised != 0)) return false;
    if (__syncthreads_or(bad) && threadIdx.x == 0) *reinterpret_cast<volatile uint32_t*>(flag) = 1u;
    return true;
}

// Key-only arrays (elem_size == W, 16-byte aligned): 128-bit streaming loads, 16 / W keys per load.
// Pairs inside a vector are compared in registers; the pair that crosses into the next vector takes
// the neighbour lane's first key by shuffle (lane 31 reads it from memory: an L1 / L2 hit, the next
// warp loads the same line).  CTAs leave as a whole, so the shuffles always see 32 lanes.
// Thread 0 of CTA 0 checks the pairs past the last full vector.
template <int W>
__device__ __forceinline__ bool sc_vec_chunk_bad(const SortedCheckParams& p, uint64_t nvec, uint64_t base) {
    constexpr int K = 16 / W;
    const uint4* src = reinterpret_cast<const uint4*>(p.data);
    const int lane = threadIdx.x & 31;
    // all loads of the chunk in flight before the first compare (the kernel is latency bound; fetching
    // lane 31's neighbour key lazily inside the compare loop serialised an L2 round trip per step:
    // u64 3.7 -> 3.1 TB/s, profiles/r3_k0_small_ladder.md)
    uint4 v[SC_VEC_ITEMS];
    uint64_t nxt[SC_VEC_ITEMS];
#pragma unroll
    for (int k = 0; k < SC_VEC_ITEMS; ++k) {
        const uint64_t idx = base + uint64_t(k) * SC_THREADS + threadIdx.x;
        v[k] = make_uint4(0, 0, 0, 0);
        nxt[k] = 0;
        if (idx < nvec) v[k] = ldg_stream(src + idx);
        if (lane == 31 && idx + 1 < nvec) nxt[k] = sc_image<W>(p, (idx + 1) * K);  // the next warp's first key
    }
    bool bad = false;
#pragma unroll
    for (int k = 0; k < SC_VEC_ITEMS; ++k) {
        const uint64_t idx = base + uint64_t(k) * SC_THREADS + threadIdx.x;
        uint64_t img[K];
#pragma unroll
        for (int j = 0; j < K; ++j) img[j] = sc_xform<W>(p, sc_unpack<W>(v[k], j));
#pragma unroll
        for (int j = 0; j + 1 < K; ++j) bad |= img[j] > img[j + 1];  // an all-zero filler vector has equal images
        uint64_t nb = __shfl_down_sync(0xffffffffu, img[0], 1);
        if (lane == 31) nb = nxt[k];
        if (idx + 1 < nvec) bad |= img[K - 1] > nb;
    }
    return bad;
}

template <int W>
__global__ void __launch_bounds__(SC_THREADS) sorted_check_vec_kernel(const SortedCheckParams p, uint64_t nvec) {
    constexpr int K = 16 / W;
    constexpr uint64_t PER_ITER = uint64_t(SC_THREADS) * SC_VEC_ITEMS;
    if (__syncthreads_or(sc_vec_chunk_bad<W>(p, nvec, 0))) {  // the probe
        if (blockIdx.x == 0 && threadIdx.x == 0) *reinterpret_cast<volatile uint32_t*>(p.flag) = 1u;
        return;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        const uint64_t first = nvec ? nvec * K - 1 : 0;
        for (uint64_t i = first; i < p.n_pairs; ++i)
            if (sc_image<W>(p, i) > sc_image<W>(p, i + 1)) *reinterpret_cast<volatile uint32_t*>(p.flag) = 1u;
    }
    for (uint64_t base = uint64_t(blockIdx.x) * PER_ITER; base < nvec; base += uint64_t(gridDim.x) * PER_ITER) {
        uint32_t raised = 0;
        if (threadIdx.x == 0) raised = *reinterpret_cast<volatile const uint32_t*>(p.flag);
        const bool bad = sc_vec_chunk_bad<W>(p, nvec, base);
        if (sc_leave(bad, raised, p.flag)) return;
    }
}

// true when record a sorts strictly after record b (levels compared from the most significant down)
__device__ __forceinline__ bool sc_greater_bytes(const SortedCheckParams& p, uint64_t ia, uint64_t ib) {
    const uint8_t* a = p.data + ia * p.elem_size;
    const uint8_t* b = p.data + ib * p.elem_size;
    uint32_t na = 0, nb = 0;
    if (p.is_float) {
        na = (a[p.sign_byte] >> 7) ? 0xffu : 0u;
        nb = (b[p.sign_byte] >> 7) ? 0xffu : 0u;
    }
    for (int l = int(p.levels) - 1; l >= 0; --l) {
        const uint32_t off = p.lv[l].byte_offset, x = p.lv[l].xor_mask;
        const uint32_t da = uint32_t(a[off]) ^ (na ? na : x);
        const uint32_t db = uint32_t(b[off]) ^ (nb ? nb : x);
        if (da != db) return da > db;
    }
    return false;
}

// W = key word bytes (1, 2, 4, 8) for the word path, 0 for the byte path
template <int W>
__device__ __forceinline__ bool sc_chunk_bad(const SortedCheckParams& p, uint64_t base) {
    bool bad = false;
    if (W != 0) {
        uint64_t a[SC_ITEMS], b[SC_ITEMS];
#pragma unroll
        for (int k = 0; k < SC_ITEMS; ++k) {
            const uint64_t i = base + uint64_t(k) * SC_THREADS + threadIdx.x;
            a[k] = b[k] = 0;
            if (i < p.n_pairs) {
                a[k] = sc_image<(W ? W : 1)>(p, i);
                b[k] = sc_image<(W ? W : 1)>(p, i + 1);
            }
        }
#pragma unroll
        for (int k = 0; k < SC_ITEMS; ++k) bad |= a[k] > b[k];
    } else {
#pragma unroll 1
        for (int k = 0; k < SC_ITEMS; ++k) {
            const uint64_t i = base + uint64_t(k) * SC_THREADS + threadIdx.x;
            if (i < p.n_pairs) bad |= sc_greater_bytes(p, i, i + 1);
        }
    }
    return bad;
}

template <int W>
__global__ void __launch_bounds__(SC_THREADS) sorted_check_kernel(const SortedCheckParams p) {
    constexpr uint64_t PER_ITER = uint64_t(SC_THREADS) * SC_ITEMS;
    if (__syncthreads_or(sc_chunk_bad<W>(p, 0))) {  // the probe
        if (blockIdx.x == 0 && threadIdx.x == 0) *reinterpret_cast<volatile uint32_t*>(p.flag) = 1u;
        return;
    }
    for (uint64_t base = uint64_t(blockIdx.x) * PER_ITER; base < p.n_pairs; base += uint64_t(gridDim.x) * PER_ITER) {
        uint32_t raised = 0;
        if (threadIdx.x == 0) raised = *reinterpret_cast<volatile const uint32_t*>(p.flag);
        const bool bad = sc_chunk_bad<W>(p, base);
        if (sc_leave(bad, raised, p.flag)) return;
    }
}

}  // namespace b200
