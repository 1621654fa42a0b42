// unstable_inst.cuh — instantiates K2-U and K3-U for one record width; included by unstable_w{1,2}.cu.
#pragma once
#include "local_count.cuh"
#include "pass_atomic.cuh"
#include "unstable.h"

namespace b200 {

template <int WORDS, bool FLOATK, int THREADS, int ITEMS, int MINB, bool VEC = false>
static MsdULaunch make_msdu(const char* name) {
    using SM = AtomSmem<WORDS, THREADS, ITEMS>;
    MsdULaunch l{};
    l.kernel = &msd_pass_atomic_kernel<WORDS, THREADS, ITEMS, FLOATK, MINB, VEC>;
    l.smem = SM::TOTAL;
    l.tile = SM::TILE;
    l.threads = THREADS;
    l.name = name;
    return l;
}

template <int WORDS, bool FLOATK>
static MsdULaunch pick_msdu(int variant) {
    if constexpr (WORDS == 1) {
        switch (variant) {
            case 0: return make_msdu<WORDS, FLOATK, 256, 32, 4>("u256_i32_occ4");
            case 1: return make_msdu<WORDS, FLOATK, 256, 32, 3>("u256_i32_occ3");
            case 2: return make_msdu<WORDS, FLOATK, 512, 24, 2>("u512_i24_occ2");
            case 3: return make_msdu<WORDS, FLOATK, 256, 16, 5>("u256_i16_occ5");
            case 4: return make_msdu<WORDS, FLOATK, 256, 32, 4, true>("v256_i32_occ4");
            case 5: return make_msdu<WORDS, FLOATK, 512, 24, 2, true>("v512_i24_occ2");
            case 6: return make_msdu<WORDS, FLOATK, 512, 16, 3, true>("v512_i16_occ3");
            case 7: return make_msdu<WORDS, FLOATK, 256, 16, 5, true>("v256_i16_occ5");
        }
    } else {
        switch (variant) {
            case 0: return make_msdu<WORDS, FLOATK, 256, 32, 2>("u256_i32_occ2");
            case 1: return make_msdu<WORDS, FLOATK, 256, 24, 3>("u256_i24_occ3");
            case 2: return make_msdu<WORDS, FLOATK, 512, 16, 2>("u512_i16_occ2");
            case 3: return make_msdu<WORDS, FLOATK, 256, 16, 4>("u256_i16_occ4");
            case 4: return make_msdu<WORDS, FLOATK, 256, 32, 2, true>("v256_i32_occ2");
            case 5: return make_msdu<WORDS, FLOATK, 256, 24, 3, true>("v256_i24_occ3");
            case 6: return make_msdu<WORDS, FLOATK, 512, 16, 2, true>("v512_i16_occ2");
            case 7: return make_msdu<WORDS, FLOATK, 256, 16, 4, true>("v256_i16_occ4");
        }
    }
    return MsdULaunch{};
}

template <int WORDS, bool FLOATK, int THREADS, int ITEMS, int CBITS, int MINB, bool AUTOTOP = false>
static LocalCountLaunch make_local_count() {
    using SM = LocalCountSmem<WORDS, THREADS, ITEMS, CBITS>;
    LocalCountLaunch l{};
    l.kernel = &local_count_kernel<WORDS, FLOATK, THREADS, ITEMS, CBITS, MINB, AUTOTOP>;
    l.smem = SM::TOTAL;
    l.threads = THREADS;
    l.cap = SM::CAP;
    l.cbits = CBITS;
    return l;
}

// size 0 / 1 / 2 = segments of <= 4096 / 8192 / 16384 records; wide = twice the threads with half the records each
// autotop (per-segment top bit) exists in the shape each width runs by default: wide for 1-word, narrow for 2-word records
template <int WORDS, bool FLOATK>
static LocalCountLaunch pick_local_count(int size, bool wide, bool autotop) {
    if (autotop) {
        if constexpr (WORDS == 1) {
            switch (size) {
                case 0: return make_local_count<WORDS, FLOATK, 256, 16, 13, 4, true>();
                case 1: return make_local_count<WORDS, FLOATK, 512, 16, 14, 2, true>();
                default: return make_local_count<WORDS, FLOATK, 1024, 16, 15, 1, true>();
            }
        } else {
            switch (size) {
                case 0: return make_local_count<WORDS, FLOATK, 128, 32, 13, 4, true>();
                case 1: return make_local_count<WORDS, FLOATK, 256, 32, 14, 2, true>();
                default: return make_local_count<WORDS, FLOATK, 512, 32, 15, 1, true>();
            }
        }
    }
    if (wide) {
        switch (size) {
            case 0: return make_local_count<WORDS, FLOATK, 256, 16, 13, 4>();
            case 1: return make_local_count<WORDS, FLOATK, 512, 16, 14, 2>();
            default: return make_local_count<WORDS, FLOATK, 1024, 16, 15, 1>();
        }
    }
    switch (size) {
        case 0: return make_local_count<WORDS, FLOATK, 128, 32, 13, 4>();
        case 1: return make_local_count<WORDS, FLOATK, 256, 32, 14, 2>();
        default: return make_local_count<WORDS, FLOATK, 512, 32, 15, 1>();
    }
}

}  // namespace b200
