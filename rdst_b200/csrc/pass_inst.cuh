// pass_inst.cuh — instantiates the K2 pass kernel for one record width (WORDS)
// over every variant; included by pass_w{1,2,3,4}.cu so the widths compile in
// parallel.
#pragma once
#include "onesweep.cuh"
#include "pass_or.cuh"
#include "pass_or16.cuh"
#include "variants.h"
#include "launch.h"

namespace b200 {

template <int WORDS, bool FLOATK, bool REMAP, int V>
static PassLaunch make_launch() {
    constexpr VariantInfo vi = kVariants[V];
    constexpr int ITEMS = vi.items[WORDS - 1];
    PassLaunch pl;
    if constexpr (vi.algo == 2) {
        constexpr int SUB = vi.sub ? vi.sub : 1;
        using SM = Or16Smem<WORDS, vi.threads, ITEMS, SUB>;
        pl.kernel = &lsd_pass_or16_kernel<WORDS, vi.threads, ITEMS, FLOATK, REMAP, vi.min_blocks, SUB>;
        pl.smem = SM::TOTAL;
        pl.tile = SM::TILE;
    } else if constexpr (vi.algo == 1) {
        using SM = OrSmem<WORDS, vi.threads, ITEMS>;
        pl.kernel = &lsd_pass_or_kernel<WORDS, vi.threads, ITEMS, FLOATK, REMAP, vi.min_blocks>;
        pl.smem = SM::TOTAL;
        pl.tile = SM::TILE;
    } else {
        using SM = PassSmem<WORDS, vi.threads, ITEMS, vi.tma>;
        pl.kernel = &onesweep_pass_kernel<WORDS, vi.threads, ITEMS, FLOATK, vi.tma, vi.match_hw, REMAP, vi.min_blocks>;
        pl.smem = SM::TOTAL;
        pl.tile = SM::TILE;
    }
    pl.threads = vi.threads;
    return pl;
}

template <int WORDS, bool FLOATK, bool REMAP>
static PassLaunch pick_launch(int v) {
    switch (v) {
        case 0: return make_launch<WORDS, FLOATK, REMAP, 0>();
        case 1: return make_launch<WORDS, FLOATK, REMAP, 1>();
        case 2: return make_launch<WORDS, FLOATK, REMAP, 2>();
        case 3: return make_launch<WORDS, FLOATK, REMAP, 3>();
        case 4: return make_launch<WORDS, FLOATK, REMAP, 4>();
        case 5: return make_launch<WORDS, FLOATK, REMAP, 5>();
        case 6: return make_launch<WORDS, FLOATK, REMAP, 6>();
        case 7: return make_launch<WORDS, FLOATK, REMAP, 7>();
        case 8: return make_launch<WORDS, FLOATK, REMAP, 8>();
        case 9: return make_launch<WORDS, FLOATK, REMAP, 9>();
        case 10: return make_launch<WORDS, FLOATK, REMAP, 10>();
        case 11: return make_launch<WORDS, FLOATK, REMAP, 11>();
        case 12: return make_launch<WORDS, FLOATK, REMAP, 12>();
        case 13: return make_launch<WORDS, FLOATK, REMAP, 13>();
        case 14: return make_launch<WORDS, FLOATK, REMAP, 14>();
        case 15: return make_launch<WORDS, FLOATK, REMAP, 15>();
        case 16: return make_launch<WORDS, FLOATK, REMAP, 16>();
        case 17: return make_launch<WORDS, FLOATK, REMAP, 17>();
        case 18: return make_launch<WORDS, FLOATK, REMAP, 18>();
        case 19: return make_launch<WORDS, FLOATK, REMAP, 19>();
        case 20: return make_launch<WORDS, FLOATK, REMAP, 20>();
        case 21: return make_launch<WORDS, FLOATK, REMAP, 21>();
        case 22: return make_launch<WORDS, FLOATK, REMAP, 22>();
        default: break;
    }
    PassLaunch none{};
    return none;
}

template <int WORDS>
static PassLaunch pick_launch_w(bool floatk, bool remap, int v) {
    if (remap) {
        // the range-partition pass exists in one variant only (kRemapVariant)
        return floatk ? make_launch<WORDS, true, true, kRemapVariant>() : make_launch<WORDS, false, true, kRemapVariant>();
    }
    return floatk ? pick_launch<WORDS, true, false>(v) : pick_launch<WORDS, false, false>(v);
}

}  // namespace b200
