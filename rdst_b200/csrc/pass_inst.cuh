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

// variants that also exist with the hot-bucket ballot path (the per-width defaults)
constexpr bool variant_has_hot(int v) { return v == 12 || v == 18 || v == 21 || v == 22; }

template <int WORDS, bool FLOATK, bool REMAP, int V>
static PassLaunch make_launch(bool hot = false, bool seg = false) {
    constexpr VariantInfo vi = kVariants[V];
    constexpr int ITEMS = vi.items[WORDS - 1];
    PassLaunch pl{};
    if constexpr (vi.algo == 2) {
        constexpr int SUB = vi.sub ? vi.sub : 1;
        using SM = Or16Smem<WORDS, vi.threads, ITEMS, SUB>;
        pl.kernel = &lsd_pass_or16_kernel<WORDS, vi.threads, ITEMS, FLOATK, REMAP, vi.min_blocks, SUB, false>;
        pl.hot = false;
        pl.seg = false;
        if constexpr (!REMAP && variant_has_hot(V)) {
            if (hot) {
                pl.kernel = &lsd_pass_or16_kernel<WORDS, vi.threads, ITEMS, FLOATK, REMAP, vi.min_blocks, SUB, true>;
                pl.hot = true;
            }
            if constexpr (WORDS <= 2 && variant_has_hot(V)) {  // the default variants only
                if (seg) {  // segmented mode (second MSD pass of the hybrid path)
                    pl.kernel = &lsd_pass_or16_kernel<WORDS, vi.threads, ITEMS, FLOATK, REMAP, vi.min_blocks, SUB, false, true>;
                    pl.hot = false;
                    pl.seg = true;
                }
            }
        }
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
static PassLaunch pick_launch(int v, bool hot, bool seg) {
    switch (v) {
        case 0: return make_launch<WORDS, FLOATK, REMAP, 0>(hot, seg);
        case 1: return make_launch<WORDS, FLOATK, REMAP, 1>(hot, seg);
        case 2: return make_launch<WORDS, FLOATK, REMAP, 2>(hot, seg);
        case 3: return make_launch<WORDS, FLOATK, REMAP, 3>(hot, seg);
        case 4: return make_launch<WORDS, FLOATK, REMAP, 4>(hot, seg);
        case 5: return make_launch<WORDS, FLOATK, REMAP, 5>(hot, seg);
        case 6: return make_launch<WORDS, FLOATK, REMAP, 6>(hot, seg);
        case 7: return make_launch<WORDS, FLOATK, REMAP, 7>(hot, seg);
        case 8: return make_launch<WORDS, FLOATK, REMAP, 8>(hot, seg);
        case 9: return make_launch<WORDS, FLOATK, REMAP, 9>(hot, seg);
        case 10: return make_launch<WORDS, FLOATK, REMAP, 10>(hot, seg);
        case 11: return make_launch<WORDS, FLOATK, REMAP, 11>(hot, seg);
        case 12: return make_launch<WORDS, FLOATK, REMAP, 12>(hot, seg);
        case 13: return make_launch<WORDS, FLOATK, REMAP, 13>(hot, seg);
        case 14: return make_launch<WORDS, FLOATK, REMAP, 14>(hot, seg);
        case 15: return make_launch<WORDS, FLOATK, REMAP, 15>(hot, seg);
        case 16: return make_launch<WORDS, FLOATK, REMAP, 16>(hot, seg);
        case 17: return make_launch<WORDS, FLOATK, REMAP, 17>(hot, seg);
        case 18: return make_launch<WORDS, FLOATK, REMAP, 18>(hot, seg);
        case 19: return make_launch<WORDS, FLOATK, REMAP, 19>(hot, seg);
        case 20: return make_launch<WORDS, FLOATK, REMAP, 20>(hot, seg);
        case 21: return make_launch<WORDS, FLOATK, REMAP, 21>(hot, seg);
        case 22: return make_launch<WORDS, FLOATK, REMAP, 22>(hot, seg);
        default: break;
    }
    PassLaunch none{};
    return none;
}

template <int WORDS>
static PassLaunch pick_launch_w(bool floatk, bool remap, int v, bool hot, bool seg = false) {
    if (remap) {
        // the range-partition pass exists in one variant only (kRemapVariant)
        return floatk ? make_launch<WORDS, true, true, kRemapVariant>() : make_launch<WORDS, false, true, kRemapVariant>();
    }
    return floatk ? pick_launch<WORDS, true, false>(v, hot, seg) : pick_launch<WORDS, false, false>(v, hot, seg);
}

}  // namespace b200
