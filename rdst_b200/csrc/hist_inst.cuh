// hist_inst.cuh — instantiates the fast K1 histogram kernels for one record
// width; included by hist_w{1,2,3,4}.cu so the widths compile in parallel.
#pragma once
#include "hist.cuh"
#include "launch.h"

namespace b200 {
template <int WORDS, int L>
void launch_hist_fast2(bool floatk, int grid, cudaStream_t st, const HistParams& hp) {
    if (floatk)
        hist_fast_kernel<WORDS, L, true><<<grid, HIST_THREADS, 0, st>>>(hp);
    else
        hist_fast_kernel<WORDS, L, false><<<grid, HIST_THREADS, 0, st>>>(hp);
}

template <int WORDS>
bool launch_hist_fast1(uint32_t L, bool floatk, int grid, cudaStream_t st, const HistParams& hp) {
    switch (L) {
#define B200_HCASE(LL)                                                   \
    case LL:                                                             \
        if constexpr ((LL + 3) / 4 <= WORDS) {                           \
            launch_hist_fast2<WORDS, LL>(floatk, grid, st, hp);     \
            return true;                                                 \
        }                                                                \
        return false;
        B200_HCASE(1)
        B200_HCASE(2)
        B200_HCASE(3)
        B200_HCASE(4)
        B200_HCASE(5)
        B200_HCASE(6)
        B200_HCASE(7)
        B200_HCASE(8)
        B200_HCASE(16)
#undef B200_HCASE
    }
    return false;
}


}  // namespace b200
