// launch.h — host-side handle on one instantiation of the K2 pass kernel.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>

namespace b200 {
struct PassParams;
struct PassLaunch {
    void (*kernel)(const PassParams);
    size_t smem;
    int tile;
    int threads;
    bool hot;  // instantiation with the hot-bucket ballot path
    bool seg;  // instantiation with the segmented mode
};
PassLaunch get_pass_launch_w1(bool floatk, bool remap, int variant, bool hot, bool seg);
PassLaunch get_pass_launch_w2(bool floatk, bool remap, int variant, bool hot, bool seg);
PassLaunch get_pass_launch_w3(bool floatk, bool remap, int variant, bool hot, bool seg);
PassLaunch get_pass_launch_w4(bool floatk, bool remap, int variant, bool hot, bool seg);
struct HistParams;
bool launch_hist_fast_w1(uint32_t L, bool floatk, int grid, cudaStream_t st, const HistParams& hp);
bool launch_hist_fast_w2(uint32_t L, bool floatk, int grid, cudaStream_t st, const HistParams& hp);
bool launch_hist_fast_w3(uint32_t L, bool floatk, int grid, cudaStream_t st, const HistParams& hp);
bool launch_hist_fast_w4(uint32_t L, bool floatk, int grid, cudaStream_t st, const HistParams& hp);
}  // namespace b200
