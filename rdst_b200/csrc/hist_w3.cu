#include "hist_inst.cuh"
namespace b200 {
bool launch_hist_fast_w3(uint32_t L, bool floatk, int grid, cudaStream_t st, const HistParams& hp) {
    return launch_hist_fast1<3>(L, floatk, grid, st, hp);
}
}  // namespace b200
