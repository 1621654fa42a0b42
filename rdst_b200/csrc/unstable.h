// unstable.h — host-side handles on the kernels of the unstable (key-only) hybrid route:
// K2-U (pass_atomic.cuh) and K3-U (local_count.cuh).  Instantiated in unstable_w{1,2}.cu.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>

namespace b200 {
struct PassParams;
struct LocalCountParams;

constexpr int NUM_MSDU_VARIANTS = 8;
struct MsdULaunch {
    void (*kernel)(const PassParams);
    size_t smem;
    int tile;
    int threads;
    const char* name;
};
MsdULaunch get_msdu_launch_w1(bool floatk, int variant);
MsdULaunch get_msdu_launch_w2(bool floatk, int variant);

// K3-U comes in three sizes: 0 = 128 threads, segments <= 4096 records, several CTAs per SM; 1 = 256 threads,
// <= 8192 records, two CTAs per SM; 2 = 512 threads, <= 16384 records, one CTA per SM.
struct LocalCountLaunch {
    void (*kernel)(const LocalCountParams);
    size_t smem;
    int threads;
    int cap;    // records per segment
    int cbits;  // widest digit
};
// autotop = the instantiation in which every segment finds its own top bit (local_count.cuh); `wide` is ignored then
LocalCountLaunch get_local_count_launch_w1(bool floatk, int size, bool wide, bool autotop = false);
LocalCountLaunch get_local_count_launch_w2(bool floatk, int size, bool wide, bool autotop = false);
}  // namespace b200
