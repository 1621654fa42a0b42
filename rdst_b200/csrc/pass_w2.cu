#include "pass_inst.cuh"
namespace b200 {
PassLaunch get_pass_launch_w2(bool floatk, bool remap, int variant, bool hot) {
    return pick_launch_w<2>(floatk, remap, variant, hot);
}
}  // namespace b200
