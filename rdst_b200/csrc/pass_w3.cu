#include "pass_inst.cuh"
namespace b200 {
PassLaunch get_pass_launch_w3(bool floatk, bool remap, int variant, bool hot, bool seg) {
    return pick_launch_w<3>(floatk, remap, variant, hot, seg);
}
}  // namespace b200
