// K2-U / K3-U instantiations for 1-word records (split per width so the widths compile in parallel).
#include "unstable_inst.cuh"
namespace b200 {
MsdULaunch get_msdu_launch_w1(bool floatk, int variant) {
    return floatk ? pick_msdu<1, true>(variant) : pick_msdu<1, false>(variant);
}
LocalCountLaunch get_local_count_launch_w1(bool floatk, int size, bool wide, bool autotop) {
    return floatk ? pick_local_count<1, true>(size, wide, autotop) : pick_local_count<1, false>(size, wide, autotop);
}
}  // namespace b200
