/* kern_q32_mis.cu -- the queue kernel running volpathmis (FT_MIS), fp32, bit-exact instances. */
#include "vp_presets.cuh"
#include "vp_launch.h"

namespace vp_launch {
const void *queue_f32_mis(uint32_t features, const char **preset, int *threads) {
    const bool generic = (features & 0x80000000u) != 0;
    return vp::pick_queue_kernel_mis<float>(features & ~0x80000000u, generic, preset, threads);
}
} // namespace vp_launch
