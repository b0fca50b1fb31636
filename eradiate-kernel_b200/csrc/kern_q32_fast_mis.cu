/* kern_q32_fast_mis.cu -- volpathmis on the queue kernel, fp32 TOLERANCE mode (flags and macros of kern_q32_fast.cu). */
#define VP_NS vp_fast
#define VP_FASTMATH 1
#include "vp_presets.cuh"
#include "vp_launch.h"

namespace vp_launch {
const void *queue_f32_fast_mis(uint32_t features, const char **preset, int *threads) {
    const bool generic = (features & 0x80000000u) != 0;
    return vp_fast::pick_queue_kernel_mis<float>(features & ~0x80000000u, generic, preset, threads);
}
} // namespace vp_launch
