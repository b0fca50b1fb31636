/* kern_q32.cu -- the queue kernel (vp_queue.cuh), fp32, bit-exact instances (scalar_mono). */
#include "vp_presets.cuh"
#include "vp_launch.h"

namespace vp_launch {
const void *queue_f32(uint32_t features, int max_threads, const char **preset, int *threads) {
    const bool generic = (features & 0x80000000u) != 0;   /* host flag: force the all-features instance */
    features &= ~0x80000000u;
    if (max_threads == 256) return vp::pick_queue_kernel<float, 256>(features, generic, preset, threads);
    return vp::pick_queue_kernel<float, 512>(features, generic, preset, threads);
}
} // namespace vp_launch
