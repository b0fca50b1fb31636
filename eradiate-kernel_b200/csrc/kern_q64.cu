/* kern_q64.cu -- the queue kernel (vp_queue.cuh), fp64 instances (scalar_mono_double). */
#include "vp_presets.cuh"
#include "vp_launch.h"

namespace vp_launch {
const void *queue_f64(uint32_t features, int max_threads, const char **preset, int *threads) {
    const bool generic = (features & 0x80000000u) != 0;
    features &= ~0x80000000u;
    if (max_threads == 256) return vp::pick_queue_kernel<double, 256>(features, generic, preset, threads);
    return vp::pick_queue_kernel<double, 512>(features, generic, preset, threads);
}
} // namespace vp_launch
