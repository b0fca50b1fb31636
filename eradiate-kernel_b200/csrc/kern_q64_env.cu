/* kern_q64_env.cu -- the queue kernel for scenes with an environment emitter (FT_ENV), volpath and volpathmis, fp64, bit-exact instances. */
#include "vp_presets.cuh"
#include "vp_launch.h"

namespace vp_launch {
const void *queue_f64_env(bool mis, const char **preset) { return vp::pick_queue_kernel_env<double>(mis, preset); }
} // namespace vp_launch
