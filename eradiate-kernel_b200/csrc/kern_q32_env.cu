/* kern_q32_env.cu -- the queue kernel for scenes with an environment emitter (FT_ENV), volpath and volpathmis, fp32, bit-exact instances. */
#include "vp_presets.cuh"
#include "vp_launch.h"

namespace vp_launch {
const void *queue_f32_env(bool mis, const char **preset) { return vp::pick_queue_kernel_env<float>(mis, preset); }
} // namespace vp_launch
