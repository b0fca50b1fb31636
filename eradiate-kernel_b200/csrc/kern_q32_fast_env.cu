/* kern_q32_fast_env.cu -- the queue kernel for scenes with an environment emitter (FT_ENV), volpath and volpathmis, fp32 TOLERANCE mode (flags and macros of kern_q32_fast.cu). */
#define VP_NS vp_fast
#define VP_FASTMATH 1
#include "vp_presets.cuh"
#include "vp_launch.h"

namespace vp_launch {
const void *queue_f32_fast_env(bool mis, const char **preset) { return vp_fast::pick_queue_kernel_env<float>(mis, preset); }
} // namespace vp_launch
