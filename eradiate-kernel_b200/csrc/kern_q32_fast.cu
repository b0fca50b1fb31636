/* kern_q32_fast.cu -- the queue kernel, fp32, TOLERANCE mode: the same kernel text as kern_q32.cu compiled with
   FMA contraction and approximate division / square root (-fmad=true -prec-div=false -prec-sqrt=false -ftz=true)
   and with the transcendental functions of vp_math.cuh mapped to the hardware approximations (VP_FASTMATH).
   Integer work (PCG32, TEA, indices, counters) is identical; floating-point results agree with the exact
   instances within the tolerance of the north star (1e-5 relative per path for paths that take the same events),
   not bit for bit.  Own namespace: the inline functions of the two builds must never be merged by the linker. */
#define VP_NS vp_fast
#define VP_FASTMATH 1
#include "vp_presets.cuh"
#include "vp_launch.h"

namespace vp_launch {
const void *queue_f32_fast(uint32_t features, const char **preset, int *threads) {
    const bool generic = (features & 0x80000000u) != 0;
    features &= ~0x80000000u;
    return vp_fast::pick_queue_kernel<float, 512>(features, generic, preset, threads);
}
} // namespace vp_launch
