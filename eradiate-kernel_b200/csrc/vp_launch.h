/*
 * vp_launch.h -- what the host API (b2vp_api.cu) sees of the render kernels.
 *
 * Every render kernel has the signature (DScene<F>, DStreams, RenderOut<F>) and is launched with cudaLaunchKernel
 * through an opaque host entry (`const void *`).  The kernels are instantiated in their own translation units
 * (kern_*.cu) so that they compile in parallel and so that the tolerance-mode instances can be built with other
 * compiler flags (FMA contraction, approximate division) than the bit-exact ones.
 */
#pragma once
#include <cstdint>

namespace vp_launch {

/* bit 31 of `features` set: force the all-features instance (tests: presets and generic give the same bits) */
constexpr uint32_t FORCE_GENERIC = 0x80000000u;

/* The queue kernel (vp_queue.cuh) for a scene-feature set: the smallest preset instance that covers `features`,
   else the all-features instance.  max_threads: 512 (product) or 256.  `*preset` names the instance, `*threads` is
   the CTA size to launch it with (max_threads, except for presets that are built for another size). */
const void *queue_f32(uint32_t features, int max_threads, const char **preset, int *threads);
const void *queue_f64(uint32_t features, int max_threads, const char **preset, int *threads);
/* Tolerance-mode fp32 instances (kern_q32_fast.cu): same kernel text, PCG32 exact, floating point with FMA
   contraction and hardware approximations.  512 threads only. */
const void *queue_f32_fast(uint32_t features, const char **preset, int *threads);
/* The same kernel running volpathmis' state machine (FT_MIS; kern_q*_mis.cu).  512 threads only. */
const void *queue_f32_mis(uint32_t features, const char **preset, int *threads);
const void *queue_f64_mis(uint32_t features, const char **preset, int *threads);
const void *queue_f32_fast_mis(uint32_t features, const char **preset, int *threads);
/* ... and for scenes with an environment emitter (FT_ENV; kern_q*_env.cu): all-features instances, 512 threads */
const void *queue_f32_env(bool mis, const char **preset);
const void *queue_f64_env(bool mis, const char **preset);
const void *queue_f32_fast_env(bool mis, const char **preset);
/* The one-thread-per-stream twin (vp_kernels.cuh) and the register-resident warp-voting wavefront (vp_wavefront.cuh) */
const void *stream_f32();
const void *stream_f64();
const void *lanes_f32();
const void *lanes_f64();

} // namespace vp_launch
