/* kern_twin.cu -- the correctness twins: one thread per stream (vp_kernels.cuh) and the register-resident
   warp-voting wavefront (vp_wavefront.cuh), fp32 and fp64. */
#include "vp_wavefront.cuh"
#include "vp_launch.h"

namespace vp_launch {
const void *stream_f32() { return reinterpret_cast<const void *>(&vp::volpath_kernel<float>); }
const void *stream_f64() { return reinterpret_cast<const void *>(&vp::volpath_kernel<double>); }
const void *lanes_f32() { return reinterpret_cast<const void *>(&vp::volpath_wf_kernel<float, 2>); }
const void *lanes_f64() { return reinterpret_cast<const void *>(&vp::volpath_wf_kernel<double, 2>); }
} // namespace vp_launch
