/*
 * vp_presets.cuh -- queue-kernel instances: the generic one and presets for the common scene classes
 * (see vp_scene.cuh, FT_*: a preset carries no code for features its scenes do not use).
 */
#pragma once
#include "vp_queue.cuh"

namespace VP_NS {

/* rectangles and small meshes (BVH staged in shared memory), no sphere (double-precision quadratic) */
constexpr uint32_t PRE_FLAT = FT_RECT | FT_MESH_SMEM | FT_MESH_ATTR;
constexpr uint32_t PRE_HOMOG = PRE_FLAT | FT_HOMOG;                                      /* C1 */
constexpr uint32_t PRE_HOMOG_BIG = FT_RECT | FT_MESH | FT_MESH_ATTR | FT_HOMOG;          /* C5: BVH in global memory */
constexpr uint32_t PRE_HET = PRE_FLAT | FT_HETERO | FT_TRILINEAR | FT_GRID_SMEM;         /* C2, C4: grid staged */
constexpr uint32_t PRE_HET_G = PRE_FLAT | FT_HETERO | FT_TRILINEAR | FT_GRID_GLOBAL;     /* large sigma_t grid, constant albedo */
constexpr uint32_t PRE_HET_ALB = PRE_FLAT | FT_HETERO | FT_TRILINEAR | FT_GRID_ALBEDO | FT_GRID_GLOBAL;   /* C3 */
/* C2 with an aerosol layer, what Eradiate's atmospheres with particles look like: molecules and
   aerosols as `tabphase` / `blendphase` (its weight a staged 1 x 1 x N grid or a constant), seen by any sensor
   including `distantflux` / `irradiancemeter` / rays starting on a shape.  The generic instance renders these at
   157-208 Mpaths/s where the plain-Rayleigh preset does 252-338 (tools/bench_aerosol.py) */
constexpr uint32_t PRE_HET_AER = PRE_HET | FT_TABPHASE | FT_BLENDPHASE | FT_DISTANTFLUX;   /* + the flux / irradiance sensors */
/* a large mesh (BVH in HBM: a canopy) under a heterogeneous atmosphere with or without particles: Eradiate's
   surface-atmosphere scenes (RAMI4ATM-like) */
constexpr uint32_t PRE_HET_BIG = (PRE_HET_AER & ~FT_MESH_SMEM) | FT_MESH;
constexpr uint32_t PRE_HET_MG = PRE_HET | FT_MAJORANT_GRID;           /* the same with majorant-grid free flights */
constexpr uint32_t PRE_HET_ALB_MG = PRE_HET_ALB | FT_MAJORANT_GRID;

template <typename F, int MAXT, uint32_t FEAT> const void *q_entry() {
    return reinterpret_cast<const void *>(&volpath_q_kernel<F, MAXT, FEAT>);
}

/* CTA size of the large-mesh preset in fp32.  BVH traversal through L2 is a chain of dependent loads; with the BVH out
   of shared memory the kernel is bound by that latency, and more warps hide more of it than the registers they give
   up cost: C5 123 Mpaths/s at 512 threads (116 registers), 142 at 640 (96), **157 at 768 (80, 30-130 bytes of spills)**,
   155 at 1024 (64).  Every other preset loses with more threads (C2 335 -> 316 at 640, C3 176 -> 142). */
constexpr int BIG_MESH_THREADS = 768;

template <typename F, int MAXT> const void *pick_queue_kernel(uint32_t ft, bool force_generic, const char **name, int *threads) {
    *threads = MAXT;
    if (force_generic) ft = FT_ALL;
    if ((ft & ~PRE_HOMOG) == 0) { *name = "homogeneous"; return q_entry<F, MAXT, PRE_HOMOG>(); }
    if ((ft & ~PRE_HOMOG_BIG) == 0) {
        *name = "homogeneous, large mesh";
        if constexpr (sizeof(F) == 4 && MAXT == 512) { *threads = BIG_MESH_THREADS; return q_entry<F, BIG_MESH_THREADS, PRE_HOMOG_BIG>(); }
        else return q_entry<F, MAXT, PRE_HOMOG_BIG>();
    }
    if ((ft & ~PRE_HET) == 0) { *name = "heterogeneous"; return q_entry<F, MAXT, PRE_HET>(); }
    if constexpr (MAXT == 512) {
        if ((ft & ~PRE_HET_AER) == 0) { *name = "heterogeneous, aerosols"; return q_entry<F, MAXT, PRE_HET_AER>(); }
    }
    if constexpr (MAXT == 512) {
        if ((ft & ~PRE_HET_BIG) == 0) {
            *name = "heterogeneous, large mesh";
            if constexpr (sizeof(F) == 4) { *threads = BIG_MESH_THREADS; return q_entry<F, BIG_MESH_THREADS, PRE_HET_BIG>(); }
            else return q_entry<F, MAXT, PRE_HET_BIG>();
        }
    }
    if ((ft & ~PRE_HET_G) == 0) { *name = "heterogeneous, grid in HBM"; return q_entry<F, MAXT, PRE_HET_G>(); }
    if ((ft & ~PRE_HET_ALB) == 0) { *name = "heterogeneous+albedo grid"; return q_entry<F, MAXT, PRE_HET_ALB>(); }
    if ((ft & ~PRE_HET_MG) == 0) { *name = "heterogeneous, majorant grid"; return q_entry<F, MAXT, PRE_HET_MG>(); }
    if ((ft & ~PRE_HET_ALB_MG) == 0) { *name = "heterogeneous+albedo grid, majorant grid"; return q_entry<F, MAXT, PRE_HET_ALB_MG>(); }
    *name = "generic";
    return q_entry<F, MAXT, FT_ALL>();
}

/* volpathmis instances (FT_MIS): 512 threads only, in their own translation units (kern_q*_mis.cu) */
template <typename F> const void *pick_queue_kernel_mis(uint32_t ft, bool force_generic, const char **name, int *threads) {
    *threads = 512;
    if (force_generic) ft = FT_ALL;
    if ((ft & ~PRE_HOMOG) == 0) { *name = "volpathmis, homogeneous"; return q_entry<F, 512, PRE_HOMOG | FT_MIS>(); }
    if constexpr (sizeof(F) == 4) {   /* C5 under volpathmis: the larger CTA of the large-mesh preset */
        if ((ft & ~PRE_HOMOG_BIG) == 0) {
            *name = "volpathmis, homogeneous, large mesh"; *threads = BIG_MESH_THREADS;
            return q_entry<F, BIG_MESH_THREADS, PRE_HOMOG_BIG | FT_MIS>();
        }
    }
    if ((ft & ~PRE_HET) == 0) { *name = "volpathmis, heterogeneous"; return q_entry<F, 512, PRE_HET | FT_MIS>(); }
    if ((ft & ~PRE_HET_AER) == 0) { *name = "volpathmis, heterogeneous, aerosols"; return q_entry<F, 512, PRE_HET_AER | FT_MIS>(); }
    if ((ft & ~PRE_HET_BIG) == 0) {
        *name = "volpathmis, heterogeneous, large mesh";
        if constexpr (sizeof(F) == 4) { *threads = BIG_MESH_THREADS; return q_entry<F, BIG_MESH_THREADS, PRE_HET_BIG | FT_MIS>(); }
        else return q_entry<F, 512, PRE_HET_BIG | FT_MIS>();
    }
    if ((ft & ~PRE_HET_ALB) == 0) { *name = "volpathmis, heterogeneous+albedo grid"; return q_entry<F, 512, PRE_HET_ALB | FT_MIS>(); }
    *name = "volpathmis, generic";
    return q_entry<F, 512, FT_ALL | FT_MIS>();
}

/* scenes with an environment emitter (FT_ENV): the all-features instance of either integrator */
template <typename F> const void *pick_queue_kernel_env(bool mis, const char **name) {
    if (mis) { *name = "volpathmis, environment, generic"; return q_entry<F, 512, FT_ALL | FT_ENV | FT_MIS>(); }
    *name = "environment, generic";
    return q_entry<F, 512, FT_ALL | FT_ENV>();
}

} // namespace VP_NS
