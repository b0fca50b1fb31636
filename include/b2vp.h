/*
 * b2vp.h -- C-ABI of the B200-native `volpath` hot path (mono variants).
 *
 * This is the drop-in boundary.  Everything below is
 * plain C: POD structs, pointers and sizes.  The reference has no FFI for this
 * path (it is a C++ plugin, `plugins/volpath.so`); the entry points here are
 * what the replacement plugin's `render()` override binds to.  Reference
 * interfaces replaced (paths relative to the reference tree):
 *
 *   b2vp_scene_create    <- the object graph an integrator sees in
 *                           SamplingIntegrator::render(Scene*, Sensor*)
 *                           src/librender/integrator.cpp:51, built by
 *                           Scene::Scene src/librender/scene.cpp:22-108
 *   b2vp_render          <- SamplingIntegrator::render / render_block /
 *                           render_sample src/librender/integrator.cpp:51-288
 *                           + VolumetricPathIntegrator::sample
 *                           src/integrators/volpath.cpp:38-257
 *   b2vp_scene_update    <- Object::parameters_changed, e.g.
 *                           src/textures/grid3d.cpp:370-391 and
 *                           TraversalCallback::put_parameter("scale"...)
 *                           src/media/heterogeneous.cpp:56-61
 *   b2vp_cancel          <- Integrator::cancel include/mitsuba/render/integrator.h:51
 *   b2vp_last_error      <- Throw()/std::runtime_error include/mitsuba/core/logger.h
 *
 * Real-valued scene parameters are carried as `double`.  For the fp32 variant
 * (scalar_mono) the host computes them in float and stores the float value
 * exactly in the double slot; the device narrows them back without rounding.
 *
 * All structs start with {abi_version, struct_size} so that a plugin built
 * against an older header is rejected instead of mis-read.
 */
#ifndef B2VP_H
#define B2VP_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2VP_ABI_VERSION 4u

/* ---- status codes ------------------------------------------------------ */
enum {
    B2VP_OK              = 0,
    B2VP_ERR_INVALID     = 1, /* malformed descriptor / argument            */
    B2VP_ERR_UNSUPPORTED = 2, /* scene content outside SURVEY.md section 8  */
    B2VP_ERR_CUDA        = 3, /* CUDA runtime error (message in last_error) */
    B2VP_ERR_NO_DEVICE   = 4, /* no usable sm_100 device: there is NO CPU fallback */
    B2VP_ERR_CANCELLED   = 5
};

/* ---- enumerations ------------------------------------------------------ */
enum { B2VP_PREC_F32 = 0 /* scalar_mono */, B2VP_PREC_F64 = 1 /* scalar_mono_double */ };

enum { B2VP_SHAPE_RECTANGLE = 0, B2VP_SHAPE_SPHERE = 1, B2VP_SHAPE_MESH = 2 };
enum { B2VP_BSDF_DIFFUSE = 0, B2VP_BSDF_NULL = 1, B2VP_BSDF_RPV = 2, B2VP_BSDF_BILAMBERTIAN = 3 };
enum { B2VP_MEDIUM_HOMOGENEOUS = 0, B2VP_MEDIUM_HETEROGENEOUS = 1 };
enum { B2VP_VOLUME_CONST = 0, B2VP_VOLUME_GRID = 1 };
enum { B2VP_FILTER_NEAREST = 0, B2VP_FILTER_TRILINEAR = 1 };
enum { B2VP_WRAP_REPEAT = 0, B2VP_WRAP_MIRROR = 1, B2VP_WRAP_CLAMP = 2 };
enum { B2VP_PHASE_ISOTROPIC = 0, B2VP_PHASE_RAYLEIGH = 1, B2VP_PHASE_HG = 2, B2VP_PHASE_TABULATED = 3, B2VP_PHASE_BLEND = 4 };
/* CONSTANT: src/emitters/constant.cpp, the environment emitter (at most one per scene, scene.cpp:40-47):
   `irradiance` holds its radiance, `direction` is unused */
enum { B2VP_EMITTER_DIRECTIONAL = 0, B2VP_EMITTER_CONSTANT = 1 };
enum {
    B2VP_SENSOR_DISTANT        = 0, /* src/sensors/distant.cpp        */
    B2VP_SENSOR_MDISTANT       = 1, /* src/sensors/mdistant.cpp       */
    B2VP_SENSOR_RADIANCEMETER  = 2, /* src/sensors/radiancemeter.cpp  */
    B2VP_SENSOR_MRADIANCEMETER = 3, /* src/sensors/mradiancemeter.cpp */
    B2VP_SENSOR_DISTANTFLUX    = 4, /* src/sensors/distantflux.cpp    */
    /* src/sensors/irradiancemeter.cpp: a sensor attached to a shape (origin_shape, a rectangle): positions are
       sampled on it, directions from the cosine hemisphere about its normal, weight pi */
    B2VP_SENSOR_IRRADIANCEMETER = 5
};
enum { B2VP_DIR_SINGLE = 0, B2VP_DIR_SAMPLE_WIDTH = 1, B2VP_DIR_SAMPLE_ALL = 2 };
enum { B2VP_TARGET_NONE = 0, B2VP_TARGET_POINT = 1, B2VP_TARGET_SHAPE = 2 };
enum { B2VP_ORIGIN_BSPHERE = 0, B2VP_ORIGIN_SHAPE = 1 };

/* Seeding of the per-stream PCG32 generators. */
enum {
    /* reference scalar_* modes: one stream per pixel, seed = block_id*bs^2 +
       morton index, all spp drawn serially (src/librender/integrator.cpp:197-208) */
    B2VP_SEED_SCALAR_BLOCKS = 0,
    /* reference wavefront mode: stream i = (tea64(seed,i), tea64(i,seed)),
       pixel = i / samples_per_pass (src/librender/sampler.cpp:89-92,
       src/librender/integrator.cpp:147-153) */
    B2VP_SEED_WAVEFRONT = 1,
    /* parity harness: caller supplies (state, inc) and the pixel per stream */
    B2VP_SEED_EXPLICIT = 2
};

/* Floating-point mode of the fp32 kernels (B2vpRenderArgs::math_mode).  Integer work -- PCG32 streams, TEA seeding,
   indices, counters -- is identical in both.
   EXACT: every operation is the IEEE operation the oracle performs, in its order (no FMA contraction, IEEE division
          and square root, polynomial log / exp / sincos / cbrt): films, RNG states, counters and per-path radiance
          are bit-identical to the CPU oracle.  The only mode in fp64.
   FAST:  the same kernels compiled with FMA contraction, approximate division / square root and the hardware
          lg2 / ex2 / sin / cos approximations.  Per-path radiance agrees with EXACT within the north star's fp32
          tolerance (1e-5 relative) on paths that take the same events; a small measured fraction of paths takes a
          different event sequence, as the reference's own fp32 build does against the oracle.
          The mode for production renders; validated against the reference's films and z-test. */
enum { B2VP_MATH_EXACT = 0, B2VP_MATH_FAST = 1 };

/* Which device kernel renders (B2vpRenderArgs::kernel).  All compute the same thing (the same per-path results in
   EXACT mode); AUTO is the product choice, the rest exist for the parity tests and A/B measurements. */
enum {
    B2VP_KERNEL_AUTO = 0,          /* queue kernel, preset instance for the scene class                     */
    B2VP_KERNEL_QUEUE_GENERIC = 1, /* queue kernel, all-features instance                                    */
    B2VP_KERNEL_QUEUE_256 = 2,     /* queue kernel with 256-thread CTAs                                      */
    B2VP_KERNEL_LANES = 3,         /* register-resident warp-voting wavefront (vp_wavefront.cuh)             */
    B2VP_KERNEL_STREAM = 4         /* one thread per stream: the transliterated reference loop (vp_kernels.cuh) */
};

/* B2vpSceneDesc::options: device data layout choices (no counterpart in the reference; results are identical) */
enum {
    B2VP_OPT_NO_CELL_RECORDS = 1u << 0,   /* large trilinear grids: plain 8-gather lookups instead of cell records */
    B2VP_OPT_BVH_LEAF_SHIFT = 8           /* bits 8..10: triangles per BVH leaf (0 = default 2)                     */
};

/* ---- scene description -------------------------------------------------- */

/* Column-major 4x4, m[c*4+r], like enoki::Matrix (transform.h:110-116). */
typedef struct B2vpMat4 { double m[16]; } B2vpMat4;

typedef struct B2vpShape {
    int32_t type;              /* B2VP_SHAPE_*                                 */
    int32_t bsdf;              /* index into bsdfs                             */
    int32_t interior_medium;   /* index into media or -1                       */
    int32_t exterior_medium;   /* index into media or -1                       */
    B2vpMat4 to_world;         /* rectangle / sphere                           */
    B2vpMat4 to_object;        /* inverse of to_world                          */
    double frame_s[3];         /* rectangle: dp_du  (rectangle.cpp:86-90)      */
    double frame_t[3];         /* rectangle: dp_dv                             */
    double frame_n[3];         /* rectangle: unit normal                       */
    double inv_surface_area;   /* rectangle / sphere                           */
    double center[3];          /* sphere (sphere.cpp:118)                      */
    double radius;             /* sphere                                       */
    int32_t flip_normals;      /* sphere                                       */
    /* mesh: ranges into the scene-wide vertex/face pools                      */
    uint32_t vertex_offset, vertex_count;
    uint32_t face_offset, face_count;
    int32_t has_vertex_normals, has_vertex_texcoords;
    /* 1: not a shape of the scene but a sensor's own target or origin shape (a <shape name="ray_target"> or
       <shape name="ray_origin"> nested in the sensor, distant.cpp:270-290): used by the sensor, never intersected by
       paths, outside the scene's bounding box.  Rectangles only */
    int32_t sample_only;
} B2vpShape;

typedef struct B2vpBsdf {
    int32_t type;              /* B2VP_BSDF_*                                  */
    int32_t reserved0;
    /* diffuse: p[0]=reflectance.  rpv: p[0]=rho_0 p[1]=k p[2]=g p[3]=rho_c   */
    double p[4];
} B2vpBsdf;

typedef struct B2vpVolume {
    int32_t type;              /* B2VP_VOLUME_*                                */
    int32_t filter;            /* B2VP_FILTER_*                                */
    int32_t wrap;              /* B2VP_WRAP_*                                  */
    int32_t channels;          /* 1 or 3 (3 => luminance in mono, grid3d.cpp:178) */
    int32_t res[3];            /* nx, ny, nz                                   */
    int32_t reserved0;
    double value;              /* constvolume value (constant3d.cpp)           */
    double max_value;          /* Volume::max()                                */
    double bbox_min[3], bbox_max[3];
    B2vpMat4 world_to_local;
    /* grid data, always float32 as on disk (volume_data.h:85-90), laid out
       [((z*ny+y)*nx+x)*channels+c].  Host pointer; copied by scene_create.   */
    const float *data;
} B2vpVolume;

typedef struct B2vpPhase {
    int32_t type;              /* B2VP_PHASE_*                                 */
    int32_t reserved0;
    double g;                  /* hg asymmetry (hg.cpp:41)                     */
    /* tabphase (src/phase/tabphase.cpp:43-66): the phase function tabulated at n_values (>= 2) regularly spaced
       cosines on [-1, 1], not normalised, in the precision the plugin holds them ((ScalarFloat) std::stod(..)).
       Host pointer; copied by scene_create.                                   */
    uint32_t n_values;
    uint32_t reserved1;
    const double *values;
    /* blendphase (src/phase/blendphase.cpp:42-140): two nested phase functions (indices into phases, not blends
       themselves) mixed by a weight volume evaluated at the scattering position, clamped to [0, 1]:
       (1 - w) * child[0] + w * child[1]; sampling picks child[0] when sample1 > w */
    int32_t child[2];
    int32_t weight_volume;     /* index into volumes                           */
    int32_t reserved2;
} B2vpPhase;

typedef struct B2vpMedium {
    int32_t type;              /* B2VP_MEDIUM_*                                */
    int32_t sigma_t;           /* index into volumes                           */
    int32_t albedo;            /* index into volumes                           */
    int32_t phase;             /* index into phases                            */
    int32_t has_spectral_extinction; /* medium.h:79, default true             */
    int32_t sample_emitters;   /* medium.cpp:29                                */
    double scale;              /* heterogeneous.cpp:27 / homogeneous.cpp:27    */
} B2vpMedium;

typedef struct B2vpEmitter {
    int32_t type;              /* B2VP_EMITTER_DIRECTIONAL                     */
    int32_t reserved0;
    double direction[3];       /* to_world * (0,0,1) (directional.cpp:112)     */
    double irradiance;
} B2vpEmitter;

#define B2VP_RFILTER_RESOLUTION 31   /* MTS_FILTER_RESOLUTION (rfilter.h:10) */
#define B2VP_RFILTER_MAX_TAPS 16     /* footprint bound of the device code: radius <= 8 pixels */

typedef struct B2vpSensor {
    int32_t type;              /* B2VP_SENSOR_*                                */
    int32_t direction_type;    /* distant: B2VP_DIR_*                          */
    int32_t target_type;       /* distant/mdistant: B2VP_TARGET_*              */
    int32_t origin_type;       /* distant: B2VP_ORIGIN_*                       */
    int32_t target_shape;      /* shape index (rectangle) when TARGET_SHAPE    */
    int32_t origin_shape;      /* shape index (rectangle) when ORIGIN_SHAPE; irradiancemeter: its parent rectangle */
    int32_t flip_directions;   /* distant.cpp:229                              */
    int32_t needs_aperture_sample; /* sensor.h:79                             */
    int32_t medium;            /* sensor->medium(): index or -1                */
    int32_t n_transforms;      /* mdistant / mradiancemeter sensor count       */
    B2vpMat4 to_world;         /* distant / radiancemeter world transform      */
    double target_point[3];
    double reference_normal[3]; /* distantflux: to_world * (0,0,1) (distantflux.cpp:173-175) */
    double shutter_open, shutter_open_time;
    const double *transforms;  /* n_transforms * 16, column-major              */
    /* film (hdrfilm.cpp / film.cpp) */
    int32_t film_size[2];
    int32_t crop_offset[2];
    int32_t crop_size[2];
    uint32_t sample_count;     /* sampler->sample_count()                      */
    uint32_t reserved0;
    uint64_t sampler_seed;     /* independent sampler `seed` property          */
    /* reconstruction filter of the film (include/mitsuba/core/rfilter.h:55-98, src/libcore/rfilter.cpp:9-21).
       rfilter_radius <= 0.5 + RayEpsilon (the default `box`): every sample lands in the one pixel under it
       (imageblock.cpp:153-160) and the other fields are not read.  Wider filters (tent, gaussian, mitchell,
       catmullrom, lanczos, a wide box): ImageBlock::put's footprint loop (imageblock.cpp:117-152) with
       eval_discretized() over this table */
    double rfilter_radius;     /* m_radius                                     */
    double rfilter_scale;      /* m_scale_factor = MTS_FILTER_RESOLUTION / m_radius, in the variant's float type */
    uint32_t rfilter_taps;     /* n of imageblock.cpp:122: ceil((radius - 2 RayEpsilon) * 2) */
    uint32_t rfilter_border;   /* m_border_size (rfilter.cpp:20)               */
    double rfilter_values[B2VP_RFILTER_RESOLUTION + 1]; /* m_values; the last one is 0 */
} B2vpSensor;

typedef struct B2vpIntegrator {
    int32_t max_depth;         /* -1 => unbounded (integrator.cpp:305-314)     */
    int32_t rr_depth;          /* default 5                                    */
    int32_t hide_emitters;
    uint32_t block_size;       /* 0 => choose like integrator.cpp:89-97        */
    uint32_t samples_per_pass; /* 0xffffffff => all                            */
    /* phase->sample(.., next_1d(), next_2d(), ..) and bsdf->sample(..) draw their
       two sample arguments in compiler-chosen order (volpath.cpp:171,215): 0 =
       right-to-left (GCC, the build SURVEY.md probed), 1 = left-to-right (clang) */
    uint32_t arg_order_ltr;
    /* 1: the volpath is nested in a `moment` integrator (src/integrators/moment.cpp:39-111): the film has
       B2VP_FILM_CHANNELS_MOMENT channels per pixel, X Y Z A W followed by the nested integrator's unweighted
       X Y Z and their squares, and ImageBlock::put skips its validity check (integrator.cpp:113-115) */
    uint32_t moment;
    /* 1: the integrator is `volpathmis` (src/integrators/volpathmis.cpp, spectral MIS over channels; one channel in
       mono) instead of `volpath`: other control flow, other random-number order.  2: volpathmis with
       use_spectral_mis=false (differs in mono only in when an emitter march stops) */
    uint32_t mis;
} B2vpIntegrator;

#define B2VP_FILM_CHANNELS 5
#define B2VP_FILM_CHANNELS_MOMENT 11

typedef struct B2vpSceneDesc {
    uint32_t abi_version;
    uint32_t struct_size;
    int32_t precision;         /* B2VP_PREC_* the host computed parameters in  */
    uint32_t options;          /* B2VP_OPT_* (0 = defaults)                    */

    uint32_t n_shapes, n_bsdfs, n_media, n_volumes, n_phases, n_emitters;
    const B2vpShape *shapes;
    const B2vpBsdf *bsdfs;
    const B2vpMedium *media;
    const B2vpVolume *volumes;
    const B2vpPhase *phases;
    const B2vpEmitter *emitters;

    /* triangle pools (mesh.h:22-28: positions/normals/texcoords always f32)  */
    uint32_t n_vertices, n_faces;
    const float *vertex_positions;  /* 3 per vertex                            */
    const float *vertex_normals;    /* 3 per vertex or NULL                    */
    const float *vertex_texcoords;  /* 2 per vertex or NULL                    */
    const uint32_t *faces;          /* 3 per face, indices local to the shape  */

    /* Scene::bbox() (scene.cpp:38) and the derived bounding sphere with the
       epsilon padding of distant.cpp:292-297 / directional.cpp:68-73.        */
    double bbox_min[3], bbox_max[3];
    double bsphere_center[3];
    double bsphere_radius;

    B2vpSensor sensor;
    B2vpIntegrator integrator;
} B2vpSceneDesc;

/* ---- render arguments & outputs ----------------------------------------- */

/* Per-stream event counters: summed over the samples
   the stream drew.  segments = n_medium_steps + n_intersect.                  */
typedef struct B2vpPathCounters {
    uint32_t n_rng_draws;      /* PCG32 advances                               */
    uint32_t n_medium_steps;   /* Medium::sample_interaction calls             */
    uint32_t n_null;           /* null collisions (main loop + both marches)   */
    uint32_t n_real;           /* real medium scattering events                */
    uint32_t n_intersect;      /* Scene::ray_intersect calls                   */
    uint32_t n_surface;        /* surface events in the main loop              */
} B2vpPathCounters;

/* Whole-launch totals. */
typedef struct B2vpCounters {
    uint64_t n_paths;
    uint64_t n_rng_draws, n_medium_steps, n_null, n_real, n_intersect, n_surface;
    uint64_t n_invalid_samples; /* dropped by ImageBlock::put (imageblock.cpp:85-109) */
    double gpu_ms;              /* CUDA-event time of the render kernels       */
    uint32_t n_launches;        /* kernels launched by this call               */
    uint32_t stopped_early;     /* 1: stopped by timeout_s, 2: by b2vp_cancel; n_paths < requested */
    /* BVH work of the launch (queue kernel): nodes whose box was tested (32 B each) and leaf triangles tested (48 B
       each) -- the algorithmic bytes of the mesh configuration's memory roofline           */
    uint64_t n_bvh_nodes, n_bvh_tris;
} B2vpCounters;

typedef struct B2vpRenderArgs {
    uint32_t abi_version;
    uint32_t struct_size;
    int32_t precision;          /* B2VP_PREC_*: arithmetic of the kernels      */
    int32_t seed_mode;          /* B2VP_SEED_*                                 */
    int32_t device;             /* CUDA device ordinal                         */
    int32_t film_on_device;     /* film pointer is a device pointer            */
    void *cuda_stream;          /* cudaStream_t or NULL for the default stream */

    /* Stream range rendered by this call: [first_stream, first_stream+n_streams)
       of the mode's global stream index space (see b2vp_stream_count).  Sharding
       across GPUs = disjoint ranges + one film sum-reduce.   */
    uint64_t first_stream;
    uint64_t n_streams;         /* 0 => all remaining                          */
    /* Interleaved sharding (many small ranges per GPU instead of one, against tail imbalance):
       with interleave_block = B > 0 the k-th stream of this call, k in [0, n_streams), is the global stream
       first_stream + ((k / B) * interleave_count + interleave_index) * B + k % B, i.e. the call renders every
       interleave_count-th block of B streams, starting with block interleave_index.  n_streams must be given.
       B = samples per pixel and GPU, count = GPUs, index = rank: every GPU renders its share of EVERY pixel. */
    uint64_t interleave_block;
    uint32_t interleave_count, interleave_index;

    /* B2VP_SEED_EXPLICIT only: per-stream start state, increment and pixel.   */
    const uint64_t *explicit_state;
    const uint64_t *explicit_inc;
    const uint32_t *explicit_pixel;   /* x + y*width                           */
    uint32_t explicit_samples_per_stream;
    uint32_t clear_film;        /* zero the film before accumulating           */

    /* Optional per-stream / per-sample outputs for the parity harness (host
       pointers, may be NULL).  out_radiance is float or double by precision
       and holds samples_per_stream entries per stream.                        */
    uint64_t *out_state;
    B2vpPathCounters *out_counters;
    void *out_radiance;
    B2vpCounters *totals;       /* may be NULL                                 */

    /* ---- ABI 3 ---- */
    int32_t math_mode;          /* B2VP_MATH_*; FAST needs precision F32 and a queue kernel                    */
    int32_t kernel;             /* B2VP_KERNEL_*                                                                */
    /* Scheduler parameters of the queue kernel; 0 = the measured default of each:
       [0] path slots per thread x 16 (default 64, or 32 when large grids are gathered through L2),
       [1] STEP keeps its batch while at least this many lanes stay (26), [2] for at most this many steps (8),
       [3] 1 + CTA phase policy (0 per warp, 1 fullest queue, 2 ray-making phases first).                       */
    uint32_t tuning[4];
    uint32_t debug_flags;       /* bit 0: print the queue kernel's scheduler statistics to stderr; bit 1: experiment,
                                   a persisting L2 access window over the queue kernel's cold path records (no gain) */
    /* Majorant grid: 0 = the reference's delta / ratio tracking against ONE global majorant scale * max(sigma_t)
       (src/media/heterogeneous.cpp:29,38) -- the only mode whose random-number consumption, event counters and
       per-path results match the reference.  Non-zero = opt-in fast mode for heterogeneous media with a clamp-mode
       single-channel sigma_t grid: free flights are sampled against a coarse grid of LOCAL majorants (1: up to 16
       cells per axis, more along a lone long axis; n > 1: n cells along the longest axis; at most 4096 cells per
       medium, staged in shared memory).  The estimator is the same and unbiased; tentative collisions are a sparser
       random sequence, so paths draw fewer random numbers and the result agrees with the global-majorant film
       statistically (z-test), not per path.  Queue kernel only; other media in the scene are unaffected.          */
    uint32_t majorant_grid;
    /* Integrator `timeout` property (include/mitsuba/render/integrator.h:143-146, src/librender/integrator.cpp:35,
       59-65,121): > 0 = stop handing out new streams once this many seconds of rendering have passed; paths in flight
       finish, the call returns B2VP_OK with the samples taken so far (the film's weight channel counts them, so it
       stays normalised) and B2vpCounters::stopped_early = 1 -- what SamplingIntegrator::render does when should_stop()
       turns true.  <= 0: no limit.                                                                                    */
    double timeout_s;
    /* Multi-GPU render behind the boundary (no counterpart in the reference, whose render is TBB on one host:
       src/librender/integrator.cpp:109-139).  n_devices > 1: devices[0] must be the scene's device (the root); the
       stream range is dealt out in interleaved blocks to the listed devices, each renders into a private film from
       its own host thread, and ONE NCCL reduce (sum) over NVLink merges the films on the root, from where `film`
       is filled.  The scene is uploaded to a device the first time it renders there and stays resident;
       b2vp_scene_update / _update_grid patch every resident copy.  Not combinable with explicit seeding, per-stream
       outputs or the caller's own interleave_*.  n_devices <= 1 (devices may be NULL): args->device only.            */
    const int32_t *devices;
    uint32_t n_devices;
    uint32_t reserved2;
} B2vpRenderArgs;

typedef struct b2vp_scene b2vp_scene_t;

/* Patch for spectral sweeps: new per-wavelength scalars without re-uploading
   grids or rebuilding the BVH.  Entries with index < 0 are ignored.           */
typedef struct B2vpParamPatch {
    uint32_t abi_version;
    uint32_t struct_size;
    int32_t medium;  double medium_scale;       /* media[medium].scale         */
    int32_t bsdf;    double bsdf_p[4];          /* bsdfs[bsdf].p               */
    int32_t emitter; double emitter_irradiance; /* emitters[emitter].irradiance */
    int32_t volume;  double volume_value;       /* const volume value          */
} B2vpParamPatch;

/* Number of PCG32 streams of a (scene, seed_mode) pair and samples per stream. */
int b2vp_stream_count(const b2vp_scene_t *scene, int32_t seed_mode,
                      uint64_t *n_streams, uint32_t *samples_per_stream);

int b2vp_scene_create(const B2vpSceneDesc *desc, int32_t device, b2vp_scene_t **out);
int b2vp_scene_update(b2vp_scene_t *scene, const B2vpParamPatch *patch);
/* New values for a grid volume of unchanged resolution (what GridVolume::parameters_changed handles after
   traverse() hands out "data": src/textures/grid3d.cpp:366-391): the grid is overwritten in HBM, derived layouts
   are rebuilt on the device, the maximum (and the majorant of every heterogeneous medium using the grid as
   sigma_t) follows the data unless keep_max is set (the reference's fixed max_value).  n_values must equal
   res[0]*res[1]*res[2]*channels.                                               */
int b2vp_scene_update_grid(b2vp_scene_t *scene, int32_t volume, const float *data, uint64_t n_values,
                           int32_t keep_max);
void b2vp_scene_destroy(b2vp_scene_t *scene);

/* Renders the stream range into `film` (crop_w*crop_h*5 reals, XYZAW
   interleaved like ImageBlock storage, float for F32 / double for F64).       */
int b2vp_render(b2vp_scene_t *scene, const B2vpRenderArgs *args, void *film);

/* b2vp_render without any device allocation.  It has NO counterpart in the reference: it was added
   for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
   serving loops, harnesses that forbid it).  Same kernels and same results as b2vp_render, but the kernels'
   scratch (launch totals, stream work counter, spiral block table, the queue kernel's cold path records) lives
   in one static device array of B2VP_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until exit;
   the call performs no cudaMalloc / cudaFree and no allocating library routine.
   Conditions: `film_device` is the caller's DEVICE buffer (args->film_on_device = 1; cleared first when
   args->clear_film); the scene must already be resident in args->precision, i.e. b2vp_render ran once on it at
   set-up; seed modes SCALAR_BLOCKS and WAVEFRONT; no per-stream outputs (out_state / out_counters /
   out_radiance NULL).  A launch whose scratch exceeds the array fails with B2VP_ERR_UNSUPPORTED and a message
   naming the sizes (on a B200 the product launch needs 46-50 MB).  All calls on one device share the array and
   it keeps the previous call's contents (what must start at zero is cleared by the call): calls that overlap on
   the GPU -- two host threads, or two streams -- RACE; serialise them.  Synchronises args->cuda_stream before
   returning, like b2vp_render.                                                    */
#define B2VP_NOALLOC_SCRATCH_BYTES (64u << 20)
int b2vp_render_noalloc(b2vp_scene_t *scene, const B2vpRenderArgs *args, void *film_device);

/* What the library built from the scene and what it ran last: the acceleration structure that stands in for the
   reference's kd-tree (include/mitsuba/render/kdtree.h:2008-2391; only the traversal RESULT is part of the
   contract, so the tree's own figures are exposed for tests and logs), and the kernel instance the last
   b2vp_render on this scene launched -- the reference logs its choices through Log(Info, ...)
   (src/librender/integrator.cpp:99-104); this is the same information as data.  Fill abi_version / struct_size. */
typedef struct B2vpSceneInfo {
    uint32_t abi_version;
    uint32_t struct_size;
    uint64_t n_triangles;
    uint32_t bvh_nodes, bvh_leaves;
    uint32_t bvh_max_depth;              /* deepest leaf (root = 0): bounds the device traversal stack use */
    uint32_t bvh_forced_median_splits;   /* splits by index median because the tree was >= 24 levels deep  */
    uint32_t bvh_degenerate_splits;      /* splits by index median because all centroids coincided         */
    uint32_t bvh_stack_capacity;         /* entries of the device traversal stack                          */
    char last_kernel[96];                /* e.g. "queue/heterogeneous f32"; empty before the first render   */
} B2vpSceneInfo;
int b2vp_scene_info(const b2vp_scene_t *scene, B2vpSceneInfo *info);

/* How the streams [0, count) of a render are dealt out to n devices: interleaved blocks of `*block` streams (a power
   of two), device `index` taking blocks index, index + n, ...; `*count_i` = how many streams that is.  Pure host
   arithmetic (callable without a GPU): the partition b2vp_render uses for args->n_devices > 1, and the values to put
   into interleave_block / interleave_count / interleave_index when sharding by hand (one process per GPU).          */
int b2vp_shard_streams(uint64_t count, uint32_t n, uint32_t index, uint64_t *block, uint64_t *count_i);

/* Asynchronous cancel (Integrator::cancel, include/mitsuba/render/integrator.h:44-51): may be called from another
   thread while b2vp_render runs (it only sets a flag, no CUDA call).  The running render stops handing out new
   streams -- the persistent kernel polls a device flag where it fetches them -- lets the paths in flight finish and
   returns B2VP_ERR_CANCELLED; the film holds what was accumulated.  Like the reference's m_stop, the flag is cleared
   when the next b2vp_render starts: the handle stays usable, and a cancel between renders has no effect.           */
void b2vp_cancel(b2vp_scene_t *scene);

/* Thread-local message of the last failing call on this thread.               */
const char *b2vp_last_error(void);

/* Library / device introspection. */
uint32_t b2vp_abi_version(void);
int b2vp_device_count(void);

/* Roofline denominator for the FP32-bound configurations (not in
   MEASURED_PEAKS.json): runs a dependent-free FMA loop on every SM of `device` and
   returns the achieved non-tensor FP32 rate in TFLOP/s (FMA = 2 flop).           */
int b2vp_measure_fp32_peak(int32_t device, double *tflops);

#ifdef __cplusplus
}
#endif
#endif /* B2VP_H */
