"""ctypes mirror of include/b2vp.h and loader of the CUDA library.

The product library is ``libb2vp.so`` (built by ``__graft_entry__.build()`` from
``csrc/``).  There is no CPU fallback: if the library is missing, loading raises.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libb2vp.so")

ABI_VERSION = 4

OK, ERR_INVALID, ERR_UNSUPPORTED, ERR_CUDA, ERR_NO_DEVICE, ERR_CANCELLED = range(6)
PREC_F32, PREC_F64 = 0, 1
SHAPE_RECTANGLE, SHAPE_SPHERE, SHAPE_MESH = 0, 1, 2
BSDF_DIFFUSE, BSDF_NULL, BSDF_RPV, BSDF_BILAMBERTIAN = 0, 1, 2, 3
PHASE_TABULATED, PHASE_BLEND = 3, 4
EMITTER_CONSTANT = 1
MEDIUM_HOMOGENEOUS, MEDIUM_HETEROGENEOUS = 0, 1
VOLUME_CONST, VOLUME_GRID = 0, 1
FILTER_NEAREST, FILTER_TRILINEAR = 0, 1
WRAP_REPEAT, WRAP_MIRROR, WRAP_CLAMP = 0, 1, 2
PHASE_ISOTROPIC, PHASE_RAYLEIGH, PHASE_HG = 0, 1, 2
EMITTER_DIRECTIONAL = 0
SENSOR_DISTANT, SENSOR_MDISTANT, SENSOR_RADIANCEMETER, SENSOR_MRADIANCEMETER, SENSOR_DISTANTFLUX = 0, 1, 2, 3, 4
SENSOR_IRRADIANCEMETER = 5
DIR_SINGLE, DIR_SAMPLE_WIDTH, DIR_SAMPLE_ALL = 0, 1, 2
TARGET_NONE, TARGET_POINT, TARGET_SHAPE = 0, 1, 2
ORIGIN_BSPHERE, ORIGIN_SHAPE = 0, 1
SEED_SCALAR_BLOCKS, SEED_WAVEFRONT, SEED_EXPLICIT = 0, 1, 2
MATH_EXACT, MATH_FAST = 0, 1
KERNEL_AUTO, KERNEL_QUEUE_GENERIC, KERNEL_QUEUE_256, KERNEL_LANES, KERNEL_STREAM = range(5)
KERNELS = {"auto": KERNEL_AUTO, "queue": KERNEL_AUTO, "generic": KERNEL_QUEUE_GENERIC, "q256": KERNEL_QUEUE_256,
           "lanes": KERNEL_LANES, "mega": KERNEL_STREAM, "stream": KERNEL_STREAM}
OPT_NO_CELL_RECORDS, OPT_BVH_LEAF_SHIFT = 1, 8

c_d, c_i32, c_u32, c_u64 = C.c_double, C.c_int32, C.c_uint32, C.c_uint64


class Mat4(C.Structure):
    _fields_ = [("m", c_d * 16)]


class Shape(C.Structure):
    _fields_ = [
        ("type", c_i32), ("bsdf", c_i32), ("interior_medium", c_i32), ("exterior_medium", c_i32),
        ("to_world", Mat4), ("to_object", Mat4),
        ("frame_s", c_d * 3), ("frame_t", c_d * 3), ("frame_n", c_d * 3),
        ("inv_surface_area", c_d), ("center", c_d * 3), ("radius", c_d),
        ("flip_normals", c_i32),
        ("vertex_offset", c_u32), ("vertex_count", c_u32), ("face_offset", c_u32), ("face_count", c_u32),
        ("has_vertex_normals", c_i32), ("has_vertex_texcoords", c_i32), ("sample_only", c_i32),
    ]


class Bsdf(C.Structure):
    _fields_ = [("type", c_i32), ("reserved0", c_i32), ("p", c_d * 4)]


class Volume(C.Structure):
    _fields_ = [
        ("type", c_i32), ("filter", c_i32), ("wrap", c_i32), ("channels", c_i32),
        ("res", c_i32 * 3), ("reserved0", c_i32),
        ("value", c_d), ("max_value", c_d), ("bbox_min", c_d * 3), ("bbox_max", c_d * 3),
        ("world_to_local", Mat4), ("data", C.POINTER(C.c_float)),
    ]


class Phase(C.Structure):
    _fields_ = [("type", c_i32), ("reserved0", c_i32), ("g", c_d),
                ("n_values", c_u32), ("reserved1", c_u32), ("values", C.POINTER(c_d)),
                ("child", c_i32 * 2), ("weight_volume", c_i32), ("reserved2", c_i32)]


class Medium(C.Structure):
    _fields_ = [
        ("type", c_i32), ("sigma_t", c_i32), ("albedo", c_i32), ("phase", c_i32),
        ("has_spectral_extinction", c_i32), ("sample_emitters", c_i32), ("scale", c_d),
    ]


class Emitter(C.Structure):
    _fields_ = [("type", c_i32), ("reserved0", c_i32), ("direction", c_d * 3), ("irradiance", c_d)]


RFILTER_RESOLUTION = 31   # MTS_FILTER_RESOLUTION (include/mitsuba/core/rfilter.h:10)
RFILTER_MAX_TAPS = 16


class Sensor(C.Structure):
    _fields_ = [
        ("type", c_i32), ("direction_type", c_i32), ("target_type", c_i32), ("origin_type", c_i32),
        ("target_shape", c_i32), ("origin_shape", c_i32), ("flip_directions", c_i32),
        ("needs_aperture_sample", c_i32), ("medium", c_i32), ("n_transforms", c_i32),
        ("to_world", Mat4), ("target_point", c_d * 3), ("reference_normal", c_d * 3),
        ("shutter_open", c_d), ("shutter_open_time", c_d),
        ("transforms", C.POINTER(c_d)),
        ("film_size", c_i32 * 2), ("crop_offset", c_i32 * 2), ("crop_size", c_i32 * 2),
        ("sample_count", c_u32), ("reserved0", c_u32), ("sampler_seed", c_u64),
        ("rfilter_radius", c_d), ("rfilter_scale", c_d), ("rfilter_taps", c_u32), ("rfilter_border", c_u32),
        ("rfilter_values", c_d * (RFILTER_RESOLUTION + 1)),
    ]


class Integrator(C.Structure):
    _fields_ = [
        ("max_depth", c_i32), ("rr_depth", c_i32), ("hide_emitters", c_i32),
        ("block_size", c_u32), ("samples_per_pass", c_u32), ("arg_order_ltr", c_u32), ("moment", c_u32), ("mis", c_u32),
    ]


class SceneDesc(C.Structure):
    _fields_ = [
        ("abi_version", c_u32), ("struct_size", c_u32), ("precision", c_i32), ("options", c_u32),
        ("n_shapes", c_u32), ("n_bsdfs", c_u32), ("n_media", c_u32), ("n_volumes", c_u32),
        ("n_phases", c_u32), ("n_emitters", c_u32),
        ("shapes", C.POINTER(Shape)), ("bsdfs", C.POINTER(Bsdf)), ("media", C.POINTER(Medium)),
        ("volumes", C.POINTER(Volume)), ("phases", C.POINTER(Phase)), ("emitters", C.POINTER(Emitter)),
        ("n_vertices", c_u32), ("n_faces", c_u32),
        ("vertex_positions", C.POINTER(C.c_float)), ("vertex_normals", C.POINTER(C.c_float)),
        ("vertex_texcoords", C.POINTER(C.c_float)), ("faces", C.POINTER(c_u32)),
        ("bbox_min", c_d * 3), ("bbox_max", c_d * 3), ("bsphere_center", c_d * 3), ("bsphere_radius", c_d),
        ("sensor", Sensor), ("integrator", Integrator),
    ]


class PathCounters(C.Structure):
    _fields_ = [("n_rng_draws", c_u32), ("n_medium_steps", c_u32), ("n_null", c_u32), ("n_real", c_u32),
                ("n_intersect", c_u32), ("n_surface", c_u32)]


class Counters(C.Structure):
    _fields_ = [
        ("n_paths", c_u64), ("n_rng_draws", c_u64), ("n_medium_steps", c_u64), ("n_null", c_u64),
        ("n_real", c_u64), ("n_intersect", c_u64), ("n_surface", c_u64), ("n_invalid_samples", c_u64),
        ("gpu_ms", c_d), ("n_launches", c_u32), ("stopped_early", c_u32),
        ("n_bvh_nodes", c_u64), ("n_bvh_tris", c_u64),
    ]


class RenderArgs(C.Structure):
    _fields_ = [
        ("abi_version", c_u32), ("struct_size", c_u32), ("precision", c_i32), ("seed_mode", c_i32),
        ("device", c_i32), ("film_on_device", c_i32), ("cuda_stream", C.c_void_p),
        ("first_stream", c_u64), ("n_streams", c_u64),
        ("interleave_block", c_u64), ("interleave_count", c_u32), ("interleave_index", c_u32),
        ("explicit_state", C.POINTER(c_u64)), ("explicit_inc", C.POINTER(c_u64)),
        ("explicit_pixel", C.POINTER(c_u32)), ("explicit_samples_per_stream", c_u32), ("clear_film", c_u32),
        ("out_state", C.POINTER(c_u64)), ("out_counters", C.POINTER(PathCounters)),
        ("out_radiance", C.c_void_p), ("totals", C.POINTER(Counters)),
        ("math_mode", c_i32), ("kernel", c_i32), ("tuning", c_u32 * 4), ("debug_flags", c_u32), ("majorant_grid", c_u32),
        ("timeout_s", c_d), ("devices", C.POINTER(c_i32)), ("n_devices", c_u32), ("reserved2", c_u32),
    ]


class ParamPatch(C.Structure):
    _fields_ = [
        ("abi_version", c_u32), ("struct_size", c_u32),
        ("medium", c_i32), ("medium_scale", c_d),
        ("bsdf", c_i32), ("bsdf_p", c_d * 4),
        ("emitter", c_i32), ("emitter_irradiance", c_d),
        ("volume", c_i32), ("volume_value", c_d),
    ]


class SceneInfo(C.Structure):
    _fields_ = [
        ("abi_version", c_u32), ("struct_size", c_u32), ("n_triangles", c_u64),
        ("bvh_nodes", c_u32), ("bvh_leaves", c_u32), ("bvh_max_depth", c_u32),
        ("bvh_forced_median_splits", c_u32), ("bvh_degenerate_splits", c_u32), ("bvh_stack_capacity", c_u32),
        ("last_kernel", C.c_char * 96),
    ]


# every symbol include/b2vp.h declares
EXPORTS = ["b2vp_stream_count", "b2vp_scene_create", "b2vp_scene_update", "b2vp_scene_update_grid", "b2vp_scene_destroy", "b2vp_render",
           "b2vp_cancel", "b2vp_last_error", "b2vp_abi_version", "b2vp_device_count", "b2vp_measure_fp32_peak",
           "b2vp_render_noalloc", "b2vp_scene_info", "b2vp_shard_streams"]

_lib = None


class B2vpError(RuntimeError):
    """Raised for a non-zero C-ABI status (the reference's Throw -> std::runtime_error)."""

    def __init__(self, code, message):
        super().__init__("b2vp error %d: %s" % (code, message))
        self.code = code


def lib():
    """Load libb2vp.so (once).  Raises if the CUDA extension has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError("CUDA library %s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU fallback)" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    L.b2vp_last_error.restype = C.c_char_p
    L.b2vp_abi_version.restype = c_u32
    L.b2vp_device_count.restype = C.c_int
    L.b2vp_scene_create.argtypes = [C.POINTER(SceneDesc), c_i32, C.POINTER(C.c_void_p)]
    L.b2vp_scene_update.argtypes = [C.c_void_p, C.POINTER(ParamPatch)]
    L.b2vp_scene_update_grid.argtypes = [C.c_void_p, c_i32, C.POINTER(C.c_float), c_u64, c_i32]
    L.b2vp_scene_destroy.argtypes = [C.c_void_p]
    L.b2vp_scene_destroy.restype = None
    L.b2vp_render.argtypes = [C.c_void_p, C.POINTER(RenderArgs), C.c_void_p]
    L.b2vp_render_noalloc.argtypes = [C.c_void_p, C.POINTER(RenderArgs), C.c_void_p]
    L.b2vp_scene_info.argtypes = [C.c_void_p, C.POINTER(SceneInfo)]
    L.b2vp_shard_streams.argtypes = [c_u64, c_u32, c_u32, C.POINTER(c_u64), C.POINTER(c_u64)]
    L.b2vp_cancel.argtypes = [C.c_void_p]
    L.b2vp_cancel.restype = None
    L.b2vp_stream_count.argtypes = [C.c_void_p, c_i32, C.POINTER(c_u64), C.POINTER(c_u32)]
    L.b2vp_measure_fp32_peak.argtypes = [c_i32, C.POINTER(c_d)]
    if L.b2vp_abi_version() != ABI_VERSION:
        raise RuntimeError("libb2vp.so ABI %d != python mirror %d" % (L.b2vp_abi_version(), ABI_VERSION))
    _lib = L
    return L


def check(code):
    if code != OK:
        raise B2vpError(code, lib().b2vp_last_error().decode())
