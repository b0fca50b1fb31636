"""Generates tests/golden/ref_golden.npz by running the UNMODIFIED reference (oracle/_ref/refprobe, built from
/root/reference by oracle/Makefile.ref) on the scenes below.  Run from the repo root, in the container that has
/root/reference:

    make -f oracle/Makefile.ref -C oracle -j8 && python tests/golden/make_ref_golden.py

Each entry is the raw film storage (X, Y, Z, A, W per pixel) HDRFilm holds after SamplingIntegrator::render
with ONE thread (the result is thread-count independent: the sampler is seeded per block, integrator.cpp:198).
Most cases use 1 sample per pixel, so that a pixel is one path and the comparison is per path.
The CPU tests compare the oracle with these films, the -m gpu tests the CUDA path (tests/test_reference_parity.py)."""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402
import scenes  # noqa: E402


def _c2_homog(**kw):
    d = scenes.c2(**kw)
    d["atm"] = {"type": "homogeneous", "sigma_t": 0.5, "albedo": 0.9, "phase": {"type": "isotropic"}}
    return d


def _c2_sphere(**kw):
    """C2 with a diffuse sphere floating in the atmosphere (analytic sphere intersection, double precision)."""
    d = scenes.c2(**kw)
    d["ball"] = {"type": "sphere", "center": [0.5, -0.25, 0.5], "radius": 0.25,
                 "bsdf": {"type": "diffuse", "reflectance": 0.75}}
    return d


def _c1_radiancemeter(spp, variant):
    d = scenes.c1(spp=spp, variant=variant)
    d["sensor"] = {"type": "radiancemeter", "origin": [0.25, 0.5, 2.0], "direction": [-0.125, -0.25, -1.0],
                   "sampler": {"type": "independent", "sample_count": spp}, "film": scenes.film(1)}
    return d


def _c3(smooth=False, **kw):
    """C3 with a ray target that is NOT on the diagonal of the cube's top face: with scenes.c3's target
    (0.5, 0.5, 1) every camera ray passes exactly through the edge shared by the two triangles of that face and
    the hit / miss decision is a rounding coin toss in any implementation.
    smooth=True replaces the 1000:1 cloud / clear-air contrast by smooth low-contrast random fields: next to a
    cloud edge a position error of one float32 ulp changes the interpolated sigma_t by 1e-3 relative, so the
    cloud field is ill-conditioned in fp32 and gets a wider per-path tolerance (TOL_OVERRIDE)."""
    d = scenes.c3(**kw)
    d["sensor"]["ray_target"] = [0.4, 0.55, 1.0]
    # The sun is given by its to_world matrix instead of `direction`: directional.cpp:55 normalises `direction`
    # in single precision with enoki's rsqrt (the SSE estimate + one Newton step, CPU-vendor dependent in the last
    # ulp); a 1e-8 tilt of the sun is invisible in the plane-parallel cases but not in a 3D field in fp64.
    T = scenes.T
    dirn = np.array(d["sun"].pop("direction"), dtype=np.float64)
    up = np.array([0.0, 1.0, 0.0])
    d["sun"]["to_world"] = T.look_at([0, 0, 0], dirn, up, np.float32)
    if smooth:
        rng = np.random.default_rng(3)
        n = d["atm"]["sigma_t"]["data"].shape[0]
        d["atm"]["sigma_t"]["data"] = (5 + 5 * rng.random((n, n, n))).astype(np.float32)
        d["atm"]["albedo"]["data"] = (0.8 + 0.19 * rng.random((n, n, n))).astype(np.float32)
    return d


def _moment(d):
    """Wraps the scene's volpath in a `moment` integrator: 11 film channels."""
    inner = dict(d["integrator"])
    outer = {"type": "moment", "nested": inner}
    for key in ("block_size", "samples_per_pass"):
        if key in inner:
            outer[key] = inner.pop(key)
    d["integrator"] = outer
    return d


def _c2_edit(fn, **kw):
    d = scenes.c2(**kw)
    fn(d)
    return d


def _crop(d):
    d["sensor"]["film"].update({"crop_offset_x": 16, "crop_offset_y": 0, "crop_width": 40, "crop_height": 1})


def _seed7(d):
    d["sensor"]["sampler"]["seed"] = 7


def _wrap(mode):
    def f(d):
        # the grid covers the lower 40 % of the atmosphere only; above it the lookups wrap
        T = scenes.T
        v = d["atm"]["sigma_t"]
        v["wrap_mode"] = mode
        v["to_world"] = T.translate([-10, -10, 0], np.float32) @ T.scale([20, 20, 0.4], np.float32)
        v["data"] = v["data"][:64]
    return f


def _rgb_grid(d):
    g = d["atm"]["sigma_t"]["data"][:, :, :, None] if d["atm"]["sigma_t"]["data"].ndim == 3 else d["atm"]["sigma_t"]["data"]
    d["atm"]["sigma_t"]["data"] = np.concatenate([g * 0.5, g, g * 1.5], axis=3).astype(np.float32)


def _sensor_in_medium(spp, variant):
    """radiancemeter inside the homogeneous atmosphere of C1 (sensor->medium())."""
    d = scenes.c1(spp=spp, variant=variant)
    d["sensor"] = {"type": "radiancemeter", "origin": [0.125, -0.25, 0.5], "direction": [0.25, 0.125, -1.0],
                   "medium": {"type": "ref", "id": "atm"},
                   "sampler": {"type": "independent", "sample_count": spp}, "film": scenes.film(1)}
    return d


def _c2_obj(face_normals, **kw):
    """C2 with a small OBJ mesh (tests/golden/data/roof.obj: quads, normals, texcoords, shared vertices) floating in
    the atmosphere, loaded by the reference's obj plugin on one side and by scene.read_obj on the other."""
    d = scenes.c2(**kw)
    T = scenes.T
    d["roof"] = {"type": "obj", "filename": os.path.join(HERE, "data", "roof.obj"), "face_normals": face_normals,
                 "to_world": T.translate([0.25, -0.5, 0.125], np.float32) @ T.scale([2.0, 1.5, 1.0], np.float32),
                 "bsdf": {"type": "diffuse", "reflectance": 0.75}}
    return d


def _c2_cpp_host(variant):
    """The scene tests/cpp/test_host.cpp::scene_c2 builds with the C++ host layer (include/b2vp.hpp), as a dictionary."""
    dt = scenes.pkg.scene.VARIANTS[variant][0]
    T = scenes.T
    nz = 200
    grid = np.array([np.float32(np.exp(-((z + 0.5) / nz) * 100.0 / 8.0)) for z in range(nz)], dtype=np.float32).reshape(nz, 1, 1)
    return {
        "type": "scene",
        "integrator": {"type": "volpath", "max_depth": -1},
        "atm": {"type": "heterogeneous",
                "sigma_t": {"type": "gridvolume", "data": grid,
                            "to_world": T.translate([-10, -10, 0], dt) @ T.scale([20, 20, 1], dt)},
                "scale": 4.0, "albedo": 1.0, "phase": {"type": "rayleigh"}},
        "cube": {"type": "cube", "to_world": T.translate([0, 0, 0.45], dt) @ T.scale([10, 10, 0.55], dt),
                 "bsdf": {"type": "null"}, "interior": {"type": "ref", "id": "atm"}},
        "ground": {"type": "rectangle", "to_world": T.scale([10, 10, 1], dt),
                   "bsdf": {"type": "rpv", "rho_0": 0.1, "k": 0.7, "g": -0.1}},
        "sun": {"type": "directional", "direction": [0.5, 0, -0.8660254037844386], "irradiance": 1.0},
        "sensor": {"type": "distant", "direction": [0, 0, 1], "orientation": [1, 0, 0], "ray_target": [0, 0, 0],
                   "sampler": {"type": "independent", "sample_count": 1}, "film": scenes.film(256)},
    }


def _tabphase(values, **kw):
    """C2 with a tabulated phase function (src/phase/tabphase.cpp; SURVEY 8(f) N3: aerosol layers)."""
    d = scenes.c2(**kw)
    d["atm"]["phase"] = {"type": "tabphase", "values": values}
    return d


def _blendphase(grid_weight, **kw):
    """C2 with molecules (rayleigh) and aerosols (tabphase) mixed by `blendphase` (src/phase/blendphase.cpp): the
    aerosol fraction is a 1D profile (gridvolume) that leaves [0, 1] at both ends (clamped), or a constant."""
    d = scenes.c2(**kw)
    T = scenes.T
    if grid_weight:
        w = np.clip(np.exp(-np.linspace(0, 1, 50) * 6) * 1.3 - 0.05, -1, 2).astype(np.float32).reshape(50, 1, 1)
        weight = {"type": "gridvolume", "data": w,
                  "to_world": T.translate([-10, -10, 0], np.float32) @ T.scale([20, 20, 1], np.float32)}
        d["atm"]["phase"] = {"type": "blendphase", "a_mol": {"type": "rayleigh"},
                             "b_aer": {"type": "tabphase", "values": TAB_AEROSOL}, "weight": weight}
    else:
        d["atm"]["phase"] = {"type": "blendphase", "p0": {"type": "hg", "g": 0.5}, "p1": {"type": "isotropic"},
                             "weight": 0.25}
    return d


def _distantflux(target, **kw):
    """C2 seen by a `distantflux` sensor (src/sensors/distantflux.cpp): a 16 x 16 film over the
    hemisphere of a tilted reference normal, with a point target or none (bounding-sphere cross section)."""
    d = scenes.c2(**kw)
    T = scenes.T
    sn = {"type": "distantflux", "to_world": T.look_at([0, 0, 0], [0.125, -0.25, 1.0], [0, 1, 0], np.float32),
          "sampler": d["sensor"]["sampler"], "film": scenes.film(16, 16)}
    if target:
        sn["target"] = [0.25, -0.5, 0.0]
    d["sensor"] = sn
    return d


def _target_shape(d, key="ray_target", shape="ground"):
    """The sensor targets a rectangle of the scene instead of a point: area sampling of the shape with the aperture
    sample, weight 1 / (pdf * area) (distant.cpp:345-354, mdistant.cpp:247-252, distantflux.cpp:228-233)."""
    d["sensor"][key] = {"type": "ref", "id": shape}
    return d


def _nested_target(d, key="ray_target"):
    """The target is a rectangle of the sensor's own, nested in the sensor: sampled like any target shape, but not a
    child of the scene -- rays pass through it (it floats inside the atmosphere, tilted) and the scene's bounding
    sphere, which places the ray origins, does not know it."""
    T = scenes.T
    d["sensor"][key] = {"type": "rectangle", "to_world": T.translate([0.5, -0.25, 0.3], np.float32) @
                        T.rotate([1, 0, 0], 20.0, np.float32) @ T.scale([2.0, 1.5, 1], np.float32)}
    return d


def _origin_shape(d, key="ray_origin", half=15.0, height=1.2, sweep=False):
    """Ray origins on a rectangle of the sensor's own above the atmosphere instead of on the scene's bounding sphere
    (distant.cpp:281-290,366-375; distantflux.cpp:171-180,245-252): the target is projected onto it along the ray.
    `sweep`: no fixed direction, the film's columns sweep the directions of the XZ half plane -- with a small rectangle
    the slanted ones miss it, their ray weight is zero and the sample counts with alpha 0."""
    T = scenes.T
    d["sensor"][key] = {"type": "rectangle", "to_world": T.translate([0.25, -0.5, height], np.float32) @ T.scale([half, half, 1], np.float32)}
    if sweep:
        del d["sensor"]["direction"], d["sensor"]["orientation"]
    return d


def _three_emitters(d):
    d["sun2"] = {"type": "directional", "direction": [-0.4, 0.1, -1], "irradiance": 0.5}
    d["sky"] = {"type": "constant", "radiance": 0.2}
    return d


def _irradiancemeter(spp, variant):
    """C2 with an `irradiancemeter` (src/sensors/irradiancemeter.cpp) on a small rectangle floating
    in the atmosphere, facing up: downwelling irradiance at that level."""
    d = scenes.c2(spp=spp, variant=variant, width=1)
    T = scenes.T
    del d["sensor"]
    d["plate"] = {"type": "rectangle", "to_world": T.translate([0.5, -0.25, 0.25], np.float32) @ T.scale([0.5, 0.25, 1], np.float32),
                  "bsdf": {"type": "diffuse", "reflectance": 0.25},
                  "sensor": {"type": "irradiancemeter", "medium": {"type": "ref", "id": "atm"},
                             "sampler": {"type": "independent", "sample_count": spp}, "film": scenes.film(1)}}
    return d


def _mis(d, **props):
    """The scene with the `volpathmis` integrator (src/integrators/volpathmis.cpp)."""
    d["integrator"] = dict(d["integrator"], type="volpathmis", **props)
    return d


def _sky(d, radiance, sun=True):
    """The scene under a `constant` environment emitter (src/emitters/constant.cpp), with or without
    the sun: emitter sampling becomes non-delta (MIS with the BSDF), escaping rays see the sky."""
    d["sky"] = {"type": "constant", "radiance": radiance}
    if not sun:
        del d["sun"]
    return d


def _c2_ply(name, face_normals, kind="ply", shape_index=None, **kw):
    """C2 with a PLY mesh (tests/golden/data/roof_bin.ply: binary, normals, s/t texcoords; roof_ascii.ply: ascii,
    positions only), loaded by the reference's ply plugin on one side and by scene.read_ply on the other."""
    d = scenes.c2(**kw)
    T = scenes.T
    d["roof"] = {"type": kind, "filename": os.path.join(HERE, "data", name), "face_normals": face_normals,
                 "to_world": T.translate([0.25, -0.5, 0.125], np.float32) @ T.scale([2.0, 1.5, 1.0], np.float32),
                 "bsdf": {"type": "diffuse", "reflectance": 0.75}}
    if shape_index is not None:
        # the same mesh as sub-mesh 1 of a `.serialized` container (src/shapes/serialized.cpp; roof.serialized:
        # version 4, float32, normals + texcoords; roof_v3_f64.serialized: version 3, float64, positions only)
        d["roof"]["shape_index"] = shape_index
    return d


_MU = np.linspace(-1, 1, 181)
TAB_AEROSOL = ((1 - 0.49) / (1 + 0.49 - 1.4 * _MU) ** 1.5 + 0.3 * np.exp(-((_MU + 1) / 0.1) ** 2)).tolist()
TAB_COARSE = [0.0, 0.0, 1.0, 0.5, 2.0]   # an interval without probability mass, a flat-to-linear mix

_BILAMB = {"type": "bilambertian", "reflectance": 0.5, "transmittance": 0.25}  # two-sided leaves

# per-path tolerance overrides (default: tests/test_reference_parity.py TOL): (tolerance, minimum fraction)
TOL_OVERRIDE = {
    "c3_cloud": (1e-3, 0.97),
    "c3_smooth": (3e-5, 0.97),
    "c3_volpathmis": (3e-5, 0.97),
}

CASES = {
    # name: (builder, variant)
    "c1": (lambda v: scenes.c1(spp=4096, variant=v), "scalar_mono"),
    "c1_sun30": (lambda v: scenes.c1(spp=1024, variant=v, sun=(0, 0.5, -1)), "scalar_mono"),
    "c1_radiancemeter": (lambda v: _c1_radiancemeter(512, v), "scalar_mono"),
    "c2_paths": (lambda v: scenes.c2(spp=1, variant=v, width=256), "scalar_mono"),
    "c2_16spp": (lambda v: scenes.c2(spp=16, variant=v, width=64), "scalar_mono"),
    "c2_tau2": (lambda v: scenes.c2(spp=1, variant=v, width=128, tau=2.0), "scalar_mono"),
    "c2_nearest": (lambda v: scenes.c2(spp=1, variant=v, width=128, filter_type="nearest"), "scalar_mono"),
    "c2_mdistant": (lambda v: scenes.c2(spp=1, variant=v, width=128, sensor="mdistant"), "scalar_mono"),
    "c2_homog_iso": (lambda v: _c2_homog(spp=1, variant=v, width=128), "scalar_mono"),
    "c2_sphere": (lambda v: _c2_sphere(spp=1, variant=v, width=128), "scalar_mono"),
    "c2_depth3": (lambda v: scenes.c2(spp=1, variant=v, width=128, max_depth=3, rr_depth=2), "scalar_mono"),
    "c3_cloud": (lambda v: _c3(spp=1, variant=v, n=16, size=32), "scalar_mono"),
    "c3_smooth": (lambda v: _c3(smooth=True, spp=1, variant=v, n=16, size=32), "scalar_mono"),
    "c5_paths": (lambda v: scenes.c5(spp=1, variant=v, n_leaves=200, n_sensors=256), "scalar_mono"),
    "c5_bilambertian": (lambda v: scenes.c5(spp=1, variant=v, n_leaves=200, n_sensors=256, leaf_bsdf=_BILAMB), "scalar_mono"),
    "c5_bilambertian_f64": (lambda v: scenes.c5(spp=1, variant=v, n_leaves=200, n_sensors=256, leaf_bsdf=_BILAMB),
                            "scalar_mono_double"),
    "c2_moment": (lambda v: _moment(scenes.c2(spp=1, variant=v, width=128)), "scalar_mono"),
    "c2_moment_256spp": (lambda v: _moment(scenes.c2(spp=256, variant=v, width=32)), "scalar_mono"),
    "c2_moment_f64": (lambda v: _moment(scenes.c2(spp=2, variant=v, width=64)), "scalar_mono_double"),
    # edge cases of the render driver and the volume lookups
    "c2_crop": (lambda v: _c2_edit(_crop, spp=1, variant=v, width=64), "scalar_mono"),
    "c2_seed7": (lambda v: _c2_edit(_seed7, spp=1, variant=v, width=128), "scalar_mono"),
    "c2_passes": (lambda v: scenes.c2(spp=8, variant=v, width=48, samples_per_pass=2, block_size=16), "scalar_mono"),
    "c2_maxdepth1": (lambda v: scenes.c2(spp=1, variant=v, width=128, max_depth=1), "scalar_mono"),
    "c2_rr1": (lambda v: scenes.c2(spp=1, variant=v, width=128, rr_depth=1), "scalar_mono"),
    "c2_hide_emitters": (lambda v: scenes.c2(spp=1, variant=v, width=64, hide_emitters=True), "scalar_mono"),
    "c2_notspectral": (lambda v: scenes.c2(spp=1, variant=v, width=128, spectral=False), "scalar_mono"),
    "c2_repeat": (lambda v: _c2_edit(_wrap("repeat"), spp=1, variant=v, width=128, nz=160), "scalar_mono"),
    "c2_mirror": (lambda v: _c2_edit(_wrap("mirror"), spp=1, variant=v, width=128, nz=160), "scalar_mono"),
    "c2_rgb_grid": (lambda v: _c2_edit(_rgb_grid, spp=1, variant=v, width=128, nz=100), "scalar_mono"),
    "c1_sensor_in_medium": (lambda v: _sensor_in_medium(512, v), "scalar_mono"),
    "c3_blocks_f64": (lambda v: _c3(spp=2, variant=v, n=16, size=20, block_size=8, samples_per_pass=1), "scalar_mono_double"),
    "c2_obj_normals": (lambda v: _c2_obj(False, spp=1, variant=v, width=128), "scalar_mono"),
    "c2_obj_face_normals": (lambda v: _c2_obj(True, spp=1, variant=v, width=128), "scalar_mono"),
    "c2_obj_normals_f64": (lambda v: _c2_obj(False, spp=1, variant=v, width=64), "scalar_mono_double"),
    "c2_cpp_host": (_c2_cpp_host, "scalar_mono"),
    "c2_cpp_host_f64": (_c2_cpp_host, "scalar_mono_double"),
    "c2_tabphase": (lambda v: _tabphase(TAB_AEROSOL, spp=1, variant=v, width=256), "scalar_mono"),
    "c2_tabphase_coarse": (lambda v: _tabphase(TAB_COARSE, spp=1, variant=v, width=256), "scalar_mono"),
    "c2_tabphase_f64": (lambda v: _tabphase(TAB_AEROSOL, spp=1, variant=v, width=128), "scalar_mono_double"),
    "c2_blendphase": (lambda v: _blendphase(True, spp=1, variant=v, width=256), "scalar_mono"),
    "c2_blendphase_const": (lambda v: _blendphase(False, spp=1, variant=v, width=256), "scalar_mono"),
    "c2_blendphase_f64": (lambda v: _blendphase(True, spp=1, variant=v, width=128), "scalar_mono_double"),
    "c2_distantflux": (lambda v: _distantflux(True, spp=1, variant=v), "scalar_mono"),
    "c2_distantflux_notarget": (lambda v: _distantflux(False, spp=1, variant=v), "scalar_mono"),
    "c2_distantflux_f64": (lambda v: _distantflux(True, spp=1, variant=v), "scalar_mono_double"),
    "c2_irradiancemeter": (lambda v: _irradiancemeter(256, v), "scalar_mono"),
    "c2_irradiancemeter_f64": (lambda v: _irradiancemeter(64, v), "scalar_mono_double"),
    "c2_volpathmis": (lambda v: _mis(scenes.c2(spp=1, variant=v, width=256)), "scalar_mono"),
    "c2_volpathmis_notspectral": (lambda v: _mis(scenes.c2(spp=1, variant=v, width=128, spectral=False)), "scalar_mono"),
    "c2_volpathmis_depth": (lambda v: _mis(scenes.c2(spp=1, variant=v, width=128, max_depth=4, rr_depth=2), use_spectral_mis=False), "scalar_mono"),
    "c3_volpathmis": (lambda v: _mis(_c3(smooth=True, spp=1, variant=v, n=16, size=16)), "scalar_mono"),
    "c5_volpathmis": (lambda v: _mis(scenes.c5(spp=1, variant=v, n_leaves=200, n_sensors=256, leaf_bsdf=_BILAMB)), "scalar_mono"),
    "c1_volpathmis_f64": (lambda v: _mis(scenes.c1(spp=1024, variant=v)), "scalar_mono_double"),
    "c2_volpathmis_f64": (lambda v: _mis(scenes.c2(spp=1, variant=v, width=128)), "scalar_mono_double"),
    "c3_volpathmis_f64": (lambda v: _mis(_c3(spp=1, variant=v, n=16, size=16)), "scalar_mono_double"),
    "c2_sky": (lambda v: _sky(scenes.c2(spp=1, variant=v, width=256), 0.25), "scalar_mono"),
    "c2_sky_only": (lambda v: _sky(scenes.c2(spp=1, variant=v, width=256), 1.0, sun=False), "scalar_mono"),
    "c5_sky": (lambda v: _sky(scenes.c5(spp=1, variant=v, n_leaves=200, n_sensors=256, leaf_bsdf=_BILAMB), 0.5), "scalar_mono"),
    "c2_sky_f64": (lambda v: _sky(scenes.c2(spp=1, variant=v, width=128), 0.25), "scalar_mono_double"),
    "c3_sky_f64": (lambda v: _sky(_c3(spp=1, variant=v, n=16, size=16), 0.25, sun=False), "scalar_mono_double"),
    "c2_sky_volpathmis": (lambda v: _mis(_sky(scenes.c2(spp=1, variant=v, width=256), 0.25)), "scalar_mono"),
    "c5_sky_volpathmis": (lambda v: _mis(_sky(scenes.c5(spp=1, variant=v, n_leaves=200, n_sensors=256), 0.5, sun=False)), "scalar_mono"),
    "c2_sky_volpathmis_f64": (lambda v: _mis(_sky(scenes.c2(spp=1, variant=v, width=128), 1.0, sun=False)), "scalar_mono_double"),
    "c2_ply_binary": (lambda v: _c2_ply("roof_bin.ply", False, spp=1, variant=v, width=128), "scalar_mono"),
    "c2_ply_ascii": (lambda v: _c2_ply("roof_ascii.ply", True, spp=1, variant=v, width=128), "scalar_mono"),
    "c2_serialized": (lambda v: _c2_ply("roof.serialized", False, "serialized", 1, spp=1, variant=v, width=128), "scalar_mono"),
    "c2_serialized_v3": (lambda v: _c2_ply("roof_v3_f64.serialized", True, "serialized", 1, spp=1, variant=v, width=64), "scalar_mono_double"),
    "c2_ply_binary_f64": (lambda v: _c2_ply("roof_bin.ply", False, spp=1, variant=v, width=64), "scalar_mono_double"),
    # sensors that target a shape (the rectangle on the ground of C2, the null rectangle above the cloud of C3)
    "c2_target_shape": (lambda v: _target_shape(scenes.c2(spp=1, variant=v, width=128)), "scalar_mono"),
    "c2_target_shape_f64": (lambda v: _target_shape(scenes.c2(spp=2, variant=v, width=64)), "scalar_mono_double"),
    "c3_target_shape_f64": (lambda v: _target_shape(_c3(spp=1, variant=v, n=16, size=16), shape="top"), "scalar_mono_double"),
    "c2_mdistant_target_shape": (lambda v: _target_shape(scenes.c2(spp=1, variant=v, width=128, sensor="mdistant"), "target"), "scalar_mono"),
    "c2_mdistant_target_shape_f64": (lambda v: _target_shape(scenes.c2(spp=1, variant=v, width=64, sensor="mdistant"), "target"), "scalar_mono_double"),
    "c2_distantflux_target_shape": (lambda v: _target_shape(_distantflux(False, spp=1, variant=v), "target"), "scalar_mono"),
    "c2_distantflux_target_shape_f64": (lambda v: _target_shape(_distantflux(False, spp=1, variant=v), "target"), "scalar_mono_double"),
    "c2_nested_target": (lambda v: _nested_target(scenes.c2(spp=1, variant=v, width=128)), "scalar_mono"),
    "c2_nested_target_f64": (lambda v: _nested_target(scenes.c2(spp=2, variant=v, width=64)), "scalar_mono_double"),
    "c2_mdistant_nested_target_f64": (lambda v: _nested_target(scenes.c2(spp=1, variant=v, width=64, sensor="mdistant"), "target"), "scalar_mono_double"),
    "c2_distantflux_nested_target": (lambda v: _nested_target(_distantflux(False, spp=1, variant=v), "target"), "scalar_mono"),
    # ray origins on a shape
    "c2_origin_shape": (lambda v: _origin_shape(scenes.c2(spp=1, variant=v, width=128)), "scalar_mono"),
    "c2_origin_shape_f64": (lambda v: _origin_shape(_target_shape(scenes.c2(spp=2, variant=v, width=64))), "scalar_mono_double"),
    "c2_origin_shape_miss": (lambda v: _origin_shape(scenes.c2(spp=1, variant=v, width=128), half=1.0, sweep=True), "scalar_mono"),
    "c2_origin_shape_miss_f64": (lambda v: _origin_shape(scenes.c2(spp=2, variant=v, width=64), half=1.0, sweep=True), "scalar_mono_double"),
    "c3_origin_ref_f64": (lambda v: _target_shape(_c3(spp=1, variant=v, n=16, size=16), "ray_origin", "top"), "scalar_mono_double"),
    "c2_distantflux_origin": (lambda v: _origin_shape(_distantflux(True, spp=1, variant=v), "origin", half=2.0), "scalar_mono"),
    "c2_distantflux_origin_f64": (lambda v: _origin_shape(_distantflux(False, spp=1, variant=v), "origin", half=4.0), "scalar_mono_double"),
    # three emitters: `ScalarFloat emitter_pdf = 1.f / m_emitters.size()` (scene.cpp:183,233) is a FLOAT quotient in
    # every variant -- (double) (1.f / 3) in scalar_mono_double, 3e-8 away from 1 / 3 (found by the random-scene test)
    "c2_three_emitters_f64": (lambda v: _three_emitters(scenes.c2(spp=2, variant=v, width=64)), "scalar_mono_double"),
    "c2_three_emitters_volpathmis_f64": (lambda v: _mis(_three_emitters(scenes.c2(spp=2, variant=v, width=64))), "scalar_mono_double"),
    "c5_three_emitters": (lambda v: _three_emitters(scenes.c5(spp=1, variant=v, n_leaves=200, n_sensors=256)), "scalar_mono"),
    # volpathmis carries the pdf phase->sample() returns into the MIS weight of the next non-delta emitter hit; for
    # tabphase that pdf is the table read at the MIRRORED cosine (tabphase.cpp:79 against :94-96)
    "c2_tabphase_sky_volpathmis_f64": (lambda v: _mis(_sky(_tabphase(TAB_AEROSOL, spp=2, variant=v, width=64), 0.25)), "scalar_mono_double"),
    "c2_tabphase_sky_volpathmis": (lambda v: _mis(_sky(_tabphase(TAB_AEROSOL, spp=1, variant=v, width=128), 0.25, sun=False)), "scalar_mono"),
    "c1_f64": (lambda v: scenes.c1(spp=1024, variant=v), "scalar_mono_double"),
    "c2_paths_f64": (lambda v: scenes.c2(spp=1, variant=v, width=128), "scalar_mono_double"),
    "c2_sphere_f64": (lambda v: _c2_sphere(spp=1, variant=v, width=64), "scalar_mono_double"),
    "c3_cloud_f64": (lambda v: _c3(spp=1, variant=v, n=16, size=16), "scalar_mono_double"),
    "c3_smooth_f64": (lambda v: _c3(smooth=True, spp=1, variant=v, n=16, size=16), "scalar_mono_double"),
    "c5_paths_f64": (lambda v: scenes.c5(spp=1, variant=v, n_leaves=200, n_sensors=64), "scalar_mono_double"),
}


def _rfilter(d, rfilter, width, height, **film):
    """The scene with a two-dimensional film under a reconstruction filter wider than a pixel
    (src/rfilters/*.cpp, src/librender/imageblock.cpp:117-152; VERDICT missing #6)."""
    d["sensor"]["film"] = dict({"type": "hdrfilm", "width": width, "height": height, "rfilter": rfilter}, **film)
    return d


_CROP = {"crop_offset_x": 3, "crop_offset_y": 2, "crop_width": 17, "crop_height": 7}

# films under wide reconstruction filters: weights and alpha are sums of products, not counts, so these are compared
# by tests/test_rfilter.py with its own check (not by check_against_reference, which wants integer weights).  Single
# precision: one sample per pixel mostly, so that a path that diverges from the reference binary
# cannot move the position of the next sample of its stream and with it a whole footprint of weights
RFILTER_CASES = {
    "rf_gaussian": (lambda v: _rfilter(scenes.c2(spp=2, variant=v, block_size=8), {"type": "gaussian"}, 24, 10), "scalar_mono"),
    "rf_gaussian_f64": (lambda v: _rfilter(scenes.c2(spp=2, variant=v, block_size=8), {"type": "gaussian"}, 24, 10), "scalar_mono_double"),
    "rf_gaussian_wide_f64": (lambda v: _rfilter(scenes.c2(spp=1, variant=v), {"type": "gaussian", "stddev": 0.8}, 16, 12), "scalar_mono_double"),
    "rf_tent_crop_f64": (lambda v: _rfilter(scenes.c2(spp=2, variant=v, block_size=4), {"type": "tent"}, 24, 12, **_CROP), "scalar_mono_double"),
    "rf_tent_crop": (lambda v: _rfilter(scenes.c2(spp=1, variant=v, block_size=4), {"type": "tent"}, 24, 12, **_CROP), "scalar_mono"),
    "rf_mitchell": (lambda v: _rfilter(scenes.c2(spp=1, variant=v, block_size=16), {"type": "mitchell"}, 20, 9), "scalar_mono"),
    "rf_mitchell_bc_f64": (lambda v: _rfilter(scenes.c2(spp=1, variant=v), {"type": "mitchell", "B": 0.2, "C": 0.4}, 20, 9), "scalar_mono_double"),
    "rf_catmullrom_f64": (lambda v: _rfilter(scenes.c2(spp=1, variant=v, block_size=8), {"type": "catmullrom"}, 20, 9), "scalar_mono_double"),
    "rf_lanczos_f64": (lambda v: _rfilter(scenes.c2(spp=1, variant=v, block_size=8), {"type": "lanczos"}, 20, 9), "scalar_mono_double"),
    "rf_lanczos2": (lambda v: _rfilter(scenes.c2(spp=1, variant=v), {"type": "lanczos", "lobes": 2}, 20, 9), "scalar_mono"),
    "rf_box_wide": (lambda v: _rfilter(scenes.c2(spp=2, variant=v, block_size=4), {"type": "box", "radius": 1.2}, 13, 10), "scalar_mono"),
    "rf_gaussian_moment": (lambda v: _moment(_rfilter(scenes.c2(spp=1, variant=v, block_size=8), {"type": "gaussian"}, 16, 8)), "scalar_mono"),
    "rf_gaussian_volpathmis_f64": (lambda v: _mis(_rfilter(scenes.c2(spp=1, variant=v), {"type": "gaussian"}, 16, 8)), "scalar_mono_double"),
    "rf_c5_tent_row": (lambda v: _rfilter(scenes.c5(spp=1, variant=v, n_leaves=200, n_sensors=64), {"type": "tent"}, 64, 1), "scalar_mono"),
    "rf_c3_gaussian_f64": (lambda v: _rfilter(_c3(spp=1, variant=v, n=16, size=16, block_size=8), {"type": "gaussian"}, 16, 16), "scalar_mono_double"),
}
ALL_CASES = dict(CASES, **RFILTER_CASES)


def scene_dict(name):
    """The dictionary both sides load: every number rounded to float32 (what the reference's XML parser and
    Properties keep, xml.cpp:100-103, properties.h:55)."""
    from oracle import ref_xml
    fn, variant = ALL_CASES[name]
    return ref_xml.round_f32(fn(variant)), variant


def main():
    from oracle import ref_xml
    pkg = scenes.pkg
    if not ref_xml.available():
        ref_xml.build()
    path = os.path.join(HERE, "ref_golden.npz")
    only = sys.argv[1:]   # names given: render those and keep the other films of the existing file
    out = dict(np.load(path)) if only else {}
    for name in (only or ALL_CASES):
        d, variant = scene_dict(name)
        film, sec = ref_xml.render(d, variant, threads=1, write_vol=pkg.scene.write_vol)
        out[name] = film
        print("%-18s %-20s %4d x %-4d mean %.7g  (%.3f s)" % (name, variant, film.shape[1], film.shape[0],
                                                             film[..., 1].sum() / film[..., 4].sum(), sec))
    np.savez_compressed(path, **{k: out[k] for k in ALL_CASES})
    print("wrote", len(out), "reference films")


if __name__ == "__main__":
    main()
