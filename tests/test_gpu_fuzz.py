"""Differential test over random combinations of the scoped plugins: every scene is drawn from a seeded generator
(medium class, phase function, surface BSDFs, emitters, sensor class with its target / origin variants, film with crop
window and reconstruction filter, integrator and its limits, seeding mode) and rendered by the CUDA path and by the
oracle; PCG32 end states, event counters and per-path radiance must agree bit for bit, films to summation order.
The named parity cases cover each feature on its own; this covers the products nobody wrote down."""
import os

import numpy as np
import pytest

import scenes

DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "data")

pytestmark = pytest.mark.gpu

T = scenes.T
TAB = [0.9, 0.5, 0.3, 0.25, 0.3, 0.6, 1.4, 3.0]


def random_scene(seed, variant):
    rng = np.random.default_rng(seed)
    pick = lambda *xs: xs[int(rng.integers(len(xs)))]
    dt = scenes.pkg.scene.VARIANTS[variant][0]
    d = {"type": "scene"}
    # ---- integrator
    integ = {"type": pick("volpath", "volpath", "volpathmis"), "max_depth": pick(-1, -1, 2, 5), "rr_depth": pick(5, 1, 3),
             "hide_emitters": bool(pick(False, False, True))}
    if integ["type"] == "volpathmis":
        integ["use_spectral_mis"] = bool(pick(True, False))
    if pick(False, False, True):
        integ["block_size"] = int(pick(2, 4))
    if pick(False, False, True):
        integ = {"type": "moment", "nested": integ}
    passes = bool(pick(False, False, True))
    d["integrator"] = integ
    # ---- phase function
    def phase(depth=0):
        kind = pick("rayleigh", "hg", "isotropic", "tabphase", *([] if depth else ["blendphase"]))
        if kind == "hg":
            return {"type": "hg", "g": float(pick(0.85, -0.3, 0.0))}
        if kind == "tabphase":
            return {"type": "tabphase", "values": pick(TAB, [float(x) for x in rng.uniform(0.1, 2.0, int(pick(2, 5, 17)))])}
        if kind == "blendphase":
            w = pick(0.3, "grid")
            if w == "grid":
                w = {"type": "gridvolume", "data": rng.uniform(0, 1, (8, 1, 1)).astype(np.float32),   # a .vol file needs 8 values
                     "to_world": T.translate([-2, -2, 0], dt) @ T.scale([4, 4, 1], dt)}
            return {"type": "blendphase", "a": phase(1), "b": phase(1), "weight": w}
        return {"type": kind}
    # ---- medium
    tw = T.translate([-2, -2, 0], dt) @ T.scale([4, 4, 1], dt)
    kind = pick("homogeneous", "heterogeneous", "heterogeneous", "cloud")
    if kind == "homogeneous":
        # (no gridvolume here: the reference evaluates a homogeneous medium's sigma_t at mi.p BEFORE it is set,
        # medium.cpp:55 -- with a grid the result depends on what the stack held)
        sigma_t = pick(0.5, 2.0, {"type": "constvolume", "color": 1.5})
        d["atm"] = {"type": "homogeneous", "sigma_t": sigma_t if isinstance(sigma_t, dict) else float(sigma_t),
                    "albedo": pick(1.0, 0.8, {"type": "constvolume", "color": 0.7}), "phase": phase()}
    elif kind == "heterogeneous":
        grid = np.exp(-np.linspace(0, 4, 24)).astype(np.float32).reshape(24, 1, 1)
        if pick(False, False, False, True):   # a 3-channel grid read in a mono variant (grid3d.cpp:163-185)
            grid = np.repeat(grid[..., None], 3, axis=3) * np.array([1.0, 0.8, 1.2], np.float32)
        d["atm"] = {"type": "heterogeneous",
                    "sigma_t": dict({"type": "gridvolume", "data": grid, "to_world": tw, "filter_type": pick("trilinear", "nearest"),
                                     "wrap_mode": pick("clamp", "clamp", "repeat", "mirror")},
                                    **pick({}, {}, {"max_value": 1.5})),   # a majorant above the data: other free flights
                    "scale": float(pick(2.0, 6.0)), "albedo": pick(1.0, 0.9, {"type": "constvolume", "color": 0.85}), "phase": phase(),
                    "has_spectral_extinction": bool(pick(True, True, False))}
    else:
        n = 6
        d["atm"] = {"type": "heterogeneous",
                    "sigma_t": dict({"type": "gridvolume", "data": rng.uniform(0, 3, (n, n, n)).astype(np.float32), "to_world": tw},
                                    **pick({}, {}, {"use_grid_bbox": True, "bbox": (0.0, 0.0, 0.0, 1.0, 1.0, 0.5)})),
                    "albedo": {"type": "gridvolume", "to_world": tw,
                               "data": rng.uniform(0.5, 1, (n, n, n) + ((3,) if pick(False, True) else ())).astype(np.float32)},
                    "phase": phase()}
    if pick(False, False, False, True):
        d["atm"]["sample_emitters"] = False          # medium.cpp:29: no emitter sampling at scattering events in this medium
    if kind == "homogeneous" and pick(False, True):
        d["atm"]["scale"] = 1.5                      # homogeneous.cpp:27
    d["box"] = {"type": "cube", "to_world": T.translate([0, 0, 0.5], dt) @ T.scale([2, 2, 0.5], dt),
                "bsdf": {"type": "null"}, "interior": {"type": "ref", "id": "atm"}}
    if pick(False, False, False, True):
        # a second medium in a small cube inside the atmosphere: medium transitions on both faces
        # (interior / exterior media of a shape, shape.cpp:34-45; si.target_medium, interaction.h:360-370)
        d["dense"] = {"type": "homogeneous", "sigma_t": 3.0, "albedo": 0.6, "phase": {"type": "hg", "g": 0.5}}
        d["blob"] = {"type": "cube", "to_world": T.translate([-0.3, 0.4, 0.5], dt) @ T.scale(0.25, dt), "bsdf": {"type": "null"},
                     "interior": {"type": "ref", "id": "dense"}, "exterior": {"type": "ref", "id": "atm"}}
    # ---- surfaces
    bsdf = pick({"type": "diffuse", "reflectance": 0.3}, {"type": "rpv", "rho_0": 0.1, "k": 0.7, "g": -0.1},
                {"type": "rpv", "rho_0": 0.2, "k": 0.9, "g": 0.1, "rho_c": 0.05},
                {"type": "bilambertian", "reflectance": 0.4, "transmittance": 0.2})
    d["ground"] = {"type": "rectangle", "to_world": T.translate([0, 0, -0.01], dt) @ T.scale([2.5, 2.5, 1], dt), "bsdf": bsdf}
    if pick(False, False, False, True):   # a ground that faces down: one-sided BSDFs reflect nothing (rectangle.cpp:80-83)
        d["ground"]["flip_normals"] = True
    if pick(False, False, True):   # a mesh file under a transform, with its own normals or face normals
        name, kind = pick(("roof.obj", "obj"), ("roof_bin.ply", "ply"), ("roof.serialized", "serialized"))
        d["roof"] = {"type": kind, "filename": os.path.join(DATA, name), "face_normals": bool(pick(False, True)),
                     "to_world": T.translate([0.25, -0.5, 0.3], dt) @ T.scale([1.0, 0.75, 1.0], dt),
                     "bsdf": {"type": "diffuse", "reflectance": 0.75}}
        if kind == "serialized":
            d["roof"]["shape_index"] = 1
    if pick(False, True):
        d["ball"] = {"type": "sphere", "center": [0.3, -0.2, 0.4], "radius": 0.25, "flip_normals": bool(pick(False, True)),
                     "bsdf": {"type": "diffuse", "reflectance": 0.6}}
        if pick(False, True):   # placed by to_world instead (a similarity: sphere.cpp:100-117)
            d["ball"] = {"type": "sphere", "to_world": T.translate([-0.4, 0.3, 0.5], dt) @ T.scale(0.2, dt),
                         "bsdf": {"type": "diffuse", "reflectance": 0.6}}
    if pick(False, True):
        v, f = scenes.canopy(int(pick(6, 40)))
        d["leaves"] = {"type": "mesh", "vertex_positions": v, "faces": f,
                       "bsdf": pick({"type": "diffuse", "reflectance": 0.5}, {"type": "bilambertian", "reflectance": 0.5, "transmittance": 0.3})}
    # ---- emitters
    d["sun"] = {"type": "directional", "direction": [0.2, pick(0.0, 0.4), -1], "irradiance": 1.0}
    if pick(False, False, True):   # the same kind of sun through to_world (directional.cpp:50-60)
        d["sun"] = {"type": "directional", "irradiance": 1.5,
                    "to_world": T.look_at([0, 0, 0], [0.3, -0.1, -1.0], [0, 1, 0], np.float32)}
    if pick(False, False, True):
        d["sun2"] = {"type": "directional", "direction": [-0.5, 0.1, -1], "irradiance": 0.5}
    if pick(False, False, True):
        d["sky"] = {"type": "constant", "radiance": 0.2}
        if pick(False, True):
            del d["sun"]
            d.pop("sun2", None)
    # ---- sensor and film
    spp = int(pick(1, 2, 4))
    if passes and spp > 1:
        d["integrator"]["samples_per_pass"] = spp // 2
    w, h = pick((8, 1), (6, 5), (1, 1), (7, 3))
    film = {"type": "hdrfilm", "width": w, "height": h,
            "rfilter": pick({"type": "box"}, {"type": "box"}, {"type": "gaussian"}, {"type": "tent"}, {"type": "mitchell"},
                            {"type": "catmullrom"}, {"type": "lanczos", "lobes": 2}, {"type": "box", "radius": 1.1})}
    if w >= 6 and pick(False, True):
        film.update(crop_offset_x=1, crop_offset_y=0, crop_width=w - 2, crop_height=h)
    own = {"type": "rectangle", "to_world": T.translate([0.1, 0.2, pick(0.0, 0.6)], dt) @ T.scale([1.2, 0.8, 1], dt)}
    origin = {"type": "rectangle", "to_world": T.translate([0, 0, 1.3], dt) @ T.scale([pick(9.0, 1.5), 9.0, 1], dt)}
    sampler = {"type": "independent", "sample_count": spp, "seed": int(pick(0, 5))}
    kind = pick("distant", "distant", "mdistant", "distantflux", "radiancemeter", "mradiancemeter", "irradiancemeter")
    target = pick(None, [0.1, -0.2, 0.0], {"type": "ref", "id": "ground"}, own)
    if kind == "distant":
        sn = {"type": "distant", "film": film}
        how = pick("none", "direction", "direction+orientation", "to_world")
        if how.startswith("direction"):
            sn.update(direction=[0.1, 0.2, 1.0])
        if how == "direction+orientation":
            sn.update(orientation=[1.0, 0.5, 0.0])
        if how == "to_world":
            sn.update(to_world=T.look_at([0, 0, 0], [0.2, -0.1, 1.0], [0, 1, 0], np.float32))
        if pick(False, False, True):
            sn["flip_directions"] = True
        if target is not None:
            sn["ray_target"] = target
        if pick(False, False, True):
            sn["ray_origin"] = origin
    elif kind == "mdistant":
        n = int(pick(3, 6))
        th = (np.arange(n) + 0.5) / n * np.pi
        sn = {"type": "mdistant", "directions": np.stack([-np.cos(th), 0 * th, -np.sin(th)], 1).reshape(-1).tolist(),
              "film": dict(film, width=n, height=1)}
        sn["film"].pop("crop_offset_x", None)
        for k in ("crop_offset_y", "crop_width", "crop_height"):
            sn["film"].pop(k, None)
        if target is not None:
            sn["target"] = target
    elif kind == "distantflux":
        sn = {"type": "distantflux", "to_world": T.look_at([0, 0, 0], [0.1, -0.2, 1.0], [0, 1, 0], np.float32), "film": film}
        if target is not None:
            sn["target"] = target
        if pick(False, True):
            sn["origin"] = origin
    elif kind == "radiancemeter":
        sn = {"type": "radiancemeter", "origin": [0.2, 0.1, pick(2.0, 0.5)], "direction": [0.1, 0.0, -1.0],
              "film": {"type": "hdrfilm", "width": 1, "height": 1, "rfilter": {"type": "box"}}}
        if sn["origin"][2] < 1:
            sn["medium"] = {"type": "ref", "id": "atm"}
    elif kind == "irradiancemeter":   # on a small plate floating in (or above) the medium, facing up or down
        sn = {"type": "irradiancemeter", "film": {"type": "hdrfilm", "width": 1, "height": 1, "rfilter": {"type": "box"}}}
        z = pick(0.4, 1.2)
        if z < 1:
            sn["medium"] = {"type": "ref", "id": "atm"}
        d["plate"] = {"type": "rectangle", "flip_normals": bool(pick(False, True)),
                      "to_world": T.translate([0.4, -0.3, z], dt) @ T.scale([0.3, 0.2, 1], dt),
                      "bsdf": {"type": "diffuse", "reflectance": 0.25}}
    else:
        n = 4
        sn = {"type": "mradiancemeter", "origins": [0.1 * i for i in range(3 * n)][:3 * n],
              "directions": [0.1, 0.0, -1.0] * n, "film": {"type": "hdrfilm", "width": n, "height": 1, "rfilter": {"type": "box"}}}
        sn["origins"] = np.stack([np.linspace(-1, 1, n), np.zeros(n), np.full(n, 2.0)], 1).reshape(-1).tolist()
    sn["sampler"] = sampler
    if kind == "irradiancemeter":
        d["plate"]["sensor"] = sn
    else:
        d["sensor"] = sn
    return d


@pytest.mark.parametrize("variant", ["scalar_mono", "scalar_mono_double"])
@pytest.mark.parametrize("seed", range(160))
def test_random_scene(pkg, oracle, seed, variant):
    abi = pkg.abi
    d = random_scene(seed, variant)
    sc = pkg.load_dict(d, variant)
    mode = abi.SEED_SCALAR_BLOCKS if seed % 3 == 0 else abi.SEED_WAVEFRONT
    ref = oracle.render(pkg, sc, mode, want_paths=True)
    got = pkg.render(sc, device=0, seed_mode=mode, want_paths=True)
    assert np.array_equal(got.state, ref.state), "PCG32 end states differ"
    assert np.array_equal(got.counters, ref.counters), "event counters differ"
    assert np.array_equal(got.radiance.view(np.uint8), ref.radiance.view(np.uint8)), "per-path radiance differs"
    assert got.totals.n_invalid_samples == ref.totals.n_invalid_samples
    tol = 3e-5 if variant == "scalar_mono" else 1e-12
    for c in range(ref.film.shape[-1]):
        scale = max(np.abs(ref.film[..., c]).max(), 1e-30)
        assert np.abs(got.film[..., c].astype(np.float64) - ref.film[..., c]).max() <= tol * scale, "film channel %d" % c
    if seed % 4 == 0:   # and the twin kernels on a quarter of the scenes
        for kernel in ("mega", "lanes"):
            other = pkg.render(sc, device=0, seed_mode=mode, want_paths=True, kernel=kernel)
            assert np.array_equal(other.state, ref.state) and np.array_equal(other.counters, ref.counters), kernel
            assert np.array_equal(other.radiance.view(np.uint8), ref.radiance.view(np.uint8)), kernel
