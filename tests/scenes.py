"""Synthetic scenes of the BASELINE.json configurations as load_dict dictionaries."""
import numpy as np

import b2vp_import

pkg = b2vp_import.load()
T = pkg.Transform


def film(w, h=1):
    return {"type": "hdrfilm", "width": w, "height": h, "rfilter": {"type": "box"}}


def c1(spp=1024, variant="scalar_mono", sun=(0, 0, -1), **integ):
    """C1: homogeneous Rayleigh cube (tau 0.5, albedo 1) over a diffuse rectangle, distant 1x1."""
    dt = pkg.scene.VARIANTS[variant][0]
    return {
        "type": "scene",
        "integrator": dict({"type": "volpath", "max_depth": -1}, **integ),
        "atm": {"type": "homogeneous", "sigma_t": 0.5, "albedo": 1.0, "phase": {"type": "rayleigh"}},
        # the cube reaches below the ground so that no two surfaces are coplanar (the appendix-A scene has the
        # cube's bottom face ON the rectangle: its result depends on how the accelerator breaks exact ties)
        "cube": {"type": "cube", "to_world": T.translate([0, 0, 0.45], dt) @ T.scale([1, 1, 0.55], dt),
                 "bsdf": {"type": "null"}, "interior": {"type": "ref", "id": "atm"}},
        "ground": {"type": "rectangle", "bsdf": {"type": "diffuse", "reflectance": 0.5}},
        "sun": {"type": "directional", "direction": list(sun), "irradiance": 1.0},
        "sensor": {"type": "distant", "direction": [0, 0, 1], "ray_target": [0, 0, 0],
                   "sampler": {"type": "independent", "sample_count": spp}, "film": film(1)},
    }


def sigma_t_1d(n=1000):
    z = (np.arange(n) + 0.5) / n
    return np.exp(-z * 100.0 / 8.0).astype(np.float32).reshape(n, 1, 1)


def c2(spp=64, variant="scalar_mono", width=256, tau=0.5, nz=1000, sensor="distant", filter_type="trilinear",
       spectral=True, **integ):
    """C2: 1D plane-parallel heterogeneous atmosphere, gridvolume 1x1xN, rayleigh, rpv, distant Wx1."""
    dt = pkg.scene.VARIANTS[variant][0]
    grid = sigma_t_1d(nz)
    scale = tau / float(grid.mean())  # column optical thickness tau over unit height
    hw = 10.0
    d = {
        "type": "scene",
        "integrator": dict({"type": "volpath", "max_depth": -1}, **integ),
        "atm": {"type": "heterogeneous",
                "sigma_t": {"type": "gridvolume", "data": grid, "filter_type": filter_type,
                            "to_world": T.translate([-hw, -hw, 0], dt) @ T.scale([2 * hw, 2 * hw, 1], dt)},
                "scale": scale, "albedo": 1.0, "phase": {"type": "rayleigh"},
                "has_spectral_extinction": spectral},
        "cube": {"type": "cube", "to_world": T.translate([0, 0, 0.45], dt) @ T.scale([hw, hw, 0.55], dt),
                 "bsdf": {"type": "null"}, "interior": {"type": "ref", "id": "atm"}},
        "ground": {"type": "rectangle", "to_world": T.scale([hw, hw, 1], dt),
                   "bsdf": {"type": "rpv", "rho_0": 0.1, "k": 0.7, "g": -0.1}},
        "sun": {"type": "directional", "direction": [0, 0.5, -1], "irradiance": 1.0},
    }
    if sensor == "distant":
        d["sensor"] = {"type": "distant", "direction": [0, 0, 1], "orientation": [1, 0, 0], "ray_target": [0, 0, 0],
                       "sampler": {"type": "independent", "sample_count": spp}, "film": film(width)}
    else:
        th = (np.arange(width) + 0.5) / width * np.pi
        dirs = np.stack([-np.cos(th), np.zeros_like(th), -np.sin(th)], axis=1)
        d["sensor"] = {"type": "mdistant", "directions": dirs.reshape(-1).tolist(), "target": [0, 0, 0],
                       "sampler": {"type": "independent", "sample_count": spp}, "film": film(width)}
    return d


def cloud_field(n=32, seed=0):
    """C3 grids: seeded Gaussian blobs thresholded to ~15 % cloud fraction (sigma_t, albedo)."""
    rng = np.random.default_rng(seed)
    ax = (np.arange(n) + 0.5) / n
    z, y, x = np.meshgrid(ax, ax, ax, indexing="ij")
    f = np.zeros((n, n, n))
    for _ in range(64):
        c = rng.random(3)
        r = 0.04 + 0.08 * rng.random()
        f += np.exp(-((x - c[0]) ** 2 + (y - c[1]) ** 2 + (z - c[2]) ** 2) / (2 * r * r))
    thr = np.quantile(f, 0.85)
    cloud = f > thr
    sig_cloud = 20.0 / 0.2   # tau ~ 20 across a 0.2-wide cloud
    sigma_t = np.where(cloud, sig_cloud, sig_cloud * 1e-3).astype(np.float32)
    albedo = np.where(cloud, 0.999, 0.9).astype(np.float32)
    return sigma_t, albedo


def c3(spp=4, variant="scalar_mono", n=32, size=16, **integ):
    """C3: 3D heterogeneous cloud field, gridvolume n^3 sigma_t/albedo, hg g=0.85, distant size x size."""
    dt = pkg.scene.VARIANTS[variant][0]
    sigma_t, albedo = cloud_field(n)
    return {
        "type": "scene",
        "integrator": dict({"type": "volpath", "max_depth": -1}, **integ),
        "atm": {"type": "heterogeneous", "sigma_t": {"type": "gridvolume", "data": sigma_t},
                "albedo": {"type": "gridvolume", "data": albedo}, "phase": {"type": "hg", "g": 0.85}},
        "cube": {"type": "cube", "to_world": T.translate([0.5, 0.5, 0.5], dt) @ T.scale(0.5, dt),
                 "bsdf": {"type": "null"}, "interior": {"type": "ref", "id": "atm"}},
        "ground": {"type": "rectangle", "to_world": T.translate([0.5, 0.5, -0.01], dt) @ T.scale([2, 2, 1], dt),
                   "bsdf": {"type": "diffuse", "reflectance": 0.1}},
        "top": {"type": "rectangle", "id": "top", "to_world": T.translate([0.5, 0.5, 1.5], dt) @ T.scale([0.5, 0.5, 1], dt),
                "bsdf": {"type": "null"}},
        "sun": {"type": "directional", "direction": [0.3, 0.2, -1], "irradiance": 1.0},
        "sensor": {"type": "distant", "ray_target": [0.5, 0.5, 1.0],
                   "sampler": {"type": "independent", "sample_count": spp}, "film": film(size, size)},
    }


def canopy(n_leaves=500, seed=0):
    rng = np.random.default_rng(seed)
    lai, area = 3.0, 4.0
    side = np.sqrt(lai * area / n_leaves)
    c = np.stack([rng.uniform(-1, 1, n_leaves), rng.uniform(-1, 1, n_leaves), rng.uniform(0.0, 0.2, n_leaves)], 1)
    nrm = rng.normal(size=(n_leaves, 3))
    nrm /= np.linalg.norm(nrm, axis=1)[:, None]
    a = np.cross(nrm, [0, 0, 1.0]) + 1e-9
    a /= np.linalg.norm(a, axis=1)[:, None]
    b = np.cross(nrm, a)
    h = side / 2
    v = np.stack([c - h * a - h * b, c + h * a - h * b, c + h * a + h * b, c - h * a + h * b], axis=1).reshape(-1, 3)
    base = np.arange(n_leaves)[:, None] * 4
    f = np.concatenate([base + [0, 1, 2], base + [0, 2, 3]], axis=1).reshape(-1, 3)
    return v.astype(np.float32), f.astype(np.uint32)


def c5(spp=4, variant="scalar_mono", n_leaves=500, n_sensors=64, leaf_bsdf=None, **integ):
    """C5: triangle-mesh canopy under a homogeneous Rayleigh slab, lambertian leaves, mradiancemeter array."""
    dt = pkg.scene.VARIANTS[variant][0]
    v, f = canopy(n_leaves)
    rng = np.random.default_rng(1)
    g = int(np.sqrt(n_sensors))
    xs = (np.arange(g) + 0.5) / g * 1.6 - 0.8
    ox, oy = np.meshgrid(xs, xs)
    origins = np.stack([ox.ravel(), oy.ravel(), np.full(g * g, 0.5)], axis=1)
    tilt = np.radians(rng.uniform(-30, 30, size=(g * g, 2)))
    dirs = np.stack([np.tan(tilt[:, 0]), np.tan(tilt[:, 1]), -np.ones(g * g)], axis=1)
    dirs /= np.linalg.norm(dirs, axis=1)[:, None]
    return {
        "type": "scene",
        "integrator": dict({"type": "volpath", "max_depth": -1}, **integ),
        "atm": {"type": "homogeneous", "sigma_t": 0.3 / 0.8, "albedo": 1.0, "phase": {"type": "rayleigh"}},
        "slab": {"type": "cube", "to_world": T.translate([0, 0, 0.6], dt) @ T.scale([1.5, 1.5, 0.4], dt),
                 "bsdf": {"type": "null"}, "interior": {"type": "ref", "id": "atm"}},
        "leaves": {"type": "mesh", "vertex_positions": v, "faces": f,
                   "bsdf": leaf_bsdf or {"type": "diffuse", "reflectance": 0.5}},
        "ground": {"type": "rectangle", "to_world": T.scale([1.5, 1.5, 1], dt),
                   "bsdf": {"type": "diffuse", "reflectance": 0.2}},
        "sun": {"type": "directional", "direction": [0.2, 0.1, -1], "irradiance": 1.0},
        "sensor": {"type": "mradiancemeter", "origins": origins.reshape(-1).tolist(),
                   "directions": dirs.reshape(-1).tolist(),
                   "sampler": {"type": "independent", "sample_count": spp}, "film": film(g * g)},
    }
