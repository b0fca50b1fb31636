"""Full-size runs of the BASELINE configurations other than the one `bench.py` quotes (C1, C3, C5).
GPU box only.  Prints one JSON line per configuration: device-timed Mpaths/s, ray segments/s, algorithmic
grid bytes (C3) and the wall time of scene upload.  usage: python tools/bench_configs.py [c1] [c3] [c4] [c5]"""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as ge
pkg = ge.build()
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import scenes
abi = pkg.abi


RENDER_KW = {}   # e.g. math="fast", majorant_grid=1 (command line: key=value)


def run(name, desc, scene_dict, reps=2, extra=None, variant="scalar_mono"):
    t0 = time.perf_counter()
    scene = pkg.load_dict(scene_dict, variant)
    t1 = time.perf_counter()
    with pkg.DeviceScene(scene, device=0) as ds:
        t2 = time.perf_counter()
        r = ds.render(seed_mode=abi.SEED_WAVEFRONT, **RENDER_KW)       # warm-up (lazy upload, kernel attributes)
        t3 = time.perf_counter()
        ms = []
        for _ in range(reps):
            r = ds.render(seed_mode=abi.SEED_WAVEFRONT, **RENDER_KW)
            ms.append(r.totals.gpu_ms)
        kernel = ds.info().last_kernel.decode()
    t = r.totals
    k = min(ms) * 1e-3
    out = {"config": name, "workload": desc, "mode": RENDER_KW, "kernel": kernel, "paths": t.n_paths, "kernel_ms": min(ms),
           "mpaths_s": t.n_paths / k / 1e6, "segments_per_s": (t.n_medium_steps + t.n_intersect) / k,
           "medium_steps_per_path": t.n_medium_steps / max(t.n_paths, 1),
           "intersections_per_path": t.n_intersect / max(t.n_paths, 1),
           "real_per_path": t.n_real / max(t.n_paths, 1), "null_per_path": t.n_null / max(t.n_paths, 1),
           "host_load_dict_s": t1 - t0, "scene_create_s": t2 - t1, "first_render_s": t3 - t2,
           "mean_radiance": float(np.mean(r.bitmap()))}
    if extra:
        out.update(extra(t, k))
    print(json.dumps(out), flush=True)


for a in [a for a in sys.argv[1:] if "=" in a]:
    k, v = a.split("=", 1)
    if k in ("stay", "iter", "cta_phase", "slots_per_thread"):   # queue-kernel scheduler parameters (B2vpRenderArgs::tuning)
        RENDER_KW.setdefault("tuning", {})[k] = float(v) if k == "slots_per_thread" else int(v)
    else:
        RENDER_KW[k] = int(v) if v.isdigit() else v
which = [a for a in sys.argv[1:] if "=" not in a] or ["c1", "c3", "c5"]
if "c1" in which:
    run("C1", "homogeneous Rayleigh cube over diffuse rectangle, distant 1x1, 1e6 spp", scenes.c1(spp=1000000))
if "c3" in which:
    n = 256
    def hbm(t, k):
        lookups = t.n_medium_steps  # one sigma_t and one albedo lookup per tentative collision
        alg = lookups * 32 * 2
        return {"grid": "2 x %d^3 f32 = %.0f MiB" % (n, 2 * n ** 3 * 4 / 2 ** 20), "algorithmic_grid_bytes": alg,
                "algorithmic_GBps": alg / k / 1e9, "sector_floor_GBps": lookups * 128 * 2 / k / 1e9}
    run("C3", "cloud field %d^3 sigma_t/albedo, hg g=0.85, distant 512x512, 64 spp" % n,
        scenes.c3(spp=64, n=n, size=512), extra=hbm)
if "c5" in which:
    run("C5", "canopy 1e6 triangles (BVH), homogeneous Rayleigh slab, diffuse leaves, mradiancemeter 4096, 256 spp",
        scenes.c5(spp=256, n_leaves=500000, n_sensors=4096))
if "c5atm" in which or "c5aer" in which:
    # what a RAMI4ATM-like Eradiate scene looks like: the C5 canopy under a HETEROGENEOUS 1D atmosphere (the C2 sigma_t
    # column over the slab), Rayleigh only ("c5atm") or with an aerosol layer mixed in by a blendphase ("c5aer")
    import make_ref_golden as _g
    for tag in [t for t in ("c5atm", "c5aer") if t in which]:
        d = scenes.c5(spp=256, n_leaves=500000, n_sensors=4096)
        T = scenes.T
        grid = scenes.sigma_t_1d(1000)
        d["atm"] = {"type": "heterogeneous",
                    "sigma_t": {"type": "gridvolume", "data": grid,
                                "to_world": T.translate([-1.5, -1.5, 0.2], np.float32) @ T.scale([3.0, 3.0, 0.8], np.float32)},
                    "scale": 0.3 / float(grid.mean()) / 0.8, "albedo": 1.0, "phase": {"type": "rayleigh"}}
        if tag == "c5aer":
            w = np.clip(np.linspace(1.0, -0.5, 64), 0, 1).astype(np.float32).reshape(64, 1, 1)
            d["atm"]["phase"] = {"type": "blendphase", "a_molecules": {"type": "rayleigh"},
                                 "b_aerosols": {"type": "tabphase", "values": _g.TAB_AEROSOL},
                                 "weight": {"type": "gridvolume", "data": w, "to_world": d["atm"]["sigma_t"]["to_world"]}}
        run(tag.upper(), "C5 canopy (1e6 triangles) under a heterogeneous 1D atmosphere%s, mradiancemeter 4096, 256 spp"
            % (" with an aerosol layer" if tag == "c5aer" else ""), d)
if "c4" in which:
    # mono spectral sweep: 64 wavelengths x the C2 atmosphere; per wavelength only scalars change (Rayleigh
    # extinction ~ lambda^-4, RPV rho_0(lambda)), applied with b2vp_scene_update -- no re-upload, no rebuild.
    # 2^20 paths per wavelength here (BASELINE: 1e8 per wavelength over 8 GPUs; the rate is what is measured).
    import time as _t
    spp, width, n_wl = 4096, 256, 64
    d = scenes.c2(spp=spp, width=width)
    scene = pkg.load_dict(d)
    base_scale = d["atm"]["scale"]
    wl = np.linspace(400.0, 1000.0, n_wl)
    films = np.zeros((n_wl, width, 5), np.float32)
    rho0 = lambda lam: float(np.float32(0.05 + 0.25 * (lam - 400.0) / 600.0))
    with pkg.DeviceScene(scene, device=0) as ds:
        ds.render(seed_mode=abi.SEED_WAVEFRONT)
        t0 = _t.perf_counter()
        k_ms, paths, segs = 0.0, 0, 0
        for i, lam in enumerate(wl):
            ds.update(medium=0, medium_scale=float(np.float32(base_scale * (550.0 / lam) ** 4)),
                      bsdf=1, bsdf_p=(rho0(lam), 0.7, -0.1, rho0(lam)))
            r = ds.render(seed_mode=abi.SEED_WAVEFRONT)
            films[i] = r.film[0]
            k_ms += r.totals.gpu_ms; paths += r.totals.n_paths; segs += r.totals.n_medium_steps + r.totals.n_intersect
        wall = _t.perf_counter() - t0
    print(json.dumps({"config": "C4", "workload": "%d wavelengths x C2 (256x1, %d spp), scene_update per wavelength" % (n_wl, spp),
                      "paths": paths, "kernel_ms": k_ms, "wall_ms": wall * 1e3, "mpaths_s": paths / (k_ms * 1e-3) / 1e6,
                      "mpaths_s_wall": paths / wall / 1e6, "segments_per_s": segs / (k_ms * 1e-3),
                      "mean_radiance_400nm": float(films[0, :, 1].sum() / films[0, :, 4].sum()),
                      "mean_radiance_1000nm": float(films[-1, :, 1].sum() / films[-1, :, 4].sum())}), flush=True)
if "c2" in which:
    run("C2", "1D heterogeneous atmosphere, 256x1, 16384 spp", scenes.c2(spp=16384, width=256))
if "c2f64" in which:
    run("C2-f64", "C2 in scalar_mono_double, 256x1, 2048 spp", scenes.c2(spp=2048, width=256, variant="scalar_mono_double"), variant="scalar_mono_double")
