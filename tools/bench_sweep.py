"""Spectral sweep through the boundary: one resident scene, per wavelength new sigma_t / albedo grids
(b2vp_scene_update_grid) + one render, against building a fresh device scene per wavelength.  C3 at a size given on
the command line (grid n^3, default 256; film 128 x 128, 16 spp).  GPU box only; prints one JSON line.
usage: python tools/bench_sweep.py [n] [wavelengths]"""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as ge
pkg = ge.build()
import scenes
abi = pkg.abi

n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
n_wl = int(sys.argv[2]) if len(sys.argv) > 2 else 4
d = scenes.c3(spp=16, n=n, size=128)
s0 = np.asarray(d["atm"]["sigma_t"]["data"], np.float32)
a0 = np.asarray(d["atm"]["albedo"]["data"], np.float32)
grids = [((s0 * np.float32(1 + 0.1 * w)).copy(), (a0 * np.float32(1 - 0.01 * w)).copy()) for w in range(n_wl)]

t0 = time.perf_counter()
scene = pkg.load_dict(d)
t_load = time.perf_counter() - t0
med = scene.desc.media[0]
fresh, patched, kernel, films_fresh, films_patched = [], [], [], [], []
for s, a in grids:                       # a device scene per wavelength (host tables rebuilt, everything uploaded)
    d["atm"]["sigma_t"]["data"], d["atm"]["albedo"]["data"] = s, a
    t0 = time.perf_counter()
    with pkg.DeviceScene(pkg.load_dict(d), device=0) as ds:
        r = ds.render(seed_mode=abi.SEED_WAVEFRONT)
    fresh.append(time.perf_counter() - t0)
    films_fresh.append(r.film)
with pkg.DeviceScene(scene, device=0) as ds:   # one resident scene, grids patched in place
    ds.render(seed_mode=abi.SEED_WAVEFRONT)
    for s, a in grids:
        t0 = time.perf_counter()
        ds.update_grid(med.sigma_t, s)
        ds.update_grid(med.albedo, a)
        t1 = time.perf_counter()
        r = ds.render(seed_mode=abi.SEED_WAVEFRONT)
        patched.append(time.perf_counter() - t0)
        kernel.append(r.totals.gpu_ms)
        films_patched.append(r.film)
        upd = t1 - t0
same = all(np.array_equal(x[..., :4], y[..., :4]) or np.allclose(x, y, rtol=1e-3, atol=1e-6) for x, y in zip(films_fresh, films_patched))
print(json.dumps({"workload": "C3 sweep: %d^3 sigma_t + albedo grids (%.0f MB each), 128x128 film, 16 spp, %d wavelengths" % (n, s0.nbytes / 1e6, n_wl),
                  "host_load_s": t_load, "fresh_scene_per_wavelength_s": min(fresh[1:] or fresh),
                  "patched_per_wavelength_s": min(patched), "of_which_update_s": upd, "kernel_ms": min(kernel),
                  "grid_bytes_per_update": int(s0.nbytes + a0.nbytes), "films_agree": bool(same)}))
