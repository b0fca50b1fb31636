"""Multi-GPU through the C-ABI (b2vp_render with a device list; one process, one host thread per GPU, one NCCL film
reduce per render).  GPU box with several GPUs (gpurun --gpus N).  One JSON line per measurement.

  strong   C2 at its bench size (256 x 1, 2^14 spp = 4 194 304 paths), the SAME work split over 1, 2, 4, ... GPUs
  weak     2^14 spp per GPU (what bench.py does with one process per GPU under torchrun), here in one process
  c4       BASELINE configs[3]: 64 wavelengths x the C2 atmosphere, 10^8 paths per wavelength, every wavelength
           sharded over all GPUs (interleaved sample blocks), per-wavelength scalars patched into every resident scene
           copy with b2vp_scene_update, 64 films of 256 x 1 x 5

usage: python tools/bench_multi.py [strong] [weak] [c4] [math=fast|exact] [paths=1e8] [wavelengths=64]"""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as ge
pkg = ge.build()
import scenes
abi = pkg.abi

opts = dict(a.split("=", 1) for a in sys.argv[1:] if "=" in a)
which = [a for a in sys.argv[1:] if "=" not in a] or ["strong", "weak", "c4"]
MATH = opts.get("math", "fast")
N = abi.lib().b2vp_device_count()
counts = [n for n in (1, 2, 4, 8) if n <= N]
WIDTH = 256


def timed(ds, reps, **kw):
    ds.render(**kw)   # warm-up: scene replicas, communicators
    best, tot = None, None
    for _ in range(reps):
        t0 = time.perf_counter()
        r = ds.render(**kw)
        dt = time.perf_counter() - t0
        if best is None or dt < best:
            best, tot = dt, r.totals
    return best, tot


if "strong" in which or "weak" in which:
    for mode in [m for m in ("strong", "weak") if m in which]:
        base = None
        for n in counts:
            spp = 16384 if mode == "strong" else 16384 * n
            sc = pkg.load_dict(scenes.c2(spp=spp, width=WIDTH))
            with pkg.DeviceScene(sc, device=0) as ds:
                wall, tot = timed(ds, 5, math=MATH, devices=list(range(n)))
            rate = tot.n_paths / wall / 1e6
            base = base or rate
            print(json.dumps({"what": mode + " scaling, C2, one process, b2vp_render(devices)", "math": MATH, "gpus": n,
                              "paths": tot.n_paths, "wall_ms": wall * 1e3, "slowest_kernel_ms": tot.gpu_ms,
                              "mpaths_s_wall": rate, "efficiency": rate / (base * n) if mode == "weak" else rate / (base * n),
                              "speedup": rate / base}), flush=True)

if "c4" in which:
    n_wl = int(opts.get("wavelengths", 64))
    paths = int(float(opts.get("paths", 1e8)))
    spp = paths // WIDTH
    d = scenes.c2(spp=spp, width=WIDTH)
    sc = pkg.load_dict(d)
    base_scale = d["atm"]["scale"]
    wl = np.linspace(400.0, 1000.0, n_wl)
    rho0 = lambda lam: float(np.float32(0.05 + 0.25 * (lam - 400.0) / 600.0))
    films = np.zeros((n_wl, WIDTH, 5), np.float32)
    devices = list(range(N))
    with pkg.DeviceScene(sc, device=0) as ds:
        ds.render(n_streams=1 << 16, math=MATH, devices=devices)       # replicas + communicators
        t0 = time.perf_counter()
        k_ms, n_paths, segs = 0.0, 0, 0
        for i, lam in enumerate(wl):
            ds.update(medium=0, medium_scale=float(np.float32(base_scale * (550.0 / lam) ** 4)),
                      bsdf=1, bsdf_p=(rho0(lam), 0.7, -0.1, rho0(lam)))
            r = ds.render(math=MATH, devices=devices)
            films[i] = r.film[0]
            k_ms += r.totals.gpu_ms; n_paths += r.totals.n_paths; segs += r.totals.n_medium_steps + r.totals.n_intersect
        wall = time.perf_counter() - t0
    mean = films[:, :, 1].sum(axis=1) / films[:, :, 4].sum(axis=1)
    print(json.dumps({"what": "C4 = BASELINE configs[3]: %d wavelengths x C2, %d paths per wavelength, every wavelength sharded "
                              "over %d GPUs, NCCL film reduce per wavelength, b2vp_scene_update per wavelength" % (n_wl, spp * WIDTH, N),
                      "math": MATH, "gpus": N, "paths": n_paths, "wall_s": wall, "sum_of_slowest_kernel_ms": k_ms,
                      "mpaths_s_wall": n_paths / wall / 1e6, "segments_per_s_wall": segs / wall,
                      "per_wavelength_ms": wall / n_wl * 1e3,
                      "film_weight_deficit_per_wavelength": float(spp * WIDTH - films[:, :, 4].astype(np.float64).sum(axis=1).max()),
                      "films_complete": bool(n_paths == n_wl * spp * WIDTH and (np.abs(films[:, :, 4].astype(np.float64).sum(axis=1) - spp * WIDTH) <= 1e-6 * spp * WIDTH).all()),
                      "mean_radiance_400nm": float(mean[0]), "mean_radiance_1000nm": float(mean[-1])}), flush=True)
