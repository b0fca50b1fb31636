"""Full-size runs of the BASELINE configurations other than the one `bench.py` quotes (C1, C3, C5).
GPU box only.  Prints one JSON line per configuration: device-timed Mpaths/s, ray segments/s, algorithmic
grid bytes (C3) and the wall time of scene upload.  usage: python tools/bench_configs.py [c1] [c3] [c5]"""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as ge
pkg = ge.build()
import scenes
abi = pkg.abi


def run(name, desc, scene_dict, reps=2, extra=None):
    t0 = time.perf_counter()
    scene = pkg.load_dict(scene_dict)
    t1 = time.perf_counter()
    with pkg.DeviceScene(scene, device=0) as ds:
        t2 = time.perf_counter()
        r = ds.render(seed_mode=abi.SEED_WAVEFRONT)       # warm-up (lazy upload, kernel attributes)
        t3 = time.perf_counter()
        ms = []
        for _ in range(reps):
            r = ds.render(seed_mode=abi.SEED_WAVEFRONT)
            ms.append(r.totals.gpu_ms)
    t = r.totals
    k = min(ms) * 1e-3
    out = {"config": name, "workload": desc, "paths": t.n_paths, "kernel_ms": min(ms),
           "mpaths_s": t.n_paths / k / 1e6, "segments_per_s": (t.n_medium_steps + t.n_intersect) / k,
           "medium_steps_per_path": t.n_medium_steps / max(t.n_paths, 1),
           "intersections_per_path": t.n_intersect / max(t.n_paths, 1),
           "host_load_dict_s": t1 - t0, "scene_create_s": t2 - t1, "first_render_s": t3 - t2,
           "mean_radiance": float(np.mean(r.bitmap()))}
    if extra:
        out.update(extra(t, k))
    print(json.dumps(out), flush=True)


which = sys.argv[1:] or ["c1", "c3", "c5"]
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
