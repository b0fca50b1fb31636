"""volpathmis at the full BASELINE sizes on the queue kernel (FT_MIS instances) next to volpath on the same scene and,
on a bounded sample, the one-thread-per-stream kernel that ran volpathmis before.  GPU box only; one JSON line per
configuration.  usage: python tools/bench_mis.py [c1] [c2] [c3] [c5] [math=fast]"""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as ge
pkg = ge.build()
import scenes
abi = pkg.abi

KW = {}
for a in [a for a in sys.argv[1:] if "=" in a]:
    k, v = a.split("=", 1)
    KW[k] = int(v) if v.isdigit() else v
which = [a for a in sys.argv[1:] if "=" not in a] or ["c1", "c2", "c3", "c5"]

CONFIGS = {
    "c1": ("C1 1e6 spp", lambda: scenes.c1(spp=1000000)),
    "c2": ("C2 64 layers, mdistant 1024 x 4096 spp", lambda: scenes.c2(spp=4096, width=1024)),
    "c3": ("C3 256^3 cloud field, 512 x 512 x 64 spp", lambda: scenes.c3(spp=64, n=256, size=512)),
    "c5": ("C5 1e6 triangles, 4096 sensors x 256 spp", lambda: scenes.c5(spp=256, n_leaves=500000, n_sensors=4096)),
}


def timed(ds, reps, **kw):
    r = ds.render(seed_mode=abi.SEED_WAVEFRONT, **kw)
    ms = [ds.render(seed_mode=abi.SEED_WAVEFRONT, **kw).totals.gpu_ms for _ in range(reps)]
    return r, min(ms), ds.info().last_kernel.decode()


for name in which:
    desc, make = CONFIGS[name]
    out = {"config": name, "workload": desc, "mode": KW}
    for integrator in ("volpath", "volpathmis"):
        d = make()
        d["integrator"] = dict(d["integrator"], type=integrator)
        sc = pkg.load_dict(d)
        with pkg.DeviceScene(sc) as ds:
            r, ms, kernel = timed(ds, 2, **KW)
            out[integrator] = {"kernel": kernel, "paths": r.totals.n_paths, "kernel_ms": ms,
                               "mpaths_s": r.totals.n_paths / ms / 1e3,
                               "steps_per_path": r.totals.n_medium_steps / r.totals.n_paths,
                               "mean_radiance": float(np.mean(r.bitmap()))}
            if integrator == "volpathmis":
                # the kernel that ran volpathmis before: 1/64 of the streams
                total, _ = ds.stream_count(abi.SEED_WAVEFRONT)
                n = max(total // 64, 1)
                r2 = ds.render(seed_mode=abi.SEED_WAVEFRONT, kernel="mega", n_streams=n)
                r2 = ds.render(seed_mode=abi.SEED_WAVEFRONT, kernel="mega", n_streams=n)
                out["volpathmis_stream_kernel"] = {"kernel": ds.info().last_kernel.decode(), "paths": r2.totals.n_paths,
                                                   "kernel_ms": r2.totals.gpu_ms,
                                                   "mpaths_s": r2.totals.n_paths / r2.totals.gpu_ms / 1e3}
    print(json.dumps(out), flush=True)
