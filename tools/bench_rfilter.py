"""C2 at full size (4 194 304 paths) under a reconstruction filter wider than a pixel: the generic queue instance with the
footprint splat (FT_RFILTER) against the box-filter preset.  usage: python tools/bench_rfilter.py"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import b2vp_import  # noqa: E402

pkg = b2vp_import.load()
import scenes  # noqa: E402

for rf in ({"type": "box"}, {"type": "tent"}, {"type": "gaussian"}, {"type": "lanczos"}):
    for math in ("exact", "fast"):
        d = scenes.c2(spp=16384, width=256)
        d["sensor"]["film"] = {"type": "hdrfilm", "width": 256, "height": 1, "rfilter": rf}
        with pkg.DeviceScene(pkg.load_dict(d)) as ds:
            best = None
            for _ in range(4):
                r = ds.render(seed_mode=pkg.abi.SEED_WAVEFRONT, math=math)
                best = r.totals.gpu_ms if best is None else min(best, r.totals.gpu_ms)
            print(json.dumps({"rfilter": rf["type"], "math": math, "kernel": ds.info().last_kernel.decode(),
                              "kernel_ms": best, "mpaths_s": r.totals.n_paths / best / 1e3,
                              "film_weight": float(r.film[..., 4].sum())}))
