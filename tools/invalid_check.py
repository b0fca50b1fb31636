"""How many samples does ImageBlock::put reject (non-finite or negative radiance, imageblock.cpp:85-110) on a C4-sized
render?  One GPU, 1e8 paths per wavelength at the two ends and the middle of the sweep, both math modes."""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as ge
pkg = ge.build()
import scenes
WIDTH, spp = 256, 100000000 // 256
d = scenes.c2(spp=spp, width=WIDTH)
sc = pkg.load_dict(d)
base = d["atm"]["scale"]
with pkg.DeviceScene(sc) as ds:
    for lam in (400.0, 550.0, 1000.0):
        rho0 = float(np.float32(0.05 + 0.25 * (lam - 400.0) / 600.0))
        ds.update(medium=0, medium_scale=float(np.float32(base * (550.0 / lam) ** 4)), bsdf=1, bsdf_p=(rho0, 0.7, -0.1, rho0))
        for math in ("exact", "fast"):
            r = ds.render(math=math)
            w = r.film[..., 4].astype(np.float64)
            print(json.dumps({"wavelength": lam, "math": math, "paths": int(r.totals.n_paths),
                              "invalid_samples": int(r.totals.n_invalid_samples),
                              "film_weight_deficit": float(spp * WIDTH - w.sum()),
                              "pixels_short": int((w != spp).sum()), "mean": float(r.film[..., 1].sum() / w.sum())}), flush=True)
