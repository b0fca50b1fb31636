"""Is the majorant-grid mode unbiased?  Two-sample z per pixel (both variances, from `moment` films) between
  control: global majorant, sampler seed 0  vs  global majorant, seed 1      (must look like N(0,1): validates the test)
  grid:    majorant grid,   sampler seed 0  vs  global majorant, seed 1
at high sample counts.  Prints the mean z (a bias shows as mean z * sqrt(pixels) far from 0), its rms (~1) and max |z|.
usage: python tools/mg_bias_check.py [majorant_grid value]   (GPU box)"""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as ge
pkg = ge.build()
import scenes
abi = pkg.abi
MG = int(sys.argv[1]) if len(sys.argv) > 1 else 1


def c3(**kw):
    d = scenes.c3(**kw)
    d["sensor"]["ray_target"] = [0.4, 0.55, 1.0]
    return d


def stats(film):
    f = film.astype(np.float64)
    n = f[..., 4]
    mean = f[..., 6] / n
    var = np.maximum(f[..., 9] / n - mean * mean, 0)
    return mean, var, n


def z(a, b):
    ma, va, na = stats(a); mb, vb, nb = stats(b)
    se = np.sqrt(va / na + vb / nb)
    ok = se > 0
    return ((ma - mb)[ok] / se[ok]), (ma * na).sum() / na.sum(), (mb * nb).sum() / nb.sum(), np.sqrt((va / na).sum() + (vb / nb).sum()) / ma.size


CASES = {"c2": lambda: scenes.c2(spp=16384, width=64), "c2_tau2": lambda: scenes.c2(spp=8192, width=64, tau=2.0),
         "c3_small": lambda: c3(spp=8192, n=32, size=12), "c3_96": lambda: c3(spp=4096, n=96, size=12)}
for name, make in CASES.items():
    def render(seed, **kw):
        d = make()
        d["integrator"] = {"type": "moment", "nested": d["integrator"]}
        d["sensor"]["sampler"]["seed"] = seed
        return pkg.render(pkg.load_dict(d), seed_mode=abi.SEED_WAVEFRONT, **kw)
    g0, g1, m0 = render(0), render(1), render(0, majorant_grid=MG)
    for label, a in (("control", g0), ("grid", m0)):
        zz, mean_a, mean_b, se_img = z(a.film, g1.film)
        print(json.dumps({"case": name, "pair": label, "pixels": int(zz.size), "mean_z": float(zz.mean()),
                          "mean_z_x_sqrt_n": float(zz.mean() * np.sqrt(zz.size)), "rms_z": float(np.sqrt((zz ** 2).mean())),
                          "max_abs_z": float(np.abs(zz).max()), "image_mean": [mean_a, mean_b],
                          "image_diff_over_se": float((mean_a - mean_b) / se_img),
                          "steps_per_path": a.totals.n_medium_steps / a.totals.n_paths}), flush=True)
