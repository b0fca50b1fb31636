"""Is B2VP_MATH_FAST unbiased?  Renders a scene in both fp32 modes with many samples (independent of the per-path
comparison of tests/test_gpu_fastmath.py) and prints the image means with their standard errors.
usage: python tools/fast_bias_check.py   (GPU box)"""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as ge
pkg = ge.build()
import scenes
abi = pkg.abi

def c3_off_edge(**kw):
    """scenes.c3 aims every camera ray at the centre of the cube's top face, i.e. exactly at the edge its two triangles
    share: whether such a ray enters the medium at all depends on the last bit of the triangle test.  Off the edge."""
    d = scenes.c3(**kw)
    d["sensor"]["ray_target"] = [0.4, 0.55, 1.0]
    return d


CASES = {"c3_small": scenes.c3(spp=256, n=32, size=48), "c3_96": scenes.c3(spp=64, n=96, size=64),
         "c3_small_off_edge": c3_off_edge(spp=256, n=32, size=48), "c3_96_off_edge": c3_off_edge(spp=64, n=96, size=64),
         "c2": scenes.c2(spp=1024, width=256), "c5_small": scenes.c5(spp=256, n_leaves=400, n_sensors=256)}
for name, d in CASES.items():
    d["integrator"] = {"type": "moment", "nested": d["integrator"]}      # second moments: per-pixel variance
    sc = pkg.load_dict(d)
    with pkg.DeviceScene(sc) as ds:
        out = {}
        for math in ("exact", "fast"):
            r = ds.render(seed_mode=abi.SEED_WAVEFRONT, math=math)
            f = r.film.astype(np.float64)
            steps = r.totals.n_medium_steps / r.totals.n_paths
            n = f[..., 4].sum()
            mean = f[..., 6].sum() / n
            var = f[..., 9].sum() / n - mean * mean
            out[math] = (mean, np.sqrt(var / n), n, steps)
    (me, se, n, st_e), (mf, sf, _, st_f) = out["exact"], out["fast"]
    # the two renders share most paths, so their difference is far smaller than the independent-sample error; a
    # difference of several standard errors of ONE render would be a bias
    print(json.dumps({"case": name, "paths": int(n), "mean_exact": me, "mean_fast": mf, "stderr": se,
                      "diff_over_stderr": (mf - me) / se,
                      "medium_steps_per_path": [st_e, st_f]}), flush=True)
