"""C2 at full size with an aerosol layer: molecules (rayleigh) and aerosols (tabphase) mixed by a blendphase whose weight
is a 1 x 1 x N grid (tests/golden/make_ref_golden.py::_blendphase), and the plain tabulated phase function.
usage: python tools/bench_aerosol.py"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import b2vp_import  # noqa: E402

pkg = b2vp_import.load()
import make_ref_golden as g  # noqa: E402
import scenes  # noqa: E402

cases = {"rayleigh": lambda: scenes.c2(spp=16384, width=256),
         "tabphase": lambda: g._tabphase(g.TAB_AEROSOL, spp=16384, width=256),
         "blendphase, grid weight": lambda: g._blendphase(True, spp=16384, width=256),
         "blendphase, constant weight": lambda: g._blendphase(False, spp=16384, width=256),
         "rayleigh, distantflux 16 x 16": lambda: g._distantflux(True, spp=16384, width=256)}
for name, make in cases.items():
    for integ in ("volpath", "volpathmis"):
        for math in ("exact", "fast"):
            d = make()
            d["integrator"] = dict(d["integrator"], type=integ)
            with pkg.DeviceScene(pkg.load_dict(d)) as ds:
                best = None
                for _ in range(3):
                    r = ds.render(seed_mode=pkg.abi.SEED_WAVEFRONT, math=math)
                    best = r.totals.gpu_ms if best is None else min(best, r.totals.gpu_ms)
                print(json.dumps({"phase": name, "integrator": integ, "math": math, "kernel": ds.info().last_kernel.decode(),
                                  "kernel_ms": best, "mpaths_s": r.totals.n_paths / best / 1e3}))
