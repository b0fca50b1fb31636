"""Where the end-to-end step of bench.py spends its time: b2vp_scene_create (descriptor -> device), the first render on a
fresh device scene, the film read-back.  One GPU; wall-clock microseconds, best of 20."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as ge
pkg = ge.build()
import scenes
abi = pkg.abi
sc = pkg.load_dict(scenes.c2(spp=16384, width=256))
pkg.render(sc, math="fast")   # CUDA up, kernels loaded
t_create, t_render, t_destroy = [], [], []
for _ in range(20):
    t0 = time.perf_counter()
    ds = pkg.DeviceScene(sc, 0)
    t1 = time.perf_counter()
    r = ds.render(math="fast")
    t2 = time.perf_counter()
    ds.close()
    t3 = time.perf_counter()
    t_create.append(t1 - t0); t_render.append(t2 - t1); t_destroy.append(t3 - t2)
print("scene_create %.0f us, render (incl. lazy upload, kernel %.0f us, read-back) %.0f us, destroy %.0f us" % (
    min(t_create) * 1e6, r.totals.gpu_ms * 1e3, min(t_render) * 1e6, min(t_destroy) * 1e6))
