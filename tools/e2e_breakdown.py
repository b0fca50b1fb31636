"""Where the end-to-end time of one render goes (scene upload / render call / teardown). GPU box only."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as ge
pkg = ge.build()
import scenes
from bench import workload
abi = pkg.abi
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
scene = pkg.load_dict(workload(spp))
for it in range(4):
    t0 = time.perf_counter()
    ds = pkg.DeviceScene(scene, device=0)
    t1 = time.perf_counter()
    r = ds.render(seed_mode=abi.SEED_WAVEFRONT)
    t2 = time.perf_counter()
    r2 = ds.render(seed_mode=abi.SEED_WAVEFRONT)
    t3 = time.perf_counter()
    ds.close()
    t4 = time.perf_counter()
    print("iter %d: create %.2f ms, first render %.2f ms (kernel %.2f), second render %.2f ms (kernel %.2f), destroy %.2f ms"
          % (it, (t1 - t0) * 1e3, (t2 - t1) * 1e3, r.totals.gpu_ms, (t3 - t2) * 1e3, r2.totals.gpu_ms, (t4 - t3) * 1e3))
