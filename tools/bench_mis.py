"""volpathmis on C2 (256 x 1, 1024 spp) on the GPU: device-timed Mpaths/s of the one-thread-per-stream kernel."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as ge
pkg = ge.build()
import scenes
d = scenes.c2(spp=1024, width=256)
d["integrator"]["type"] = "volpathmis"
with pkg.DeviceScene(pkg.load_dict(d), device=0) as ds:
    ds.render(seed_mode=pkg.abi.SEED_WAVEFRONT)
    ms = [ds.render(seed_mode=pkg.abi.SEED_WAVEFRONT).totals.gpu_ms for _ in range(3)]
    t = ds.render(seed_mode=pkg.abi.SEED_WAVEFRONT).totals
print(json.dumps({"config": "C2 volpathmis", "paths": t.n_paths, "kernel_ms": min(ms), "mpaths_s": t.n_paths / min(ms) / 1e3,
                  "segments_per_s": (t.n_medium_steps + t.n_intersect) / (min(ms) * 1e-3)}))
