"""Tolerance mode (math=fast) on the random scenes of tests/test_gpu_fuzz.py: per scene, the fraction of paths whose
event counters equal the exact mode's and, among those, the fraction within 1e-5 on radiance.  GPU box only.
usage: python tools/fuzz_fast_check.py [n_scenes]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import b2vp_import  # noqa: E402

pkg = b2vp_import.load()
import importlib.util  # noqa: E402

spec = importlib.util.spec_from_file_location("fz", os.path.join(ROOT, "tests", "test_gpu_fuzz.py"))
fz = importlib.util.module_from_spec(spec)
spec.loader.exec_module(fz)

n = int(sys.argv[1]) if len(sys.argv) > 1 else 160
paths = same = close = samples = 0
worst = []
for seed in range(n):
    sc = pkg.load_dict(fz.random_scene(seed, "scalar_mono"), "scalar_mono")
    a = pkg.render(sc, device=0, seed_mode=pkg.abi.SEED_WAVEFRONT, want_paths=True)
    b = pkg.render(sc, device=0, seed_mode=pkg.abi.SEED_WAVEFRONT, want_paths=True, math="fast")
    eq = (a.counters == b.counters).all(axis=1) & (a.state == b.state)
    ra, rb = a.radiance[eq].astype(np.float64), b.radiance[eq].astype(np.float64)
    ok = np.abs(ra - rb) <= 1e-5 * np.maximum(np.abs(ra), 1e-3 * max(np.abs(ra).max(), 1e-30)) if len(ra) else np.ones(0, bool)
    paths += len(eq); same += int(eq.sum()); close += int(ok.sum()); samples += int(ra.size)
    ma, mb = a.film[..., 1].sum(), b.film[..., 1].sum()
    if eq.mean() < 0.9 or (len(ok) and ok.mean() < 0.9):
        worst.append((seed, round(float(eq.mean()), 3), round(float(ok.mean()) if len(ok) else 1.0, 3), float(ma), float(mb)))
print("scenes %d, streams %d: same event sequence %.4f; %d samples of those, within 1e-5: %.4f" % (n, paths, same / paths, samples, close / max(samples, 1)))
print("scenes below 0.9:", worst)
