"""Random scenes (tests/test_gpu_fuzz.py::random_scene) rendered by the reference binary and, from the same XML through the
plugin's describe(), by the CPU oracle -- the by-hand version of tests/test_reference_plugin.py::
test_random_scenes_against_the_reference for thousands of seeds.  Needs oracle/_ref (the build container).
usage: python tools/fuzz_vs_reference.py N [variant]      prints every scene whose film is off by more than 1e-9"""
import sys, numpy as np, time
import os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
import b2vp_import; pkg = b2vp_import.load()
from oracle import oracle, ref_xml
import importlib.util
spec = importlib.util.spec_from_file_location("fz", os.path.join(ROOT, "tests", "test_gpu_fuzz.py"))
fz = importlib.util.module_from_spec(spec); spec.loader.exec_module(fz)
variant = sys.argv[2] if len(sys.argv) > 2 else "scalar_mono_double"
t0=time.time(); worst=0; bad=[]; n=0
for seed in range(int(sys.argv[1])):
    d = ref_xml.round_f32(fz.random_scene(seed, variant))
    try:
        want, _ = ref_xml.render(d, variant, write_vol=pkg.scene.write_vol)
    except Exception as e:
        print("REF FAIL", seed, str(e)[-160:].replace("\n"," ")); continue
    try:
        got, _ = ref_xml.render(ref_xml.with_plugin(d, seeding="scalar"), variant, write_vol=pkg.scene.write_vol, describe_oracle=True)
    except Exception as e:
        print("PLUGIN FAIL", seed, str(e)[-200:].replace("\n"," ")); continue
    n += 1
    scale = np.abs(want).max(axis=(0,1)) + 1e-30
    err = (np.abs(got - want).max(axis=(0,1)) / scale).max()
    worst = max(worst, err)
    if err > 1e-9: bad.append(seed); print("MISMATCH", seed, err, d.get("sensor", {}).get("type"), d["integrator"]["type"], d["atm"]["type"], d.get("sensor", {}).get("film", {}).get("rfilter"))
print("compared", n, "worst", worst, "bad", bad, round(time.time()-t0,1), "s")
