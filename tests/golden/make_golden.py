"""Generates tests/golden/golden.json from the CPU oracle (run from the repo root:
`python tests/golden/make_golden.py`).  The vectors pin per-stream PCG32 end states, event counters and
radiance bit patterns; the -m gpu tests compare the CUDA path against the same fixtures."""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402
import scenes  # noqa: E402

BASE_CASES = {
    "c1": lambda v: scenes.c1(spp=48, variant=v),
    "c2": lambda v: scenes.c2(spp=2, variant=v, width=24, nz=100),
    "c2_mdistant": lambda v: scenes.c2(spp=2, variant=v, width=16, nz=50, sensor="mdistant"),
    "c3": lambda v: scenes.c3(spp=1, variant=v, n=8, size=6),
    "c5": lambda v: scenes.c5(spp=3, variant=v, n_leaves=60, n_sensors=16),
}


def main():
    from oracle import oracle
    pkg = scenes.pkg
    out = {}
    for name, fn in BASE_CASES.items():
        for variant in ("scalar_mono", "scalar_mono_double"):
            sc = pkg.load_dict(fn(variant), variant)
            r = oracle.render(pkg, sc, pkg.abi.SEED_WAVEFRONT, want_paths=True)
            bits = r.radiance.view(np.uint32 if r.radiance.dtype == np.float32 else np.uint64).ravel()
            key = name + ("_f64" if variant.endswith("double") else "")
            out[key] = {"variant": variant, "state": [int(x) for x in r.state], "counters": r.counters.tolist(),
                        "radiance_bits": [int(x) for x in bits]}
    with open(os.path.join(HERE, "golden.json"), "w") as f:
        json.dump(out, f)
    print("wrote", len(out), "cases")


# golden.json keys "<case>" / "<case>_f64" both map to the same scene builder
CASES = dict(BASE_CASES)
for _k in BASE_CASES:
    CASES[_k + "_f64"] = BASE_CASES[_k]

if __name__ == "__main__":
    main()
