"""The C++ host layer (include/b2vp.hpp) through the C-ABI: tests/cpp/test_host.cpp builds scenes with the
plugin-style builder, renders them with libb2vp.so and checks them against the oracle."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "tests", "cpp", "_build")
EXE = os.path.join(BUILD, "test_host")


def build_exe(pkg, oracle):
    oracle.lib()  # makes sure oracle/libb2vp_oracle.so exists
    src = os.path.join(ROOT, "tests", "cpp", "test_host.cpp")
    deps = [src, os.path.join(ROOT, "include", "b2vp.hpp"), os.path.join(ROOT, "include", "b2vp.h")]
    if os.path.exists(EXE) and all(os.path.getmtime(EXE) > os.path.getmtime(d) for d in deps):
        return EXE
    os.makedirs(BUILD, exist_ok=True)
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"), src, "-o", EXE,
                           "-L" + os.path.join(ROOT, "eradiate-kernel_b200"), "-lb2vp",
                           "-L" + os.path.join(ROOT, "oracle"), "-lb2vp_oracle",
                           "-Wl,-rpath,$ORIGIN/../../../eradiate-kernel_b200", "-Wl,-rpath,$ORIGIN/../../../oracle"])
    return EXE


def test_cpp_host_builds_and_fails_loudly_without_gpu(pkg, oracle):
    exe = build_exe(pkg, oracle)
    out = subprocess.run([exe, "cpu"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "C1 (C++ host, oracle)" in out.stdout


def ext_files(pkg, tmp_path):
    """The files tests/cpp/test_host.cpp::scene_ext loads: a 1 x 1 x 64 `.vol` column and the OBJ fixture."""
    import shutil
    import numpy as np
    z = (np.arange(64) + 0.5) / 64
    pkg.write_vol(str(tmp_path / "col.vol"), np.exp(-z * 100.0 / 8.0).astype(np.float32).reshape(64, 1, 1))
    shutil.copy(os.path.join(ROOT, "tests", "golden", "data", "roof.obj"), str(tmp_path / "roof.obj"))


def ext_dict(pkg, tmp_path, variant, sensor, spp):
    """scene_ext of tests/cpp/test_host.cpp as a `load_dict` dictionary."""
    import numpy as np
    import scenes
    T, dt = scenes.T, pkg.scene.VARIANTS[variant][0]
    n = 8
    d = {
        "type": "scene",
        "integrator": {"type": "volpath", "max_depth": -1},
        "atm": {"type": "heterogeneous",
                "sigma_t": {"type": "gridvolume", "filename": str(tmp_path / "col.vol"),
                            "to_world": T.translate([-10, -10, 0], dt) @ T.scale([20, 20, 1], dt)},
                "scale": 4.0, "albedo": 1.0, "phase": {"type": "rayleigh"}},
        "cube": {"type": "cube", "to_world": T.translate([0, 0, 0.45], dt) @ T.scale([10, 10, 0.55], dt),
                 "bsdf": {"type": "null"}, "interior": {"type": "ref", "id": "atm"}},
        "ground": {"type": "rectangle", "to_world": T.scale([10, 10, 1], dt),
                   "bsdf": {"type": "rpv", "rho_0": 0.1, "k": 0.7, "g": -0.1}},
        "ball": {"type": "sphere", "center": [0.5, -0.25, 0.5], "radius": 0.25,
                 "bsdf": {"type": "diffuse", "reflectance": 0.75}},
        "roof": {"type": "obj", "filename": str(tmp_path / "roof.obj"),
                 "to_world": T.translate([0.25, -0.5, 0.125], dt) @ T.scale([2.0, 1.5, 1.0], dt),
                 "bsdf": {"type": "diffuse", "reflectance": 0.75}},
        "sun": {"type": "directional", "direction": [0.5, 0, -0.8660254037844386], "irradiance": 1.0},
    }
    film = {"type": "hdrfilm", "width": n, "height": 1, "rfilter": {"type": "box"}}
    sampler = {"type": "independent", "sample_count": spp}
    if sensor == "mdistant":
        th = (np.arange(n) + 0.5) / n * np.pi
        dirs = np.stack([-np.cos(th), np.zeros_like(th), -np.sin(th)], axis=1)
        d["sensor"] = {"type": "mdistant", "directions": dirs.reshape(-1).tolist(), "target": [0, 0, 0],
                       "sampler": sampler, "film": film}
    else:
        i = np.arange(n)
        org = np.stack([-1.0 + 0.25 * i, 0.125 * i - 0.5, np.full(n, 2.0)], axis=1)
        dirs = np.stack([0.0625 * i, -0.03125 * i, np.full(n, -1.0)], axis=1)
        d["sensor"] = {"type": "mradiancemeter", "origins": org.reshape(-1).tolist(), "directions": dirs.reshape(-1).tolist(),
                       "sampler": sampler, "film": film}
    return d


@pytest.mark.parametrize("sensor", ["mdistant", "mradiancemeter"])
@pytest.mark.parametrize("prec", ["f32", "f64"])
def test_cpp_host_loaders_and_sensors_match_the_python_host(pkg, oracle, prec, sensor, tmp_path):
    """read_vol, load_obj + add_mesh(to_world), add_sphere and the mdistant / mradiancemeter sensors of b2vp.hpp give
    the scene `load_dict` gives (which is pinned on the reference binary): same film through the oracle.  The films
    the C++ host develops (PFM, OpenEXR) hold Y / W of that film."""
    import numpy as np
    exe = build_exe(pkg, oracle)
    ext_files(pkg, tmp_path)
    out = tmp_path / "film.bin"
    subprocess.check_call([exe, "ext", str(tmp_path), prec, sensor, str(out)])
    variant = "scalar_mono_double" if prec == "f64" else "scalar_mono"
    dtype = np.float64 if prec == "f64" else np.float32
    got = np.fromfile(str(out), dtype=dtype).reshape(1, 8, 5)
    sc = pkg.load_dict(ext_dict(pkg, tmp_path, variant, sensor, 64), variant)
    ref = oracle.render(pkg, sc, pkg.abi.SEED_WAVEFRONT).film
    assert np.array_equal(got[..., 3:], ref[..., 3:])                      # alpha and weights
    assert np.allclose(got[..., :3], ref[..., :3], rtol=1e-12 if prec == "f64" else 1e-5, atol=0), \
        np.abs(got[..., 1] / ref[..., 1] - 1).max()
    lum = (got[0, :, 1] / got[0, :, 4]).astype(np.float32)
    exr = pkg.read_exr(str(tmp_path / "film.exr"))
    assert list(exr) == ["Y"] and np.array_equal(exr["Y"][0], lum)
    raw = open(str(tmp_path / "film.pfm"), "rb").read()
    assert raw.startswith(b"Pf\n8 1\n-1\n") and np.array_equal(np.frombuffer(raw[len(b"Pf\n8 1\n-1\n"):], "<f4"), lum)


@pytest.mark.gpu
def test_cpp_host_matches_oracle_on_gpu(pkg, oracle, tmp_path):
    exe = build_exe(pkg, oracle)
    ext_files(pkg, tmp_path)
    out = subprocess.run([exe, "gpu", str(tmp_path)], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout + out.stderr
    assert out.stdout.count("-> ok") == 9, out.stdout      # C1 / C2 x 2 precisions, scene_ext x 2 sensors x 2, options


@pytest.mark.parametrize("prec", ["f32", "f64"])
def test_cpp_host_scene_matches_reference_binary(pkg, oracle, prec, tmp_path):
    """The scene the C++ host layer builds (SceneBuilder<Float>: derived quantities computed in C++) gives the film
    the reference binary renders from the same parameters (fixture c2_cpp_host*, tests/golden/make_ref_golden.py)."""
    import numpy as np
    from test_reference_parity import REF, check_against_reference
    exe = build_exe(pkg, oracle)
    out = tmp_path / "film.bin"
    subprocess.check_call([exe, "dump", prec, str(out)])
    name = "c2_cpp_host" + ("_f64" if prec == "f64" else "")
    film = np.fromfile(str(out), dtype=np.float64 if prec == "f64" else np.float32).reshape(1, 256, 5)
    check_against_reference(name, film, REF[name], "scalar_mono_double" if prec == "f64" else "scalar_mono")
