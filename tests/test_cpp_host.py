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


@pytest.mark.gpu
def test_cpp_host_matches_oracle_on_gpu(pkg, oracle):
    exe = build_exe(pkg, oracle)
    out = subprocess.run([exe, "gpu"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    assert out.stdout.count("-> ok") == 4


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
