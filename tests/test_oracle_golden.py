"""CPU tests: the oracle against the reference's own golden vectors / known-answer tests,
and the committed golden fixtures under tests/golden/."""
import ctypes as C
import json
import os

import numpy as np
import pytest

import scenes

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_tea_float32_reference_values(oracle):
    """src/libcore/tests/test_random.py:6-16"""
    L = oracle.lib()
    exp = {(1, 1): 0.5424730777740479, (1, 2): 0.5079904794692993, (1, 3): 0.4171961545944214,
           (1, 4): 0.008385419845581055, (1, 5): 0.8085528612136841, (2, 1): 0.6939879655838013,
           (3, 1): 0.6978365182876587, (4, 1): 0.4897364377975464}
    for (a, b), v in exp.items():
        assert L.b2vp_oracle_tea_float32(a, b, 4) == np.float32(v)


def test_tea_float64_reference_values(oracle):
    """src/libcore/tests/test_random.py:19-28"""
    L = oracle.lib()
    exp = {(1, 1): 0.5424730799533735, (1, 2): 0.5079905082233922, (1, 3): 0.4171962610608142,
           (1, 4): 0.008385529523330604, (1, 5): 0.80855288317879, (2, 1): 0.6939880404156831,
           (3, 1): 0.6978365636630994, (4, 1): 0.48973647949223253}
    for (a, b), v in exp.items():
        assert L.b2vp_oracle_tea_float64(a, b, 4) == v


def test_pcg32_reference_stream(oracle):
    """The published PCG32 demo stream for seed (42, 54), which enoki::PCG32 reproduces
    (ext/enoki/include/enoki/random.h; src/samplers/tests/test_independent.py:31-37 pins sampler == PCG32)."""
    out = np.zeros(6, dtype=np.uint32)
    f32 = np.zeros(6, dtype=np.float32)
    f64 = np.zeros(6, dtype=np.float64)
    oracle.lib().b2vp_oracle_pcg32(42, 54, 6, out.ctypes.data, f32.ctypes.data, f64.ctypes.data)
    assert [hex(x) for x in out] == ["0xa15c02b7", "0x7b47f409", "0xba1d3330", "0x83d2f293", "0xbfa4784b", "0xcbed606e"]
    # next_float32 / next_float64 bit constructions (random.h:113-141)
    assert np.array_equal(f32, ((out >> 9) | 0x3f800000).astype(np.uint32).view(np.float32) - np.float32(1))
    assert np.array_equal(f64, ((out.astype(np.uint64) << np.uint64(20)) | np.uint64(0x3ff0000000000000)).view(np.float64) - 1.0)


def rpv_reference(rho_0, rho_0_hotspot, g, k, theta_i, phi_i, theta_o, phi_o):
    """src/bsdfs/tests/test_rpv.py:35-58"""
    sini, ui = np.sin(theta_i), np.cos(theta_i)
    tan_i = sini / ui
    sino, uo = np.sin(theta_o), np.cos(theta_o)
    tan_o = sino / uo
    cosphi = np.cos(phi_i - phi_o)
    K1 = np.power(ui * uo * (ui + uo), k - 1.)
    cos_g = ui * uo + sini * sino * cosphi
    Fg = (1. - g * g) / np.power(1. + g * g + 2. * g * cos_g, 1.5)
    G = np.sqrt(tan_i * tan_i + tan_o * tan_o - 2. * tan_i * tan_o * cosphi)
    K3 = 1. + (1. - rho_0_hotspot) / (1. + G)
    return K1 * Fg * K3 * rho_0 * np.abs(uo) / np.pi


def test_rpv_eval_against_reference_formula(oracle):
    """src/bsdfs/tests/test_rpv.py:77-107: 27 parameter triples x directions, rtol 1e-3"""
    L = oracle.lib()
    rng = np.random.default_rng(0)
    for rho_0 in (0.004, 0.1, 0.497):
        for k in (0.543, 0.634, 0.851):
            for g in (-0.29, 0.086, 0.2):
                for _ in range(20):
                    ti, to = rng.uniform(0, 1.3, 2)
                    pi_, po = rng.uniform(0, 2 * np.pi, 2)
                    wi = (C.c_double * 3)(np.sin(ti) * np.cos(pi_), np.sin(ti) * np.sin(pi_), np.cos(ti))
                    wo = (C.c_double * 3)(np.sin(to) * np.cos(po), np.sin(to) * np.sin(po), np.cos(to))
                    got = L.b2vp_oracle_rpv_eval(rho_0, k, g, rho_0, wi, wo)
                    ref = rpv_reference(rho_0, rho_0, g, k, ti, pi_, to, po)
                    assert np.isclose(got, ref, rtol=1e-3, atol=1e-9)


def test_oracle_math_close_to_libm(oracle):
    """The fixed IEEE operation sequences of oracle_math.h stay within a few ulp of libm."""
    L = oracle.lib()
    rng = np.random.default_rng(1)
    n = 20000
    for fn, x, ref, tol in [(0, rng.uniform(1e-6, 1.0, n), np.log, 2e-7), (1, rng.uniform(-80, 5, n), np.exp, 3e-7),
                            (3, rng.uniform(-30, 30, n), np.cbrt, 3e-7)]:
        x32 = x.astype(np.float32); o = np.zeros(n, np.float32); o2 = np.zeros(n, np.float32)
        L.b2vp_oracle_math_f32(fn, n, x32.ctypes.data_as(C.c_void_p), None, o.ctypes.data_as(C.c_void_p),
                               o2.ctypes.data_as(C.c_void_p))
        r = ref(x32.astype(np.float64))
        assert np.max(np.abs(o - r) / np.maximum(np.abs(r), 1e-30)) < tol or fn == 0 and np.allclose(o, r, rtol=2e-6, atol=2e-7)
        x64 = x.astype(np.float64); d = np.zeros(n); d2 = np.zeros(n)
        L.b2vp_oracle_math_f64(fn, n, x64.ctypes.data_as(C.c_void_p), None, d.ctypes.data_as(C.c_void_p),
                               d2.ctypes.data_as(C.c_void_p))
        assert np.allclose(d, ref(x64), rtol=1e-14, atol=1e-15)
    x = rng.uniform(-7, 7, n)
    s = np.zeros(n); c = np.zeros(n)
    L.b2vp_oracle_math_f64(2, n, x.ctypes.data_as(C.c_void_p), None, s.ctypes.data_as(C.c_void_p), c.ctypes.data_as(C.c_void_p))
    assert np.allclose(s, np.sin(x), atol=1e-15) and np.allclose(c, np.cos(x), atol=1e-15)


def test_surface_radiance_analytic(pkg, oracle):
    """src/sensors/tests/test_distant.py:313-475 / test_radiancemeter.py:114: directional emitter + diffuse
    rectangle + no medium => L = E cos(theta_e) rho / pi, here through volpath (SURVEY: probe 0.15917827)."""
    for variant in ("scalar_mono", "scalar_mono_double"):
        d = {"type": "scene", "integrator": {"type": "volpath"},
             "ground": {"type": "rectangle", "bsdf": {"type": "diffuse", "reflectance": 0.5}},
             "sun": {"type": "directional", "direction": [0, 0, -1], "irradiance": 1.0},
             "sensor": {"type": "radiancemeter", "origin": [0, 0, 1], "direction": [0, 0, -1],
                        "sampler": {"type": "independent", "sample_count": 2048}, "film": scenes.film(1)}}
        r = oracle.render(pkg, pkg.load_dict(d, variant), pkg.abi.SEED_SCALAR_BLOCKS)
        assert np.isclose(r.bitmap()[0, 0], 0.5 / np.pi, rtol=1e-4)
        # 60 degree sun: cos factor
        d["sun"]["direction"] = [np.sin(np.pi / 3), 0, -np.cos(np.pi / 3)]
        r = oracle.render(pkg, pkg.load_dict(d, variant), pkg.abi.SEED_SCALAR_BLOCKS)
        assert np.isclose(r.bitmap()[0, 0], 0.5 * 0.5 / np.pi, rtol=1e-4)


def test_single_scattering_homogeneous_slab(pkg, oracle):
    """Known answer: max_depth=2 over a black ground, isotropic phase, sun and view at nadir gives the
    single-scattering radiance  L = albedo/(4 pi) * (1 - exp(-2 tau)) / 2  (plane-parallel)."""
    tau, albedo = 0.4, 0.8
    T = pkg.Transform
    d = {"type": "scene", "integrator": {"type": "volpath", "max_depth": 2},
         "atm": {"type": "homogeneous", "sigma_t": tau, "albedo": albedo, "phase": {"type": "isotropic"}},
         "cube": {"type": "cube", "to_world": T.translate([0, 0, 0.5]) @ T.scale([1e4, 1e4, 0.5]),
                  "bsdf": {"type": "null"}, "interior": {"type": "ref", "id": "atm"}},
         "ground": {"type": "rectangle", "to_world": T.translate([0, 0, -1.0]) @ T.scale([2e4, 2e4, 1]),
                    "bsdf": {"type": "diffuse", "reflectance": 0.0}},
         "sun": {"type": "directional", "direction": [0, 0, -1], "irradiance": 1.0},
         "sensor": {"type": "distant", "direction": [0, 0, 1], "ray_target": [0, 0, 0.5],
                    "sampler": {"type": "independent", "sample_count": 200000}, "film": scenes.film(1)}}
    r = oracle.render(pkg, pkg.load_dict(d), pkg.abi.SEED_WAVEFRONT, want_paths=True)
    x = r.radiance.ravel().astype(np.float64)
    expect = albedo / (4 * np.pi) * (1 - np.exp(-2 * tau)) / 2
    assert abs(x.mean() - expect) < 4 * x.std() / np.sqrt(len(x)), (x.mean(), expect)


def test_golden_fixtures(pkg, oracle):
    """Committed golden vectors (tests/golden/make_golden.py): per-stream PCG32 end states, event counters and
    radiance bit patterns of small scenes; guards the oracle itself against drift."""
    with open(os.path.join(GOLDEN, "golden.json")) as f:
        gold = json.load(f)
    import make_golden
    for name, g in gold.items():
        sc = pkg.load_dict(make_golden.CASES[name](g["variant"]), g["variant"])
        r = oracle.render(pkg, sc, pkg.abi.SEED_WAVEFRONT, want_paths=True)
        assert [int(x) for x in r.state] == g["state"], name
        assert r.counters.tolist() == g["counters"], name
        bits = r.radiance.view(np.uint32 if r.radiance.dtype == np.float32 else np.uint64).ravel()
        assert [int(x) for x in bits] == g["radiance_bits"], name


def test_abi_library_exports(pkg):
    """The C-ABI library loads and exports every symbol include/b2vp.h declares (no compute without a GPU)."""
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    header = open(os.path.join(root, "include", "b2vp.h")).read()
    declared = set(re.findall(r"\b(b2vp_[a-z0-9_]+)\s*\(", header)) - {"b2vp_scene_t"}
    assert declared == set(pkg.abi.EXPORTS)
    L = pkg.abi.lib()
    for name in declared:
        assert getattr(L, name) is not None
    assert L.b2vp_abi_version() == pkg.abi.ABI_VERSION
    # struct sizes agree with the header as compiled by gcc
    import subprocess
    import tempfile
    src = ('#include "%s"\n#include <stdio.h>\nint main(){printf("%%zu %%zu %%zu %%zu %%zu %%zu %%zu %%zu %%zu %%zu %%zu\\n",'
           'sizeof(B2vpSceneDesc),sizeof(B2vpRenderArgs),sizeof(B2vpShape),sizeof(B2vpSensor),sizeof(B2vpParamPatch),'
           'sizeof(B2vpVolume),sizeof(B2vpMedium),sizeof(B2vpBsdf),sizeof(B2vpCounters),sizeof(B2vpPathCounters),sizeof(B2vpSceneInfo));}'
           % os.path.join(root, "include", "b2vp.h"))
    with tempfile.TemporaryDirectory() as td:
        open(os.path.join(td, "s.c"), "w").write(src)
        subprocess.check_call(["gcc", os.path.join(td, "s.c"), "-o", os.path.join(td, "s")])
        sizes = [int(x) for x in subprocess.check_output([os.path.join(td, "s")]).split()]
    a = pkg.abi
    assert sizes == [C.sizeof(t) for t in (a.SceneDesc, a.RenderArgs, a.Shape, a.Sensor, a.ParamPatch, a.Volume,
                                           a.Medium, a.Bsdf, a.Counters, a.PathCounters, a.SceneInfo)]


def test_no_device_fails_loudly(pkg):
    """Without a CUDA device the product path must fail, not fall back to the CPU."""
    if pkg.abi.lib().b2vp_device_count() > 0:
        pytest.skip("a CUDA device is present")
    sc = pkg.load_dict(scenes.c1(spp=4))
    with pytest.raises(pkg.B2vpError) as e:
        pkg.render(sc)
    assert e.value.code == pkg.abi.ERR_NO_DEVICE


def test_load_dict_host_logic(pkg):
    """Host-side plugin parsing mirrors the reference's defaults and errors."""
    sc = pkg.load_dict(scenes.c2(spp=4, width=8))
    d = sc.desc
    assert (d.n_shapes, d.n_media, d.n_emitters) == (2, 1, 1)
    assert d.sensor.direction_type == pkg.abi.DIR_SAMPLE_WIDTH and d.sensor.needs_aperture_sample == 1
    assert d.bsdfs[1].type == pkg.abi.BSDF_RPV and d.bsdfs[1].p[3] == d.bsdfs[1].p[0]   # rho_c defaults to rho_0
    assert d.media[0].has_spectral_extinction == 1 and d.media[0].sample_emitters == 1
    assert d.integrator.max_depth == -1 and d.integrator.rr_depth == 5
    # bounding sphere padding (distant.cpp:292-297)
    ext = np.array(d.bbox_max[:]) - np.array(d.bbox_min[:])
    assert d.bsphere_radius > 0.5 * np.linalg.norm(ext)
    with pytest.raises(RuntimeError):
        pkg.load_dict({"type": "scene", "s": {"type": "perspective"}})
    with pytest.raises(RuntimeError):
        bad = scenes.c1(spp=4)
        bad["integrator"]["rr_depth"] = 0
        pkg.load_dict(bad)
    with pytest.raises(RuntimeError):
        pkg.load_dict(scenes.c1(spp=4), variant="scalar_rgb")


def test_vol_file_roundtrip(pkg, tmp_path):
    """.vol layout of src/textures/volume_data.h:42-102: x fastest, then y, then z."""
    data = np.arange(2 * 3 * 4, dtype=np.float32).reshape(4, 3, 2)   # [z, y, x]
    p = str(tmp_path / "g.vol")
    pkg.write_vol(p, data, bbox=(0, 0, 0, 2, 3, 4))
    raw = open(p, "rb").read()
    assert raw[:4] == b"VOL\x03" and np.frombuffer(raw, "<i4", 5, 4).tolist() == [1, 2, 3, 4, 1]
    back, bbox = pkg.read_vol(p)
    assert back.shape == (4, 3, 2, 1) and np.array_equal(back[..., 0], data) and bbox == (0, 0, 0, 2, 3, 4)
    d = scenes.c2(spp=2, width=4)
    d["atm"]["sigma_t"] = {"type": "gridvolume", "filename": p}
    sc = pkg.load_dict(d)
    assert list(sc.desc.volumes[1].res) == [2, 3, 4] and sc.desc.volumes[1].max_value == 23.0


def test_stream_partition_oracle(pkg, oracle):
    """Sharding contract: disjoint stream ranges reproduce the full render."""
    sc = pkg.load_dict(scenes.c2(spp=4, width=8, nz=50))
    full = oracle.render(pkg, sc, pkg.abi.SEED_WAVEFRONT, want_paths=True)
    a = oracle.render(pkg, sc, pkg.abi.SEED_WAVEFRONT, first_stream=0, n_streams=11, want_paths=True)
    b = oracle.render(pkg, sc, pkg.abi.SEED_WAVEFRONT, first_stream=11, n_streams=21, want_paths=True)
    assert np.array_equal(np.concatenate([a.radiance, b.radiance]), full.radiance)
    assert np.allclose(a.film + b.film, full.film, rtol=1e-5, atol=1e-7)
