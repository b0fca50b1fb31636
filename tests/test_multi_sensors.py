"""The reference's known-answer tests for the two multi-sensor plugins, restated:
src/sensors/tests/test_mdistant.py:100-118 (ray directions per film column) and
src/sensors/tests/test_mradiancemeter.py:57-107,156-249 (construction errors, which sensor a film position selects,
a constant environment seen by every sensor, three rectangles of different reflectance under one sun).  The reference
runs them in scalar_rgb with `path`; here they run in the mono variants with `volpath` -- no medium is involved, so the
integrators coincide -- on the CPU through the oracle and on the device through the C-ABI."""
import numpy as np
import pytest

from scenes import T, film

SAMPLER = {"type": "independent", "sample_count": 1}


def mdistant_scene(pkg, directions, n, variant="scalar_mono_double", **kw):
    d = {"type": "scene", "integrator": {"type": "volpath"}, "sun": {"type": "directional", "direction": [0, 0, -1]},
         "shape": {"type": "rectangle"},
         "sensor": dict({"type": "mdistant", "directions": directions, "film": film(n), "sampler": SAMPLER}, **kw)}
    return pkg.load_dict(d, variant)


def mradiancemeter_dict(origins, directions, n, spp=1, **scene):
    d = {"type": "scene", "integrator": {"type": "volpath"},
         "sensor": {"type": "mradiancemeter", "origins": origins, "directions": directions, "film": film(n),
                    "sampler": {"type": "independent", "sample_count": spp}}}
    d.update(scene)
    return d


def test_mdistant_sample_ray_direction(pkg, oracle):
    """test_mdistant.py:100-118: film column i looks along directions[i]"""
    sc = mdistant_scene(pkg, [0, 0, -1, -1, -1, 0, -2, 0, 0], 3, target=[0, 0, 0])
    for s1, s2, expected in [([0.32, 0.87], [0.16, 0.44], [0, 0, -1]), ([0.17, 0.44], [0.22, 0.81], [0, 0, -1]),
                             ([0.51, 0.82], [0.99, 0.42], [-1, -1, 0]), ([0.72, 0.40], [0.01, 0.61], [-2, 0, 0])]:
        _, d, _ = oracle.sensor_ray(sc, s1[0], s1[1], s2[0], s2[1])
        assert np.allclose(d, np.asarray(expected) / np.linalg.norm(expected), atol=1e-7)


def test_mdistant_film_size_must_match(pkg):
    """mdistant.cpp:168-172: the film must be [sensor_count, 1]"""
    with pytest.raises(RuntimeError, match="Film size"):
        mdistant_scene(pkg, [0, 0, -1, -1, -1, 0], 3)


def test_mradiancemeter_construction(pkg):
    """test_mradiancemeter.py:57-82"""
    for origins, directions in (([0, 0, 0], [1, 0, 0]), ([0, 0, 0] * 2, [1, 0, 0, -1, 0, 0])):
        assert pkg.load_dict(mradiancemeter_dict(origins, directions, len(origins) // 3, sky={"type": "constant"}))
    with pytest.raises(RuntimeError):
        pkg.load_dict(mradiancemeter_dict([0, 0, 0], [1, 0, 0], 2, sky={"type": "constant"}))   # film wider than the sensor count
    with pytest.raises((RuntimeError, ValueError)):
        pkg.load_dict(mradiancemeter_dict([0, 0], [1, 0], 1, sky={"type": "constant"}))        # not triples
    with pytest.raises((RuntimeError, ValueError)):
        pkg.load_dict(mradiancemeter_dict([0, 0, 0], [1, 0], 1, sky={"type": "constant"}))     # unequal counts


def test_mradiancemeter_ray_sampling(pkg, oracle):
    """test_mradiancemeter.py:85-107: the first film coordinate selects the sensor"""
    import random
    sc = pkg.load_dict(mradiancemeter_dict([0, 0, 0, 1, 0, 1], [1, 0, 0, 1, 1, 1], 2, sky={"type": "constant"}),
                       "scalar_mono_double")
    random.seed(42)
    for _ in range(10):
        random.random()
        px, py = random.random(), random.random()
        o, d, w = oracle.sensor_ray(sc, px, py, 0.0, 0.0)
        if px < 0.5:
            assert np.allclose(o, [0, 0, 0]) and np.allclose(d, [1, 0, 0], atol=1e-7)
        else:
            assert np.allclose(o, [1, 0, 1]) and np.allclose(d, np.ones(3) / np.sqrt(3), atol=1e-7)
        assert w == 1.0


def _render(pkg, oracle, sc, on_device):
    if on_device:
        return pkg.render(sc, device=0, seed_mode=pkg.abi.SEED_WAVEFRONT)
    return oracle.render(pkg, sc, pkg.abi.SEED_WAVEFRONT)


def check_constant_radiance(pkg, oracle, radiance, variant, on_device):
    """test_mradiancemeter.py:156-168: every sensor sees the environment's radiance"""
    d = mradiancemeter_dict([1, 0, 0, 0, 1, 0, 0, 0, 1], [1, 0, 0, 0, -1, 0, 0, 0, -1], 3,
                            sky={"type": "constant", "radiance": radiance})
    img = _render(pkg, oracle, pkg.load_dict(d, variant), on_device).bitmap()
    assert np.allclose(img, np.float32(radiance) if variant == "scalar_mono" else float(np.float32(radiance)), rtol=1e-6)


def check_three_rectangles(pkg, oracle, variant, on_device):
    """test_mradiancemeter.py:171-249: reflectances 1, 0.5, 0 under one sun -> radiances 2 : 1 : 0"""
    rect = lambda x, rho: {"type": "rectangle", "to_world": T.translate([x, 0, 0]), "bsdf": {"type": "diffuse", "reflectance": rho}}
    d = mradiancemeter_dict([-2, 0, 1, 0, 0, 1, 2, 0, 1], [0, 0, -1, 0, 0, -1, 0, 0, -1], 3, spp=16,
                            emitter={"type": "directional", "direction": [0, 0, -1], "irradiance": 1.0},
                            light_rectangle=rect(-2, 1.0), medium_rectangle=rect(0, 0.5), dark_rectangle=rect(2, 0.0))
    data = np.squeeze(_render(pkg, oracle, pkg.load_dict(d, variant), on_device).bitmap())
    assert np.isclose(data[0] / data[1], 2, atol=1e-3) and data[2] == 0
    assert np.isclose(data[0], 1 / np.pi, rtol=1e-5)          # L = E rho / pi


@pytest.mark.parametrize("variant", ["scalar_mono", "scalar_mono_double"])
@pytest.mark.parametrize("radiance", [10.0 ** x for x in range(-3, 4)])
def test_mradiancemeter_render_constant(pkg, oracle, radiance, variant):
    check_constant_radiance(pkg, oracle, radiance, variant, False)


@pytest.mark.parametrize("variant", ["scalar_mono", "scalar_mono_double"])
def test_mradiancemeter_render_complex(pkg, oracle, variant):
    check_three_rectangles(pkg, oracle, variant, False)


@pytest.mark.gpu
@pytest.mark.parametrize("variant", ["scalar_mono", "scalar_mono_double"])
def test_mradiancemeter_renders_on_the_device(pkg, oracle, variant):
    for radiance in (1e-3, 1.0, 1e3):
        check_constant_radiance(pkg, oracle, radiance, variant, True)
    check_three_rectangles(pkg, oracle, variant, True)
