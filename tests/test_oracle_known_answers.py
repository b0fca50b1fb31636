"""Known-answer tests of the reference for sensors and shapes, restated against the oracle (CPU only).
Each test cites the reference test it follows; the scene dictionaries are the reference's."""
import os

import numpy as np
import pytest

import scenes  # noqa: F401  (sets up the package for `T`)
from scenes import T, film

SAMPLES = [([0.32, 0.87], [0.16, 0.44]), ([0.17, 0.44], [0.22, 0.81]),
           ([0.12, 0.82], [0.99, 0.42]), ([0.72, 0.40], [0.01, 0.61])]


def sensor_scene(pkg, sensor, shapes=None, variant="scalar_mono_double"):
    d = {"type": "scene", "integrator": {"type": "volpath"},
         "sun": {"type": "directional", "direction": [0, 0, -1]}, "sensor": sensor}
    d.update(shapes or {"shape": {"type": "rectangle"}})
    return pkg.load_dict(d, variant)


def distant(w, h, **kw):
    return dict({"type": "distant", "film": film(w, h), "sampler": {"type": "independent", "sample_count": 1}}, **kw)


@pytest.mark.parametrize("direction", [[0.0, 0.0, 1.0], [-1.0, -1.0, 0.0], [2.0, 0.0, 0.0]])
def test_distant_direction_1x1(pkg, oracle, direction):
    """src/sensors/tests/test_distant.py:141-156: ray.d == -normalize(direction)"""
    sc = sensor_scene(pkg, distant(1, 1, direction=direction))
    for s1, s2 in SAMPLES:
        _, d, _ = oracle.sensor_ray(sc, s1[0], s1[1], s2[0], s2[1])
        assert np.allclose(d, -np.asarray(direction) / np.linalg.norm(direction), atol=1e-12)


@pytest.mark.parametrize("orientation", [[1.0, 0.0, 0.0], [0.0, 1.0, 0.0], [1.0, 1.0, 0.0]])
def test_distant_direction_32x1(pkg, oracle, orientation):
    """test_distant.py:159-182: with an N x 1 film the directions stay in the plane (direction, orientation)"""
    direction = [0.0, 0.0, 1.0]
    sc = sensor_scene(pkg, distant(32, 1, direction=direction, orientation=orientation))
    excluded = np.cross(direction, orientation)
    for s1, s2 in SAMPLES:
        _, d, _ = oracle.sensor_ray(sc, s1[0], s1[1], s2[0], s2[1])
        assert abs(np.dot(d, excluded)) < 1e-12


def test_distant_flip_directions(pkg, oracle):
    """test_distant.py:185-210"""
    a = sensor_scene(pkg, distant(32, 32, direction=[0, 0, 1], flip_directions=False))
    b = sensor_scene(pkg, distant(32, 32, direction=[0, 0, 1], flip_directions=True))
    for s1, s2 in SAMPLES:
        _, d1, _ = oracle.sensor_ray(a, s1[0], s1[1], s2[0], s2[1])
        _, d2, _ = oracle.sensor_ray(b, s1[0], s1[1], s2[0], s2[1])
        assert np.allclose(d1, -d2, atol=1e-12)


def test_distant_origin_on_bounding_sphere(pkg, oracle):
    """test_distant.py:218-251: default origins lie on the scene's bounding sphere"""
    sc = sensor_scene(pkg, distant(1, 1, direction=[0, 0, 1]))
    radius = np.linalg.norm([1.0, 1.0, 0.0])  # bbox of the unit rectangle: centre 0, max (1, 1, 0)
    for s1, s2 in SAMPLES:
        o, _, _ = oracle.sensor_ray(sc, s1[0], s1[1], s2[0], s2[1])
        assert np.isclose(o[2], radius, rtol=1e-3)


@pytest.mark.parametrize("direction", [[0.0, 0.0, 1.0], [-1.0, -1.0, 0.0], [2.0, 0.0, 0.0]])
@pytest.mark.parametrize("origin", [[0.0, 0.0, 0.0], [-1.0, -1.0, 0.5], [4.0, 1.0, 0.0]])
def test_radiancemeter_sample_ray(pkg, oracle, direction, origin):
    """src/sensors/tests/test_radiancemeter.py:92-103: ray.o == origin, ray.d == normalize(direction)"""
    sc = sensor_scene(pkg, {"type": "radiancemeter", "origin": origin, "direction": direction, "film": film(1, 1),
                            "sampler": {"type": "independent", "sample_count": 1}})
    o, d, w = oracle.sensor_ray(sc, 0.32, 0.87, 0.16, 0.44)
    assert np.allclose(o, origin, atol=1e-12)
    assert np.allclose(d, np.asarray(direction) / np.linalg.norm(direction), atol=1e-12)
    assert w == 1.0


def test_radiancemeter_construct_errors(pkg):
    """test_radiancemeter.py:82-90: missing origin or direction, or a film that is not 1x1, raise"""
    for bad in ({"direction": [0, 1, 0]}, {"origin": [0, 1, 0]}):
        with pytest.raises(RuntimeError):
            sensor_scene(pkg, dict({"type": "radiancemeter", "film": film(1, 1)}, **bad))
    with pytest.raises(RuntimeError):
        sensor_scene(pkg, {"type": "radiancemeter", "origin": [0, 0, 0], "direction": [0, 0, 1], "film": film(2, 1)})


@pytest.mark.parametrize("variant", ["scalar_mono", "scalar_mono_double"])
def test_rectangle_ray_intersect(pkg, oracle, variant):
    """src/shapes/tests/test_rectangle.py:37-63: scale (2, 0.5, 1); rays from (a, a, 5) downwards hit iff |a| <= 0.5"""
    sc = sensor_scene(pkg, distant(1, 1, direction=[0, 0, 1]),
                      {"foo": {"type": "rectangle", "to_world": T.scale([2.0, 0.5, 1.0])}}, variant)
    coords = np.linspace(-1, 1, 15)
    valid = 0
    for a in coords:
        h = oracle.ray_intersect(sc, [a, a, 5], [0, 0, -1], mint=0.0)
        assert (h["t"] is not None) == (abs(a) <= 0.5)
        if h["t"] is not None:
            valid += 1
            assert np.isclose(h["t"], 5.0) and np.allclose(h["p"], [a, a, 0], atol=1e-6)
            assert np.allclose(h["n"], [0, 0, 1]) and h["shape"] == 0
    assert valid == 7


@pytest.mark.parametrize("r", [1, 3])
def test_sphere_ray_intersect_transform(pkg, oracle, r):
    """src/shapes/tests/test_sphere.py:52-75: radius r, translate(0,1,0) * rotate(y, 30 deg); rays along +z
    hit iff x^2 + y^2 <= r^2, at t = 8 - sqrt(r^2 - x^2 - y^2)"""
    sc = sensor_scene(pkg, distant(1, 1, direction=[0, 0, 1]),
                      {"s": {"type": "sphere", "radius": r,
                             "to_world": T.translate([0, 1, 0]) @ T.rotate([0, 1, 0], 30.0)}})
    n = 21
    for x in range(n):
        for y in range(n):
            xc, yc = r * (2 * (x / n) - 1), r * (2 * (y / n) - 1)
            h = oracle.ray_intersect(sc, [xc, yc + 1, -8], [0, 0, 1], mint=0.0)
            rr = xc * xc + yc * yc
            assert (h["t"] is not None) == (rr <= r * r) or abs(rr - r * r) < 1e-8
            if h["t"] is not None and rr < r * r - 1e-6:
                assert np.isclose(h["t"], 8 - np.sqrt(r * r - rr), rtol=1e-6)  # to_world is built in float32
                assert np.isclose(np.linalg.norm(h["p"] - [0, 1, 0]), r, rtol=1e-6)


def test_cube_faces(pkg, oracle):
    """src/shapes/tests/test_cube.py:25-40 (bbox [-1,1]^3, area 24) as ray casts: every axis-aligned ray through
    the centre enters at distance 4 through the face whose outward normal opposes it"""
    sc = sensor_scene(pkg, distant(1, 1, direction=[0, 0, 1]), {"c": {"type": "cube"}})
    assert np.allclose(sc.desc.bbox_min[:], [-1, -1, -1]) and np.allclose(sc.desc.bbox_max[:], [1, 1, 1])
    for axis in range(3):
        for sign in (-1.0, 1.0):
            o = np.zeros(3); o[axis] = -5 * sign
            d = np.zeros(3); d[axis] = sign
            h = oracle.ray_intersect(sc, o + [0.3, 0.2, 0.1] * (1 - np.abs(d)), d, mint=0.0)
            assert np.isclose(h["t"], 4.0) and np.allclose(h["n"], -d)


def test_film_develop_pfm(pkg, tmp_path):
    """RenderResult.write_pfm: luminance = Y / W, float32, scanlines bottom to top (bitmap.cpp:2206-2230)."""
    film = np.zeros((2, 3, 5), dtype=np.float32)
    film[..., 1] = np.arange(6, dtype=np.float32).reshape(2, 3) * 2
    film[..., 4] = 2
    film[1, 2, 4] = 0          # a pixel without samples develops to 0
    r = pkg.RenderResult(film, None)
    path = tmp_path / "out.pfm"
    r.write_pfm(str(path))
    raw = path.read_bytes()
    assert raw.startswith(b"Pf\n3 2\n-1\n")
    data = np.frombuffer(raw[len(b"Pf\n3 2\n-1\n"):], dtype="<f4").reshape(2, 3)
    assert np.array_equal(data, np.array([[3, 4, 0], [0, 1, 2]], dtype=np.float32))


# ---- `.serialized` mesh containers (src/shapes/serialized.cpp) -----------------------------------------
def _roof(pkg):
    return pkg.scene.read_ply(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "data", "roof_bin.ply"))


@pytest.mark.parametrize("version,double", [(4, False), (4, True), (3, False), (3, True)])
def test_serialized_round_trip(pkg, tmp_path, version, double):
    """Every (container version, precision) pair, three sub-meshes, each reached through the offset dictionary."""
    pos, faces, nrm, uv = _roof(pkg)
    tri = (np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], np.float32), np.array([[0, 1, 2]], np.uint32))
    path = str(tmp_path / "m.serialized")
    pkg.scene.write_serialized(path, [("a", tri[0], tri[1], None, None), ("roof", pos, faces, nrm, uv),
                                      ("b", pos, faces, None, uv)], version=version, double=double)
    p, f, n, t = pkg.scene.read_serialized(path, 1)
    assert np.array_equal(p, pos) and np.array_equal(f, faces) and np.array_equal(n, nrm) and np.array_equal(t, uv)
    p, f, n, t = pkg.scene.read_serialized(path, 2)
    assert np.array_equal(p, pos) and n is None and np.array_equal(t, uv)
    p, f, n, t = pkg.scene.read_serialized(path)   # shape_index 0: read from the start of the file
    assert np.array_equal(p, tri[0]) and np.array_equal(f, tri[1]) and n is None and t is None
    assert p.dtype == np.float32 and f.dtype == np.uint32


def test_serialized_errors(pkg, tmp_path):
    """The reference's failure cases (serialized.cpp:184-221): missing file, bad magic, bad version, bad index."""
    rd = pkg.scene.read_serialized
    with pytest.raises(RuntimeError, match="file not found"):
        rd(str(tmp_path / "absent.serialized"))
    path = str(tmp_path / "m.serialized")
    pos, faces, nrm, uv = _roof(pkg)
    pkg.scene.write_serialized(path, [("roof", pos, faces, nrm, uv)])
    with pytest.raises(RuntimeError, match="shape index must be nonnegative"):
        rd(path, -1)
    with pytest.raises(RuntimeError, match="out of range"):
        rd(path, 2)
    raw = bytearray(open(path, "rb").read())
    bad = str(tmp_path / "bad.serialized")
    open(bad, "wb").write(bytes([0, 0]) + bytes(raw[2:]))
    with pytest.raises(RuntimeError, match="invalid file format"):
        rd(bad)
    open(bad, "wb").write(bytes(raw[:2]) + bytes([5, 0]) + bytes(raw[4:]))
    with pytest.raises(RuntimeError, match="incompatible file version"):
        rd(bad)


@pytest.mark.parametrize("features", ["normals", "uv", "normals_uv"])
def test_serialized_reference_files(pkg, features):
    """The reference's own test meshes and the values its test asserts (src/librender/tests/test_mesh.py:130-172,
    test06_load_various_features).  The files live in the reference tree: skipped where that is absent."""
    path = "/root/reference/resources/data/tests/serialized/rectangle_%s.serialized" % features
    if not os.path.exists(path):
        pytest.skip("reference tree not present")
    pos, faces, nrm, uv = pkg.scene.read_serialized(path)
    assert faces.shape == (2, 3)
    assert np.allclose(pos[0], [-2.85, 0.0, -7.6], atol=1e-3)
    assert np.allclose(pos[2], [2.85, 0.0, 0.599999], atol=1e-3)
    assert np.allclose(pos[3], [2.85, 0.0, -7.6], atol=1e-3)
    if "uv" in features:
        assert np.allclose(uv[0], [0.950589, 1 - 0.988416], atol=1e-3)
        assert np.allclose(uv[2], [0.025105, 1 - 0.689127], atol=1e-3)
        assert np.allclose(uv[3], [0.950589, 1 - 0.689127], atol=1e-3)
    else:
        assert uv is None
    if "normals" in features:
        assert np.allclose(nrm[[0, 2, 3]], [0.0, 1.0, 0.0])
    else:
        assert nrm is None
