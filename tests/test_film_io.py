"""film->develop() equivalents of the Python host: OpenEXR (hdrfilm's default file format,
src/films/hdrfilm.cpp:251-343) and PFM, written without any image library.  The OpenEXR files are read back by the
host's own reader and -- where the image has it -- by OpenCV's reader, which is built on the OpenEXR library itself."""
import os

import numpy as np
import pytest


def _result(pkg, h=7, w=13, seed=1):
    rng = np.random.default_rng(seed)
    film = np.zeros((h, w, 5), np.float32)
    film[..., :3] = rng.uniform(0, 2, (h, w, 1))
    film[..., 3] = 1
    film[..., 4] = rng.integers(1, 5, (h, w))
    film[2, 3, :] = 0          # a pixel without samples develops to 0, not NaN
    return pkg.RenderResult(film, None)


@pytest.mark.parametrize("component_format", ["float16", "float32"])
def test_exr_luminance_round_trip(pkg, tmp_path, component_format):
    r = _result(pkg)
    path = str(tmp_path / "y.exr")
    r.write_exr(path, component_format=component_format)
    back = pkg.read_exr(path)
    assert list(back) == ["Y"]
    assert back["Y"].dtype == np.dtype("<f2" if component_format == "float16" else "<f4")
    assert np.array_equal(back["Y"], r.bitmap().astype(back["Y"].dtype))
    assert back["Y"][2, 3] == 0
    head = open(path, "rb").read(8)
    assert head == bytes([0x76, 0x2f, 0x31, 0x01, 2, 0, 0, 0])     # OpenEXR magic, version 2, scan lines


def test_exr_raw_channels(pkg, tmp_path):
    r = _result(pkg, 5, 4)
    path = str(tmp_path / "raw.exr")
    r.write_exr(path, component_format="float32", pixel_format="xyzaw")
    back = pkg.read_exr(path)
    assert sorted(back) == ["A", "W", "X", "Y", "Z"]               # OpenEXR stores channels alphabetically
    for i, n in enumerate("XYZAW"):
        assert np.array_equal(back[n], r.film[..., i])
    with pytest.raises(ValueError):
        r.write_exr(path, pixel_format="rgb")


@pytest.mark.parametrize("component_format", ["float16", "float32"])
def test_exr_is_read_by_the_openexr_library(pkg, tmp_path, component_format):
    os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
    cv2 = pytest.importorskip("cv2")
    r = _result(pkg, 9, 17, seed=3)
    path = str(tmp_path / "y.exr")
    r.write_exr(path, component_format=component_format)
    img = cv2.imread(path, cv2.IMREAD_UNCHANGED)
    if img is None:
        pytest.skip("this OpenCV build has no OpenEXR support")
    want = r.bitmap().astype(np.float16 if component_format == "float16" else np.float32).astype(np.float32)
    assert img.shape == want.shape and np.array_equal(img.astype(np.float32), want)


def test_pfm(pkg, tmp_path):
    r = _result(pkg, 3, 5)
    path = str(tmp_path / "y.pfm")
    r.write_pfm(path)
    raw = open(path, "rb").read()
    assert raw.startswith(b"Pf\n5 3\n-1\n")
    data = np.frombuffer(raw[len(b"Pf\n5 3\n-1\n"):], dtype="<f4").reshape(3, 5)
    assert np.array_equal(data[::-1], r.bitmap())                  # PFM stores the bottom scan line first


def _film_scene(pkg, **film):
    import scenes
    d = scenes.c1(spp=1)
    d["sensor"]["film"] = dict({"type": "hdrfilm", "width": 32, "height": 21, "rfilter": {"type": "box"}}, **film)
    return pkg.load_dict(d)


def test_hdrfilm_crops_and_parameters(pkg):
    """src/films/tests/test_hdrfilm.py:7-33 (construction errors) and :35-74 (crop windows), through load_dict"""
    sn = _film_scene(pkg, crop_width=11, crop_height=5, crop_offset_x=2, crop_offset_y=3).desc.sensor
    assert list(sn.film_size) == [32, 21] and list(sn.crop_size) == [11, 5] and list(sn.crop_offset) == [2, 3]
    # the crop size does not adjust itself: an offset alone pushes the window out of bounds
    with pytest.raises(RuntimeError, match="Invalid crop window"):
        _film_scene(pkg, crop_offset_x=30, crop_offset_y=20)
    sn = _film_scene(pkg, crop_offset_x=30, crop_offset_y=20, crop_width=2, crop_height=1).desc.sensor
    assert list(sn.film_size) == [32, 21] and list(sn.crop_size) == [2, 1] and list(sn.crop_offset) == [30, 20]
    with pytest.raises(RuntimeError, match="component_format"):
        _film_scene(pkg, component_format="uint8")
    with pytest.raises(RuntimeError, match="pixel_format"):
        _film_scene(pkg, pixel_format="brga")
    with pytest.raises(RuntimeError, match="file_format"):
        _film_scene(pkg, file_format="png")
    with pytest.raises(RuntimeError, match="radius of 74 pixels"):
        _film_scene(pkg, rfilter={"type": "gaussian", "stddev": 18.5})   # footprints of up to 16 x 16 pixels are carried (tests/test_rfilter.py)
