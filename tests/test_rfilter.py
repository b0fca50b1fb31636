"""Reconstruction filters wider than a pixel (src/rfilters/{tent,gaussian,mitchell,catmullrom,lanczos,box}.cpp,
src/libcore/rfilter.cpp:9-21, ImageBlock::put's footprint loop src/librender/imageblock.cpp:117-152).

The films of tests/golden/ref_golden.npz under the names of make_ref_golden.RFILTER_CASES were rendered by the
reference binary.  With these filters the W and A channels are sums of weight products, not counts, so the films are
compared with a summation-order tolerance instead of check_against_reference's exact channels."""
import os
import sys

import numpy as np
import pytest

import scenes

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_ref_golden  # noqa: E402

REF = np.load(os.path.join(HERE, "golden", "ref_golden.npz"))
CASES = sorted(make_ref_golden.RFILTER_CASES)


def check_film(name, got, ref, variant, paths_may_diverge=True):
    """W: every sample lands on the same pixels with the same weights (table, footprint and clipping are exact; only
    the order of the sums differs).  Values: the per-path tolerance of the north star; in single precision a few
    paths diverge from the reference binary and each one touches a whole footprint."""
    assert got.shape == ref.shape, name
    got, ref = got.astype(np.float64), ref.astype(np.float64)
    wtol = 2e-6 if variant == "scalar_mono" else 1e-13
    wscale = np.abs(ref[..., 4]).max()
    assert np.abs(got[..., 4] - ref[..., 4]).max() <= wtol * wscale, "%s: filter weights differ" % name
    if variant == "scalar_mono_double" or not paths_may_diverge:
        tol = 1e-11 if variant == "scalar_mono_double" else 2e-5
        for c in range(ref.shape[-1]):
            scale = max(np.abs(ref[..., c]).max(), 1e-30)
            assert np.abs(got[..., c] - ref[..., c]).max() <= tol * scale, "%s: channel %d" % (name, c)
    else:
        scale = np.abs(ref[..., 1]).max()
        ok = np.abs(got[..., 1] - ref[..., 1]) <= 2e-5 * scale
        assert ok.mean() >= 0.5, "%s: %.3f of the pixels agree" % (name, ok.mean())
        mean_got, mean_ref = got[..., 1].sum() / got[..., 4].sum(), ref[..., 1].sum() / ref[..., 4].sum()
        assert abs(mean_got - mean_ref) <= 0.03 * abs(mean_ref), name
    for c in (0, 2):   # X = Y = Z in mono: the same addends; on the device each channel is its own chain of atomic adds
        assert np.allclose(got[..., c], got[..., 1], rtol=1e-5 if variant == "scalar_mono" else 1e-12, atol=1e-7 * np.abs(got[..., 1]).max())


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_films(pkg, oracle, name):
    d, variant = make_ref_golden.scene_dict(name)
    r = oracle.render(pkg, pkg.load_dict(d, variant), pkg.abi.SEED_SCALAR_BLOCKS)
    check_film(name, r.film, REF[name], variant)


def test_tables_follow_init_discretization(pkg):
    """rfilter.cpp:9-21: 31 samples of eval() over [0, radius) and a closing zero; radius, border and footprint of the
    six plugins with their default parameters."""
    want = {"box": (0.5, 0, 1), "tent": (1.0, 1, 2), "gaussian": (2.0, 2, 4), "mitchell": (2.0, 2, 4),
            "catmullrom": (2.0, 2, 4), "lanczos": (3.0, 3, 6)}
    for dtype in (np.float32, np.float64):
        for t, (radius, border, taps) in want.items():
            rf = pkg.scene.rfilter_table({"type": t}, dtype)
            assert abs(rf["radius"] - radius) < 1e-3 and rf["border"] == border and rf["taps"] == taps, t
            assert len(rf["values"]) == 32 and rf["values"][31] == 0.0
            assert rf["scale"] == float(dtype(31) / dtype(rf["radius"]))
        v = pkg.scene.rfilter_table({"type": "tent"}, dtype)["values"]
        assert v[0] == 1.0 and np.allclose(v[:31], 1.0 - np.arange(31) / 31.0, atol=1e-6)
        g = pkg.scene.rfilter_table({"type": "gaussian", "stddev": 1.0}, dtype)
        assert g["radius"] == 4.0 and g["taps"] == 8 and g["border"] == 4
        x = 4.0 * np.arange(31) / 31.0
        assert np.allclose(g["values"][:31], np.exp(-0.5 * x * x) - np.exp(-8.0), atol=1e-6)
        m = pkg.scene.rfilter_table({"type": "mitchell"}, dtype)["values"]
        assert abs(m[0] - 8.0 / 9.0) < 1e-6 and min(m) < 0     # (6 - 2B) / 6 at B = 1/3; negative lobe
        c = pkg.scene.rfilter_table({"type": "catmullrom"}, dtype)["values"]
        assert abs(c[0] - 1.0) < 1e-6 and min(c) < 0
        la = pkg.scene.rfilter_table({"type": "lanczos", "lobes": 2}, dtype)
        assert la["radius"] == 2.0 and la["values"][0] == 1.0
    # single precision rounds what double precision computes
    a = pkg.scene.rfilter_table({"type": "gaussian"}, np.float32)["values"]
    b = pkg.scene.rfilter_table({"type": "gaussian"}, np.float64)["values"]
    assert np.allclose(a, b, atol=1e-6) and all(float(np.float32(x)) == x for x in a)


def test_descriptor_and_refusals(pkg):
    d = scenes.c2(spp=1, width=8)
    d["sensor"]["film"] = {"type": "hdrfilm", "width": 8, "height": 4, "rfilter": {"type": "gaussian"}}
    sn = pkg.load_dict(d).desc.sensor
    assert sn.rfilter_radius == 2.0 and sn.rfilter_taps == 4 and sn.rfilter_border == 2 and sn.rfilter_scale == 15.5
    assert sn.rfilter_values[31] == 0.0 and 0.99 < sn.rfilter_values[0] < 1.0
    d["sensor"]["film"]["rfilter"] = {"type": "box"}
    sn = pkg.load_dict(d).desc.sensor
    assert 0.5 < sn.rfilter_radius < 0.5001 and sn.rfilter_border == 0
    for rf, msg in (({"type": "gaussian", "stddev": 18.5}, "radius of 74 pixels"),
                    ({"type": "sinc"}, "reconstruction filter plugin 'sinc'"),
                    ({"type": "tent", "radius": 2.0}, "Unreferenced property 'radius'"),
                    ({"type": "lanczos", "lobes": 9}, "radius of 9 pixels")):
        d["sensor"]["film"]["rfilter"] = rf
        with pytest.raises(RuntimeError, match=msg):
            pkg.load_dict(d)


def test_weights_sum_to_the_footprint(pkg, oracle):
    """A property that needs no fixture.  The tent table is read at floor(31 |x|), so the two taps of an axis, at
    distances x and 1 - x, add up to 1 or 1 + 1/31: away from the border of the film every sample spreads a total
    weight between 1 and (32/31)^2, and alpha follows the weights wherever every ray is valid."""
    variant = "scalar_mono_double"
    d = scenes.c2(spp=2, variant=variant, block_size=8)
    d["sensor"]["film"] = {"type": "hdrfilm", "width": 20, "height": 12, "rfilter": {"type": "tent"}}
    full = oracle.render(pkg, pkg.load_dict(d, variant), pkg.abi.SEED_WAVEFRONT).film
    per_sample = full[..., 4].sum() / (20 * 12 * 2)
    assert 0.9 < per_sample <= (32.0 / 31.0) ** 2          # footprints cut by the border of the film lose weight
    inner = full[2:-2, 2:-2, 4].mean() / 2                 # two samples per pixel
    assert 0.95 <= inner <= 1.12                          # an average over pixels, not a bound per pixel
    assert np.allclose(full[..., 3], full[..., 4], rtol=1e-13)
    # the same streams in the reference's block seeding, blocks of 8 with a border of one pixel: the same film
    blocks = oracle.render(pkg, pkg.load_dict(d, variant), pkg.abi.SEED_SCALAR_BLOCKS).film
    assert abs(blocks[..., 4].sum() / full[..., 4].sum() - 1) < 0.05


@pytest.mark.parametrize("name", CASES)
def test_plugin_describes_the_filter_of_the_reference_film(pkg, oracle, name):
    """The reference's own film and filter objects, read by the plugin through ReconstructionFilter's public interface
    (volpath_b200.cpp: radius(), border_size(), eval_discretized()), rendered by the oracle: the reference's film.  In
    double precision the Python host describes the very same table (same bits)."""
    from test_reference_plugin import ref_xml_or_skip
    ref_xml = ref_xml_or_skip()
    d, variant = make_ref_golden.scene_dict(name)
    film, _ = ref_xml.render(ref_xml.with_plugin(d), variant, write_vol=pkg.scene.write_vol, describe_oracle=True)
    check_film(name, film, REF[name], variant)
    if variant == "scalar_mono_double":
        r = oracle.render(pkg, pkg.load_dict(d, variant), pkg.abi.SEED_SCALAR_BLOCKS)
        assert np.array_equal(r.film, film), name


# ---------------------------------------------------------------- GPU parity ----------------------------------------
def gpu_compare(pkg, oracle, sc, mode, **kw):
    got = pkg.render(sc, device=0, seed_mode=mode, want_paths=True, **kw)
    ref = oracle.render(pkg, sc, mode, want_paths=True)
    assert got.totals.n_paths == ref.totals.n_paths
    assert np.array_equal(got.state, ref.state), "PCG32 end states differ"
    assert np.array_equal(got.counters, ref.counters), "event counters differ"
    assert np.array_equal(got.radiance.view(np.uint8), ref.radiance.view(np.uint8)), "per-path radiance differs"
    # film: the same addends (weight products included) in another order
    tol = 2e-5 if got.film.dtype == np.float32 else 1e-12
    for c in range(got.film.shape[-1]):
        scale = max(np.abs(ref.film[..., c]).max(), 1e-30)
        assert np.abs(got.film[..., c].astype(np.float64) - ref.film[..., c]).max() <= tol * scale, "channel %d" % c
    return got


@pytest.mark.gpu
@pytest.mark.parametrize("kernel", ["queue", "generic", "lanes", "mega"])
@pytest.mark.parametrize("name", CASES)
def test_cuda_matches_oracle(pkg, oracle, name, kernel, monkeypatch):
    """Every device kernel that splats (the product queue kernel with its presets and its generic instance, the
    warp-voting wavefront, the one-thread-per-stream twin) against the oracle, in the reference's block seeding
    (blocks with borders) and in wavefront seeding (one block over the whole crop window)."""
    monkeypatch.setitem(pkg.render_defaults, "kernel", kernel)
    d, variant = make_ref_golden.scene_dict(name)
    sc = pkg.load_dict(d, variant)
    got = gpu_compare(pkg, oracle, sc, pkg.abi.SEED_SCALAR_BLOCKS)
    check_film(name, got.film, REF[name], variant)          # and straight against the reference's film
    if kernel in ("queue", "mega"):
        gpu_compare(pkg, oracle, sc, pkg.abi.SEED_WAVEFRONT)


@pytest.mark.gpu
def test_cuda_large_film_in_global_memory(pkg, oracle):
    """A film too large for shared memory is accumulated in HBM: same footprints."""
    for variant in ("scalar_mono", "scalar_mono_double"):
        d = scenes.c2(spp=1, variant=variant, block_size=32)
        d["sensor"]["film"] = {"type": "hdrfilm", "width": 96, "height": 80, "rfilter": {"type": "gaussian"}}
        gpu_compare(pkg, oracle, pkg.load_dict(d, variant), pkg.abi.SEED_SCALAR_BLOCKS)
        d["integrator"] = {"type": "moment", "nested": d["integrator"]}
        d["sensor"]["film"].update(crop_offset_x=5, crop_offset_y=7, crop_width=70, crop_height=60,
                                   rfilter={"type": "lanczos"})
        gpu_compare(pkg, oracle, pkg.load_dict(d, variant), pkg.abi.SEED_WAVEFRONT)


@pytest.mark.gpu
def test_cuda_fast_math_keeps_the_weights(pkg, oracle):
    """math=fast changes path arithmetic, not the filter table or the footprints: W agrees with the exact film."""
    d = scenes.c2(spp=4, block_size=8)
    d["sensor"]["film"] = {"type": "hdrfilm", "width": 24, "height": 10, "rfilter": {"type": "mitchell"}}
    sc = pkg.load_dict(d)
    exact = pkg.render(sc, device=0, seed_mode=pkg.abi.SEED_WAVEFRONT)
    fast = pkg.render(sc, device=0, seed_mode=pkg.abi.SEED_WAVEFRONT, math="fast")
    assert np.allclose(fast.film[..., 4], exact.film[..., 4], rtol=1e-5, atol=1e-5)
    assert abs(fast.film[..., 1].sum() / exact.film[..., 1].sum() - 1) < 0.05


@pytest.mark.gpu
def test_descriptor_is_validated(pkg):
    d = scenes.c2(spp=1, width=8)
    d["sensor"]["film"] = {"type": "hdrfilm", "width": 8, "height": 4, "rfilter": {"type": "gaussian"}}
    sc = pkg.load_dict(d)
    sc.desc.sensor.rfilter_taps = 40
    with pytest.raises(pkg.B2vpError, match="reconstruction filter"):
        pkg.render(sc, device=0)
    sc = pkg.load_dict(d)
    sc.desc.sensor.rfilter_values[3] = float("nan")
    with pytest.raises(pkg.B2vpError, match="not finite"):
        pkg.render(sc, device=0)


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_reference_binary_renders_wide_filters_through_the_plugin(pkg, name):
    """mitsuba's own scene loader, film and filter plugins; the integrator is the product: the film the reference's
    volpath develops, with film->put(block) on a border-less block (volpath_b200.cpp render())."""
    from test_reference_plugin import ref_xml_or_skip
    ref_xml = ref_xml_or_skip()
    d, variant = make_ref_golden.scene_dict(name)
    film, _ = ref_xml.render(ref_xml.with_plugin(d, math="exact"), variant, write_vol=pkg.scene.write_vol)
    check_film(name, film, REF[name], variant)
