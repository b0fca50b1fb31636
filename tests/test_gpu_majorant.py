"""B2vpRenderArgs::majorant_grid: opt-in free-flight sampling against a grid of LOCAL majorants.

The reference has one global majorant per medium (src/media/heterogeneous.cpp:29,38), so this mode has no
per-path counterpart there: tentative collisions are a different, sparser random sequence with the same distribution of
real collisions.  It is therefore never part of the bit-exact suite; it is accepted the way SURVEY section 7 prescribes:
statistically against the GLOBAL-majorant estimator -- a two-sample z per pixel from `moment` films (mean z: no bias;
rms z: the right spread; max |z|: no outlier), with a control pair (global vs global, other streams) put to the same test
-- and by what it is for: far fewer tentative collisions per path.  Where it does not apply (homogeneous media, wrapped grids) the flag must change nothing."""
import numpy as np
import pytest

import scenes

pytestmark = pytest.mark.gpu


def _moment(d):
    d["integrator"] = {"type": "moment", "nested": d["integrator"]}
    return d


def _c3(**kw):
    d = scenes.c3(**kw)
    d["sensor"]["ray_target"] = [0.4, 0.55, 1.0]   # off the edge shared by the cube's top triangles (see test_gpu_fastmath)
    return d


CASES = {
    # name: (scene(spp, variant, seed), spp, minimum reduction of tentative collisions per path -- measured 3.3-3.5x,
    # 8x, 1.3x, 2.2x; the count includes the VIRTUAL collisions that bound the cell walk per visit, vp_scene.cuh)
    "c2": (lambda spp, v: scenes.c2(spp=spp, width=64, variant=v), 2048, 2.5),
    "c2_tau2": (lambda spp, v: scenes.c2(spp=spp, width=64, tau=2.0, variant=v), 1024, 5.0),
    "c2_notspectral": (lambda spp, v: scenes.c2(spp=spp, width=64, spectral=False, variant=v), 2048, 2.5),
    # the cloud field gains less: 41-64 % of the coarse cells touch a cloud, and inside one the local majorant IS the global one
    "c3_small": (lambda spp, v: _c3(spp=spp, n=32, size=12, variant=v), 2048, 1.2),
    "c3_cells": (lambda spp, v: _c3(spp=spp, n=96, size=12, variant=v), 1024, 1.8),
}


def pixel_stats(film):
    """Per-pixel mean, variance and sample count of the nested integrator's value from a `moment` film
    (bitmap_extract of src/librender/tests/test_renders.py:55-60)."""
    f = film.astype(np.float64)
    n = f[..., 4]
    mean = f[..., 6] / n
    return mean, np.maximum(f[..., 9] / n - mean * mean, 0), n


def two_sample_z(a, b):
    """z per pixel between two INDEPENDENT estimates, each with its own variance.  The reference's z-test
    (test_renders.py:62-83) takes the variance from a converged reference image only; on the cloud field a pixel's
    samples are mostly ~0 with rare bright ones, its variance estimate at a few thousand samples is itself noise, and
    even two renders of the SAME estimator fail that test (measured; tools/mg_bias_check.py)."""
    ma, va, na = pixel_stats(a)
    mb, vb, nb = pixel_stats(b)
    se = np.sqrt(va / na + vb / nb)
    ok = se > 0
    return (ma - mb)[ok] / se[ok]


def assert_same_distribution(a, b, what):
    """z ~ N(0, 1) per pixel: no bias (mean z), right spread (rms z), no outlier (max |z|; P(|z| > 4.5) = 7e-6)."""
    z = two_sample_z(a, b)
    n = z.size
    assert n > 30
    assert abs(z.mean()) * np.sqrt(n) < 4.0, "%s: mean z %.3f over %d pixels: biased" % (what, z.mean(), n)
    assert 0.75 < np.sqrt((z ** 2).mean()) < 1.3, "%s: rms z %.3f" % (what, np.sqrt((z ** 2).mean()))
    assert np.abs(z).max() < 4.5, "%s: max |z| %.2f" % (what, np.abs(z).max())
    return z


@pytest.mark.parametrize("math", ["exact", "fast"])
@pytest.mark.parametrize("variant", ["scalar_mono", "scalar_mono_double"])
@pytest.mark.parametrize("name", sorted(CASES))
def test_majorant_grid_is_statistically_the_reference_estimator(pkg, name, variant, math):
    if math == "fast" and variant == "scalar_mono_double":
        pytest.skip("B2VP_MATH_FAST is an fp32 mode")
    abi = pkg.abi
    make, spp, min_gain = CASES[name]

    def scene(seed):
        d = _moment(make(spp, variant))
        d["sensor"]["sampler"]["seed"] = seed
        return pkg.load_dict(d, variant)
    other = pkg.render(scene(1), seed_mode=abi.SEED_WAVEFRONT)             # global majorant, other random streams
    with pkg.DeviceScene(scene(0)) as ds:
        base = ds.render(seed_mode=abi.SEED_WAVEFRONT)                     # global majorant: the reference's estimator
        got = ds.render(seed_mode=abi.SEED_WAVEFRONT, majorant_grid=1, math=math)
        assert "majorant grid" in ds.info().last_kernel.decode()
        again = ds.render(seed_mode=abi.SEED_WAVEFRONT, majorant_grid=1, math=math)
    assert got.totals.n_paths == base.totals.n_paths and got.totals.n_invalid_samples == 0
    assert np.array_equal(got.film[..., 4], base.film[..., 4])             # sample bookkeeping is exact
    assert_same_distribution(base.film, other.film, "control (global vs global)")   # the test itself is sound
    z = assert_same_distribution(got.film, other.film, "majorant grid vs global")
    ref = base
    gain = ref.totals.n_medium_steps / got.totals.n_medium_steps
    m_ref, m_got = ref.film[..., 6].sum() / ref.film[..., 4].sum(), got.film[..., 6].sum() / got.film[..., 4].sum()
    print("\nmajorant grid %-15s %-18s %-5s tentative collisions per path %.1f -> %.1f (x%.1f), image mean %.6g vs %.6g, "
          "z: mean %.3f rms %.3f max %.2f" % (
        name, variant, math, ref.totals.n_medium_steps / ref.totals.n_paths, got.totals.n_medium_steps / got.totals.n_paths,
        gain, m_got, m_ref, z.mean(), np.sqrt((z ** 2).mean()), np.abs(z).max()))
    assert gain >= min_gain, "only %.2fx fewer tentative collisions" % gain
    # real scattering events have the same distribution under any majorant: their count per path agrees statistically
    r_ref, r_got = ref.totals.n_real / ref.totals.n_paths, got.totals.n_real / got.totals.n_paths
    assert abs(r_got - r_ref) <= 0.02 * r_ref + 5 * np.sqrt(r_ref / ref.totals.n_paths), (r_got, r_ref)
    # deterministic: the same launch twice gives the same counters
    for k in ("n_rng_draws", "n_medium_steps", "n_null", "n_real", "n_intersect", "n_surface"):
        assert getattr(again.totals, k) == getattr(got.totals, k), k


def test_majorant_grid_changes_nothing_where_it_does_not_apply(pkg):
    """Homogeneous media and wrapped (repeat) grids keep the global majorant: same bits with and without the flag."""
    abi = pkg.abi
    d = scenes.c2(spp=8, width=32)
    d["atm"]["sigma_t"]["wrap_mode"] = "repeat"
    for dd in (scenes.c1(spp=256), scenes.c5(spp=4, n_leaves=200, n_sensors=64), d):
        sc = pkg.load_dict(dd)
        with pkg.DeviceScene(sc) as ds:
            a = ds.render(seed_mode=abi.SEED_WAVEFRONT, want_paths=True)
            b = ds.render(seed_mode=abi.SEED_WAVEFRONT, want_paths=True, majorant_grid=1)
            assert "majorant grid" not in ds.info().last_kernel.decode()
        assert np.array_equal(a.state, b.state) and np.array_equal(a.counters, b.counters)
        assert np.array_equal(a.radiance.view(np.uint8), b.radiance.view(np.uint8))


def test_majorant_grid_follows_updates(pkg):
    """b2vp_scene_update (medium scale) and b2vp_scene_update_grid (new values): the majorant grid follows -- same
    counters and film as a fresh scene built from the new values."""
    abi = pkg.abi
    d = _c3(spp=8, n=32, size=16)
    rng = np.random.default_rng(5)
    new_sigma = (d["atm"]["sigma_t"]["data"] * rng.uniform(0.5, 2.0, size=d["atm"]["sigma_t"]["data"].shape)).astype(np.float32)
    sc = pkg.load_dict(d)
    with pkg.DeviceScene(sc) as ds:
        ds.render(seed_mode=abi.SEED_WAVEFRONT, majorant_grid=1)
        ds.update_grid(sc.desc.media[0].sigma_t, new_sigma)
        ds.update(medium=0, medium_scale=0.5)
        got = ds.render(seed_mode=abi.SEED_WAVEFRONT, majorant_grid=1, want_paths=True)
    d2 = _c3(spp=8, n=32, size=16)
    d2["atm"]["sigma_t"]["data"] = new_sigma
    d2["atm"]["scale"] = 0.5
    fresh = pkg.render(pkg.load_dict(d2), seed_mode=abi.SEED_WAVEFRONT, majorant_grid=1, want_paths=True)
    assert np.array_equal(got.state, fresh.state) and np.array_equal(got.counters, fresh.counters)
    assert np.array_equal(got.radiance.view(np.uint8), fresh.radiance.view(np.uint8))


def test_majorant_grid_resolution_request(pkg):
    """majorant_grid = n > 1 asks for n cells along the longest axis.  Every resolution beats the global majorant; finer
    is not always fewer, because a walk is cut into visits of 8 cells by virtual collisions (64 layers: more of them)."""
    abi = pkg.abi
    sc = pkg.load_dict(scenes.c2(spp=64, width=64))
    steps = {}
    with pkg.DeviceScene(sc) as ds:
        for n in (0, 4, 16, 64):
            r = ds.render(seed_mode=abi.SEED_WAVEFRONT, majorant_grid=n)
            assert r.totals.n_paths == 64 * 64
            steps[n] = r.totals.n_medium_steps
    assert steps[0] > 2 * steps[4] and steps[0] > 2 * steps[16] and steps[0] > steps[64], steps
    with pytest.raises(pkg.B2vpError, match="queue kernel"):
        pkg.render(sc, majorant_grid=1, kernel="mega")
