"""GPU parity proper: the CUDA path (through the C-ABI) against the CPU oracle on identical scenes and seeds.
Integer results (PCG32 end state, draw and event counters) and per-path radiance must be bit-exact in both
precisions; films are compared with a summation-order tolerance."""
import numpy as np
import pytest

import scenes

pytestmark = pytest.mark.gpu

CASES = {
    "c1": lambda v: scenes.c1(spp=4096, variant=v),
    "c1_sun30": lambda v: scenes.c1(spp=2048, variant=v, sun=(0.5, 0, -0.8660254)),
    "c1_depth3": lambda v: scenes.c1(spp=2048, variant=v, max_depth=3, rr_depth=2),
    "c2": lambda v: scenes.c2(spp=16, variant=v),
    "c2_nearest_tau2": lambda v: scenes.c2(spp=8, variant=v, tau=2.0, filter_type="nearest"),
    "c2_notspectral": lambda v: scenes.c2(spp=8, variant=v, spectral=False),
    "c2_mdistant": lambda v: scenes.c2(spp=8, variant=v, sensor="mdistant", width=64),
    "c3_small": lambda v: scenes.c3(spp=2, variant=v, n=32, size=16),
    "c5_small": lambda v: scenes.c5(spp=8, variant=v, n_leaves=400, n_sensors=64),
    "c2_tabphase": lambda v: _tabphase(v),
    "c2_sky_volpathmis": lambda v: _mis(_sky(scenes.c2(spp=8, variant=v, width=64), 0.25)),
    "c2_sky": lambda v: _sky(scenes.c2(spp=8, variant=v, width=64), 0.25),
    "c5_sky_only": lambda v: _sky(scenes.c5(spp=4, variant=v, n_leaves=200, n_sensors=64), 0.5, sun=False),
    "c2_volpathmis": lambda v: _mis(scenes.c2(spp=8, variant=v, width=64)),
    "c3_volpathmis": lambda v: _mis(scenes.c3(spp=2, variant=v, n=16, size=12)),
    "c5_volpathmis": lambda v: _mis(scenes.c5(spp=4, variant=v, n_leaves=200, n_sensors=64)),
    # volpathmis on the queue kernel: the other implementation (use_spectral_mis = false, volpathmis.cpp:35-52), media
    # without spectral free-flight weights, a blend phase (its pdf enters the weights as pdf / pdf), a depth limit
    "c2_volpathmis_nospectralmis": lambda v: _mis(scenes.c2(spp=8, variant=v, width=64), use_spectral_mis=False),
    "c2_volpathmis_notspectral": lambda v: _mis(scenes.c2(spp=8, variant=v, width=64, spectral=False)),
    "c3_volpathmis_nospectralmis": lambda v: _mis(scenes.c3(spp=2, variant=v, n=16, size=12), use_spectral_mis=False),
    "c2_volpathmis_blendphase": lambda v: _mis(_blendphase(v, True)),
    "c1_volpathmis_depth3": lambda v: _mis(scenes.c1(spp=2048, variant=v, max_depth=3, rr_depth=2)),
    # two suns: the emitter pdf is 1/2, volpathmis works with weight * pdf (volpathmis.cpp:341-343)
    "c2_two_suns": lambda v: _two_suns(scenes.c2(spp=8, variant=v, width=64)),
    "c2_two_suns_volpathmis": lambda v: _mis(_two_suns(scenes.c2(spp=8, variant=v, width=64))),
    "c5_two_suns_volpathmis": lambda v: _mis(_two_suns(scenes.c5(spp=4, variant=v, n_leaves=200, n_sensors=64))),
    # environment emitter on the queue kernel (FT_ENV): with and without a sun, hidden, both integrators, surfaces
    "c2_sky_hidden": lambda v: _hide(_sky(scenes.c2(spp=8, variant=v, width=64), 0.25)),
    "c2_sky_hidden_volpathmis": lambda v: _mis(_hide(_sky(scenes.c2(spp=8, variant=v, width=64), 0.25))),
    "c2_sky_only_volpathmis": lambda v: _mis(_sky(scenes.c2(spp=8, variant=v, width=64), 0.5, sun=False)),
    "c1_sky_depth3": lambda v: _sky(scenes.c1(spp=2048, variant=v, max_depth=3, rr_depth=2), 0.25),
    "c3_sky": lambda v: _sky(scenes.c3(spp=2, variant=v, n=16, size=12), 0.25),
    "c3_sky_volpathmis": lambda v: _mis(_sky(scenes.c3(spp=2, variant=v, n=16, size=12), 0.25)),
    "c5_sky": lambda v: _sky(scenes.c5(spp=4, variant=v, n_leaves=200, n_sensors=64), 0.5),
    "c5_sky_volpathmis": lambda v: _mis(_sky(scenes.c5(spp=4, variant=v, n_leaves=200, n_sensors=64), 0.5)),
    "c5_sky_only_volpathmis_nospectralmis": lambda v: _mis(_sky(scenes.c5(spp=4, variant=v, n_leaves=200, n_sensors=64), 0.5,
                                                                sun=False), use_spectral_mis=False),
    "c5_sky_two_suns_bilambertian": lambda v: _two_suns(_sky(scenes.c5(
        spp=8, variant=v, n_leaves=400, n_sensors=64,
        leaf_bsdf={"type": "bilambertian", "reflectance": 0.5, "transmittance": 0.25}), 0.5)),
    "c5_volpathmis_bilambertian": lambda v: _mis(scenes.c5(
        spp=8, variant=v, n_leaves=400, n_sensors=64,
        leaf_bsdf={"type": "bilambertian", "reflectance": 0.5, "transmittance": 0.25})),
    "c2_irradiancemeter": lambda v: _irradiancemeter(v),
    "c2_distantflux": lambda v: _distantflux(v, True),
    "c2_distantflux_notarget": lambda v: _distantflux(v, False),
    "c2_blendphase": lambda v: _blendphase(v, True),
    "c2_blendphase_const": lambda v: _blendphase(v, False),
    "c5_bilambertian": lambda v: scenes.c5(spp=8, variant=v, n_leaves=400, n_sensors=64,
                                           leaf_bsdf={"type": "bilambertian", "reflectance": 0.5, "transmittance": 0.25}),
    # sensors that target a shape: a rectangle of the scene, or one nested in the sensor (sampled, never intersected)
    "c2_target_shape": lambda v: _golden()._target_shape(scenes.c2(spp=8, variant=v, width=64)),
    "c3_target_shape": lambda v: _golden()._target_shape(scenes.c3(spp=2, variant=v, n=16, size=12), shape="top"),
    "c2_nested_target": lambda v: _golden()._nested_target(scenes.c2(spp=8, variant=v, width=64)),
    "c2_mdistant_nested_target": lambda v: _golden()._nested_target(scenes.c2(spp=8, variant=v, width=64, sensor="mdistant"), "target"),
    "c2_distantflux_target_shape": lambda v: _golden()._target_shape(_distantflux(v, False), "target"),
    # ray origins on a shape (distant.cpp:366-375): all rays start on it; a small one that slanted rays miss (weight 0);
    # a null rectangle of the scene itself; distantflux
    "c2_origin_shape": lambda v: _golden()._origin_shape(scenes.c2(spp=8, variant=v, width=64)),
    "c2_origin_shape_miss": lambda v: _golden()._origin_shape(scenes.c2(spp=8, variant=v, width=64), half=1.0, sweep=True),
    "c3_origin_ref": lambda v: _golden()._target_shape(scenes.c3(spp=2, variant=v, n=16, size=12), "ray_origin", "top"),
    "c2_distantflux_origin": lambda v: _golden()._origin_shape(_distantflux(v, True), "origin", half=2.0),
}


def _golden():
    import make_ref_golden
    return make_ref_golden


def _tabphase(variant):
    import make_ref_golden
    d = scenes.c2(spp=8, variant=variant, width=64)
    d["atm"]["phase"] = {"type": "tabphase", "values": make_ref_golden.TAB_AEROSOL}
    return d


def _blendphase(variant, grid):
    import make_ref_golden
    return make_ref_golden._blendphase(grid, spp=8, variant=variant, width=64)


def _distantflux(variant, target):
    import make_ref_golden
    return make_ref_golden._distantflux(target, spp=4, variant=variant)


def _irradiancemeter(variant):
    import make_ref_golden
    return make_ref_golden._irradiancemeter(2048, variant)


def _sky(d, radiance, sun=True):
    d["sky"] = {"type": "constant", "radiance": radiance}
    if not sun:
        del d["sun"]
    return d


def _two_suns(d):
    d["sun2"] = {"type": "directional", "direction": [-0.4, 0.1, -1], "irradiance": 0.5}
    return d


def _hide(d):
    d["integrator"] = dict(d["integrator"], hide_emitters=True)
    return d


def _mis(d, **props):
    d["integrator"] = dict(d["integrator"], type="volpathmis", **props)
    return d


def compare(pkg, oracle, sc, mode, **kw):
    got = pkg.render(sc, device=0, seed_mode=mode, want_paths=True, **kw)
    ref = oracle.render(pkg, sc, mode, want_paths=True, **kw)
    assert got.totals.n_paths == ref.totals.n_paths
    assert np.array_equal(got.state, ref.state), "PCG32 end states differ"
    assert np.array_equal(got.counters, ref.counters), "event counters differ"
    assert np.array_equal(got.radiance.view(np.uint8), ref.radiance.view(np.uint8)), "per-path radiance differs"
    for k in ("n_rng_draws", "n_medium_steps", "n_null", "n_real", "n_intersect", "n_surface", "n_invalid_samples"):
        assert getattr(got.totals, k) == getattr(ref.totals, k), k
    # film: same addends, different summation order
    tol = 2e-4 if got.film.dtype == np.float32 else 1e-12
    assert np.allclose(got.film, ref.film, rtol=tol, atol=tol * 1e-2)
    assert got.totals.n_launches >= 1 and got.totals.gpu_ms > 0
    return got


@pytest.fixture(params=["queue", "generic", "q256", "lanes", "mega"])
def kernel(request, monkeypatch, pkg):
    """B2vpRenderArgs::kernel selects the device kernel: the queue-based wavefront product kernel (two CTA sizes), the
    register-resident warp-voting wavefront or its one-thread-per-path twin; "generic" forces the product
    kernel's all-features instance instead of the scene-class preset.  All must give identical results."""
    monkeypatch.setitem(pkg.render_defaults, "kernel", request.param)
    return request.param


@pytest.mark.parametrize("variant", ["scalar_mono", "scalar_mono_double"])
@pytest.mark.parametrize("case", sorted(CASES))
def test_wavefront_seeding_bit_exact(pkg, oracle, case, variant, kernel):
    if kernel in ("q256", "lanes", "generic") and variant == "scalar_mono_double":
        pytest.skip("variant covered in fp32")
    sc = pkg.load_dict(CASES[case](variant), variant)
    compare(pkg, oracle, sc, pkg.abi.SEED_WAVEFRONT)


@pytest.mark.parametrize("variant", ["scalar_mono", "scalar_mono_double"])
def test_volpathmis_runs_on_the_queue_kernel(pkg, oracle, variant):
    """volpathmis is a set of phase bodies of the product kernel (FT_MIS instances), not a job for the slow
    one-thread-per-stream kernel: B2vpSceneInfo::last_kernel says which instance a launch took."""
    abi = pkg.abi
    for make, preset in ((lambda: scenes.c1(spp=256, variant=variant), "homogeneous"),
                         (lambda: scenes.c2(spp=4, variant=variant, width=32), "heterogeneous"),
                         (lambda: scenes.c3(spp=1, variant=variant, n=48, size=8), "heterogeneous+albedo grid"),
                         (lambda: _sky(scenes.c2(spp=4, variant=variant, width=32), 0.25), "environment, generic"),
                         (lambda: scenes.c5(spp=2, variant=variant, n_leaves=100, n_sensors=16),   # BVH in HBM
                          "homogeneous, large mesh" if variant == "scalar_mono" else "generic")):
        sc = pkg.load_dict(_mis(make()), variant)
        with pkg.DeviceScene(sc) as ds:
            ds.render(seed_mode=abi.SEED_WAVEFRONT)
            assert ds.info().last_kernel.decode().startswith("queue/volpathmis, " + preset), ds.info().last_kernel.decode()
            ds.render(seed_mode=abi.SEED_WAVEFRONT, kernel="generic")
            assert "generic" in ds.info().last_kernel.decode()
            ds.render(seed_mode=abi.SEED_WAVEFRONT, kernel="mega")
            assert ds.info().last_kernel.decode().startswith("stream")
            with pytest.raises(pkg.B2vpError, match="volpathmis"):
                ds.render(seed_mode=abi.SEED_WAVEFRONT, majorant_grid=1)
            if variant == "scalar_mono":
                ds.render(seed_mode=abi.SEED_WAVEFRONT, math="fast")
                assert ds.info().last_kernel.decode().endswith("fast f32")


@pytest.mark.parametrize("variant", ["scalar_mono", "scalar_mono_double"])
def test_aerosol_and_flux_scenes_take_their_preset(pkg, oracle, variant):
    """Atmospheres with particles (tabphase, blendphase with a grid or constant weight) and the flux sensors have a
    preset of their own (PRE_HET_AER, both integrators) instead of the all-features instance: same bits, 1.4x faster."""
    abi = pkg.abi
    for make in (lambda: _tabphase(variant), lambda: _blendphase(variant, True), lambda: _blendphase(variant, False),
                 lambda: _distantflux(variant, True), lambda: _irradiancemeter(variant)):
        for integ, name in (("volpath", "queue/heterogeneous, aerosols"), ("volpathmis", "queue/volpathmis, heterogeneous, aerosols")):
            d = make()
            d["integrator"] = dict(d["integrator"], type=integ)
            sc = pkg.load_dict(d, variant)
            with pkg.DeviceScene(sc) as ds:
                got = ds.render(seed_mode=abi.SEED_WAVEFRONT, want_paths=True)
                assert ds.info().last_kernel.decode().startswith(name), ds.info().last_kernel.decode()
                generic = ds.render(seed_mode=abi.SEED_WAVEFRONT, want_paths=True, kernel="generic")
                assert "generic" in ds.info().last_kernel.decode()
            assert np.array_equal(got.state, generic.state) and np.array_equal(got.counters, generic.counters)
            assert np.array_equal(got.radiance.view(np.uint8), generic.radiance.view(np.uint8))


def _canopy_under_atmosphere(variant, aerosols):
    """The C5 canopy under a heterogeneous 1D atmosphere (the C2 column over the slab), optionally with an aerosol layer
    mixed in by a blendphase: the shape of Eradiate's surface-atmosphere scenes."""
    import make_ref_golden
    T = scenes.T
    d = scenes.c5(spp=4, variant=variant, n_leaves=150, n_sensors=16)
    grid = scenes.sigma_t_1d(64)
    tw = T.translate([-1.5, -1.5, 0.2], np.float32) @ T.scale([3.0, 3.0, 0.8], np.float32)
    d["atm"] = {"type": "heterogeneous", "sigma_t": {"type": "gridvolume", "data": grid, "to_world": tw},
                "scale": 0.3 / float(grid.mean()) / 0.8, "albedo": 1.0, "phase": {"type": "rayleigh"}}
    if aerosols:
        w = np.clip(np.linspace(1.0, -0.5, 16), 0, 1).astype(np.float32).reshape(16, 1, 1)
        d["atm"]["phase"] = {"type": "blendphase", "a_molecules": {"type": "rayleigh"},
                             "b_aerosols": {"type": "tabphase", "values": make_ref_golden.TAB_AEROSOL},
                             "weight": {"type": "gridvolume", "data": w, "to_world": tw}}
    return d


@pytest.mark.parametrize("variant", ["scalar_mono", "scalar_mono_double"])
@pytest.mark.parametrize("aerosols", [False, True])
def test_canopy_under_a_heterogeneous_atmosphere(pkg, oracle, variant, aerosols):
    """Preset PRE_HET_BIG (a mesh whose BVH is read from HBM under a heterogeneous atmosphere with or without
    particles; 768-thread CTAs in fp32): the oracle's bits, the generic instance's bits, both integrators."""
    abi = pkg.abi
    for integ in ("volpath", "volpathmis"):
        d = _canopy_under_atmosphere(variant, aerosols)
        d["integrator"] = dict(d["integrator"], type=integ)
        sc = pkg.load_dict(d, variant)
        ref = oracle.render(pkg, sc, abi.SEED_WAVEFRONT, want_paths=True)
        with pkg.DeviceScene(sc) as ds:
            got = ds.render(seed_mode=abi.SEED_WAVEFRONT, want_paths=True)
            name = ds.info().last_kernel.decode()
            assert "heterogeneous, large mesh" in name or "heterogeneous, aerosols" in name, name
            generic = ds.render(seed_mode=abi.SEED_WAVEFRONT, want_paths=True, kernel="generic")
        for other in (generic, ref):
            assert np.array_equal(got.state, other.state) and np.array_equal(got.counters, other.counters)
            assert np.array_equal(got.radiance.view(np.uint8), other.radiance.view(np.uint8))


@pytest.mark.parametrize("variant", ["scalar_mono", "scalar_mono_double"])
def test_scalar_block_seeding_bit_exact(pkg, oracle, variant):
    """Reference scalar-mode seeding: one stream per pixel, spp drawn serially (integrator.cpp:197-208)."""
    sc = pkg.load_dict(scenes.c2(spp=32, variant=variant, width=48, block_size=16), variant)
    compare(pkg, oracle, sc, pkg.abi.SEED_SCALAR_BLOCKS)
    sc = pkg.load_dict(scenes.c3(spp=2, variant=variant, n=16, size=20, block_size=8, samples_per_pass=1), variant)
    compare(pkg, oracle, sc, pkg.abi.SEED_SCALAR_BLOCKS)


@pytest.mark.parametrize("variant", ["scalar_mono", "scalar_mono_double"])
def test_moment_integrator(pkg, oracle, variant, kernel):
    """volpath nested in `moment` (src/integrators/moment.cpp): 11 film channels, no validity check on put."""
    if kernel in ("q256", "generic") and variant == "scalar_mono_double":
        pytest.skip("variant covered in fp32")
    d = scenes.c2(spp=8, variant=variant, width=48)
    d["integrator"] = {"type": "moment", "nested": d["integrator"]}
    sc = pkg.load_dict(d, variant)
    assert sc.film_channels == 11
    got = compare(pkg, oracle, sc, pkg.abi.SEED_WAVEFRONT)
    assert got.film.shape == (1, 48, 11) and (got.film[..., 8] > 0).any()
    d = scenes.c3(spp=2, variant=variant, n=16, size=72)   # 72 x 72 x 11 film: accumulated in global memory
    d["integrator"] = {"type": "moment", "nested": d["integrator"]}
    compare(pkg, oracle, pkg.load_dict(d, variant), pkg.abi.SEED_SCALAR_BLOCKS)


@pytest.mark.parametrize("cells", ["1", "0"])
def test_large_grid_layouts(pkg, oracle, cells):
    """Grids too large for shared memory are read from HBM either through trilinear cell records (8 corner
    values per 32-byte record, the default) or through eight plain gathers (B2VP_OPT_NO_CELL_RECORDS): same bits."""
    options = 0 if cells == "1" else pkg.abi.OPT_NO_CELL_RECORDS
    for variant in ("scalar_mono", "scalar_mono_double"):
        sc = pkg.load_dict(scenes.c3(spp=2, variant=variant, n=40, size=12), variant)
        sc.desc.options = options
        compare(pkg, oracle, sc, pkg.abi.SEED_WAVEFRONT)
    d = scenes.c3(spp=2, n=40, size=12)
    d["atm"]["albedo"] = 0.9   # unpaired: sigma_t grid alone
    sc = pkg.load_dict(d)
    sc.desc.options = options
    compare(pkg, oracle, sc, pkg.abi.SEED_WAVEFRONT)


def test_explicit_states(pkg, oracle):
    sc = pkg.load_dict(scenes.c2(spp=4, width=32))
    rng = np.random.default_rng(3)
    n = 777
    st = rng.integers(0, 2**63, n, dtype=np.uint64)
    inc = rng.integers(0, 2**63, n, dtype=np.uint64) | np.uint64(1)
    pix = rng.integers(0, 32, n, dtype=np.uint32)
    compare(pkg, oracle, sc, pkg.abi.SEED_EXPLICIT, explicit=(st, inc, pix, 3))


def test_stream_ranges_partition(pkg, oracle):
    """Disjoint stream ranges sum to the full render (the multi-GPU sharding contract)."""
    sc = pkg.load_dict(scenes.c2(spp=16, width=32))
    with pkg.DeviceScene(sc) as ds:
        total, _ = ds.stream_count(pkg.abi.SEED_WAVEFRONT)
        full = ds.render(seed_mode=pkg.abi.SEED_WAVEFRONT, want_paths=True)
        a = ds.render(seed_mode=pkg.abi.SEED_WAVEFRONT, first_stream=0, n_streams=total // 3, want_paths=True)
        b = ds.render(seed_mode=pkg.abi.SEED_WAVEFRONT, first_stream=total // 3, n_streams=total - total // 3,
                      want_paths=True)
        assert np.array_equal(np.concatenate([a.radiance, b.radiance]), full.radiance)
        assert np.array_equal(np.concatenate([a.state, b.state]), full.state)
        assert np.allclose(a.film + b.film, full.film, rtol=2e-4, atol=1e-6)
        # accumulate into an existing film
        acc = ds.render(seed_mode=pkg.abi.SEED_WAVEFRONT, first_stream=total // 3, n_streams=total - total // 3,
                        film=a.film.copy())
        assert np.allclose(acc.film, full.film, rtol=2e-4, atol=1e-6)


@pytest.mark.parametrize("spp,ranks", [(12, 3), (32, 4)])   # blocks of 4 and of 8 streams (powers of two)
def test_interleaved_ranges_partition(pkg, oracle, kernel, spp, ranks):
    """Interleaved sharding (B2vpRenderArgs::interleave_*): each rank a share of every pixel's samples; the films add up
    to the full render, every rank's paths are bit-exact against the oracle (whose index map is the general one, with
    divisions; the device's is shift / multiply / mask, which is why blocks must be powers of two)."""
    sc = pkg.load_dict(scenes.c2(spp=spp, width=24))
    with pkg.DeviceScene(sc) as ds:
        full = ds.render(seed_mode=pkg.abi.SEED_WAVEFRONT)
        acc = np.zeros_like(full.film)
        for rank in range(ranks):
            kw = pkg.shard_interleaved(spp, 24, rank, ranks)
            got = ds.render(seed_mode=pkg.abi.SEED_WAVEFRONT, want_paths=True, **kw)
            ref = oracle.render(pkg, sc, pkg.abi.SEED_WAVEFRONT, want_paths=True, **kw)
            assert np.array_equal(got.state, ref.state) and np.array_equal(got.radiance, ref.radiance)
            assert np.array_equal(got.counters, ref.counters)
            assert (got.film[..., 4] == spp // ranks).all()
            acc += got.film
        assert np.allclose(acc, full.film, rtol=2e-4, atol=1e-6)
        # the partition b2vp_render itself uses for a multi-device render (b2vp_shard_streams), by hand on one device
        n, _ = ds.stream_count(pkg.abi.SEED_WAVEFRONT)
        acc = np.zeros_like(full.film)
        paths = 0
        for rank in range(3):
            block, count = pkg.shard_streams(n, 3, rank)
            got = ds.render(seed_mode=pkg.abi.SEED_WAVEFRONT, n_streams=count, interleave=(block, 3, rank), want_paths=True)
            ref = oracle.render(pkg, sc, pkg.abi.SEED_WAVEFRONT, n_streams=count, interleave=(block, 3, rank), want_paths=True)
            assert np.array_equal(got.state, ref.state) and np.array_equal(got.counters, ref.counters)
            acc += got.film
            paths += got.totals.n_paths
        assert paths == full.totals.n_paths and np.allclose(acc, full.film, rtol=2e-4, atol=1e-6)
        with pytest.raises(pkg.B2vpError, match="power of two"):
            ds.render(seed_mode=pkg.abi.SEED_WAVEFRONT, n_streams=6 * 24, interleave=(6, 3, 0))


def test_scene_update_matches_fresh_scene(pkg, oracle):
    d = scenes.c2(spp=8, width=16)
    sc = pkg.load_dict(d)
    with pkg.DeviceScene(sc) as ds:
        ds.render(seed_mode=pkg.abi.SEED_WAVEFRONT)
        new_scale = float(np.float32(d["atm"]["scale"] * 1.7))
        ds.update(medium=0, medium_scale=new_scale, bsdf=1, bsdf_p=(0.2, 0.7, -0.1, 0.2))
        got = ds.render(seed_mode=pkg.abi.SEED_WAVEFRONT, want_paths=True)
    d["atm"]["scale"] = new_scale
    d["ground"]["bsdf"] = {"type": "rpv", "rho_0": 0.2, "k": 0.7, "g": -0.1}
    sc2 = pkg.load_dict(d)
    assert sc2.desc.bsdfs[1].type == pkg.abi.BSDF_RPV
    ref = oracle.render(pkg, sc2, pkg.abi.SEED_WAVEFRONT, want_paths=True)
    assert np.array_equal(got.radiance, ref.radiance) and np.array_equal(got.state, ref.state)


def test_grid_update_matches_fresh_scene(pkg, oracle):
    """b2vp_scene_update_grid (the reference's traverse() -> "data" -> GridVolume::parameters_changed,
    grid3d.cpp:366-391): new grid values in place == a scene built from those values.  Covers a grid staged in shared
    memory (C2), a grid read from HBM through cell records with its paired albedo grid (C3, n = 40), both
    precisions, the majorant following the data, and keep_max."""
    for variant in ("scalar_mono", "scalar_mono_double"):
        d = scenes.c2(spp=8, width=16, variant=variant)
        g0 = np.asarray(d["atm"]["sigma_t"]["data"], np.float32)
        g1 = (g0[::-1] * np.float32(1.75) + np.float32(0.125)).copy()     # other profile, other maximum
        sc = pkg.load_dict(d, variant)
        vol = sc.desc.media[0].sigma_t
        with pkg.DeviceScene(sc) as ds:
            ds.render(seed_mode=pkg.abi.SEED_WAVEFRONT)
            ds.update_grid(vol, g1)
            got = ds.render(seed_mode=pkg.abi.SEED_WAVEFRONT, want_paths=True)
            ds.update_grid(vol, g0)                                          # and back: the first film again
            back = ds.render(seed_mode=pkg.abi.SEED_WAVEFRONT, want_paths=True)
        d["atm"]["sigma_t"]["data"] = g1
        ref = oracle.render(pkg, pkg.load_dict(d, variant), pkg.abi.SEED_WAVEFRONT, want_paths=True)
        assert np.array_equal(got.radiance, ref.radiance) and np.array_equal(got.state, ref.state), variant
        d["atm"]["sigma_t"]["data"] = g0
        ref = oracle.render(pkg, pkg.load_dict(d, variant), pkg.abi.SEED_WAVEFRONT, want_paths=True)
        assert np.array_equal(back.radiance, ref.radiance), variant

        d = scenes.c3(spp=2, variant=variant, n=40, size=12)
        s0 = np.asarray(d["atm"]["sigma_t"]["data"], np.float32)
        a0 = np.asarray(d["atm"]["albedo"]["data"], np.float32)
        s1 = (s0.transpose(2, 1, 0) * np.float32(0.5)).copy()
        a1 = (np.float32(1) - (np.float32(1) - a0) * np.float32(3)).clip(0, 1).astype(np.float32)
        sc = pkg.load_dict(d, variant)
        with pkg.DeviceScene(sc) as ds:
            ds.render(seed_mode=pkg.abi.SEED_WAVEFRONT)
            m = sc.desc.media[0]
            ds.update_grid(m.sigma_t, s1)
            ds.update_grid(m.albedo, a1)
            got = ds.render(seed_mode=pkg.abi.SEED_WAVEFRONT, want_paths=True)
        d["atm"]["sigma_t"]["data"], d["atm"]["albedo"]["data"] = s1, a1
        ref = oracle.render(pkg, pkg.load_dict(d, variant), pkg.abi.SEED_WAVEFRONT, want_paths=True)
        assert np.array_equal(got.radiance, ref.radiance) and np.array_equal(got.state, ref.state), variant

    # keep_max: the majorant of the first grid stays (the reference's fixed max_value)
    d = scenes.c2(spp=8, width=16)
    g0 = np.asarray(d["atm"]["sigma_t"]["data"], np.float32)
    g1 = (g0 * np.float32(0.5)).copy()
    sc = pkg.load_dict(d)
    with pkg.DeviceScene(sc) as ds:
        ds.update_grid(sc.desc.media[0].sigma_t, g1, keep_max=True)
        got = ds.render(seed_mode=pkg.abi.SEED_WAVEFRONT, want_paths=True)
    d["atm"]["sigma_t"]["data"] = g1
    d["atm"]["sigma_t"]["max_value"] = float(g0.max())
    ref = oracle.render(pkg, pkg.load_dict(d), pkg.abi.SEED_WAVEFRONT, want_paths=True)
    assert np.array_equal(got.radiance, ref.radiance)


def test_grid_update_errors(pkg):
    d = scenes.c2(spp=1, width=4)
    sc = pkg.load_dict(d)
    vol = sc.desc.media[0].sigma_t
    with pkg.DeviceScene(sc) as ds:
        g = np.asarray(d["atm"]["sigma_t"]["data"], np.float32)
        with pytest.raises(pkg.abi.B2vpError, match="Unsupported GridVolume data size update"):
            ds.update_grid(vol, g[:-1])
        with pytest.raises(pkg.abi.B2vpError, match="constant volume"):
            ds.update_grid(1 - vol, g)
        with pytest.raises(pkg.abi.B2vpError, match="out of range"):
            ds.update_grid(7, g)


def test_cuda_matches_golden_fixtures(pkg):
    """The CUDA path against the committed golden vectors (no oracle involved)."""
    import json
    import os
    import make_golden
    gold = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden.json")))
    for name, g in gold.items():
        sc = pkg.load_dict(make_golden.CASES[name](g["variant"]), g["variant"])
        r = pkg.render(sc, seed_mode=pkg.abi.SEED_WAVEFRONT, want_paths=True)
        assert [int(x) for x in r.state] == g["state"], name
        assert r.counters.tolist() == g["counters"], name
        bits = r.radiance.view(np.uint32 if r.radiance.dtype == np.float32 else np.uint64).ravel()
        assert [int(x) for x in bits] == g["radiance_bits"], name


def test_full_size_c2_properties(pkg, oracle):
    """BASELINE configs[1] at the size bench.py times (256 x 1 film, 2^14 spp = 4 194 304 paths), checked through
    size-independent properties: the event counters of any partition of the stream space add up exactly to those
    of the whole render, the films add up to summation accuracy, every pixel received exactly spp samples, and the
    image agrees with the oracle's estimate from an independent subsample of the same streams."""
    abi = pkg.abi
    spp, width = 16384, 256
    sc = pkg.load_dict(scenes.c2(spp=spp, width=width))
    keys = ("n_paths", "n_rng_draws", "n_medium_steps", "n_null", "n_real", "n_intersect", "n_surface", "n_invalid_samples")
    with pkg.DeviceScene(sc) as ds:
        n, _ = ds.stream_count(abi.SEED_WAVEFRONT)
        assert n == spp * width
        full = ds.render(seed_mode=abi.SEED_WAVEFRONT)
        cuts = [0, n // 3, n // 3 + 12345, n]
        parts = [ds.render(seed_mode=abi.SEED_WAVEFRONT, first_stream=a, n_streams=b - a) for a, b in zip(cuts, cuts[1:])]
        again = ds.render(seed_mode=abi.SEED_WAVEFRONT)
    for k in keys:
        assert sum(getattr(p.totals, k) for p in parts) == getattr(full.totals, k), k
        assert getattr(again.totals, k) == getattr(full.totals, k), k          # idempotent
    assert full.totals.n_paths == n and full.totals.n_invalid_samples == 0
    # W counts samples: spp per pixel, except that in float32 `pixel + u` rounds up to the next pixel for u close
    # to 1 (integrator.cpp:240-246 adds them in Float), which moves a sample to the neighbour -- the total is exact
    assert float(full.film[..., 4].astype(np.float64).sum()) == n
    assert np.all(np.abs(full.film[..., 4] - spp) <= 4)
    assert np.all(full.film[..., 3] <= full.film[..., 4])                           # alpha <= weight
    assert np.allclose(sum(p.film.astype(np.float64) for p in parts), full.film, rtol=2e-4, atol=1e-6)
    assert np.allclose(again.film, full.film, rtol=2e-4, atol=1e-6)
    # oracle estimate from every 64th sample of each pixel (same streams, so the same expectation)
    sub = 256
    ranges = [(px * spp + j * (spp // sub), 1) for px in range(0, width, 16) for j in range(sub)]
    est = {}
    for a, cnt in ranges:
        r = oracle.render(pkg, sc, abi.SEED_WAVEFRONT, first_stream=a, n_streams=cnt)
        px = a // spp
        est.setdefault(px, []).append(float(r.film[0, px, 1]))
    for px, vals in est.items():
        mean, sem = np.mean(vals), np.std(vals) / np.sqrt(len(vals))
        got = full.film[0, px, 1] / spp
        assert abs(got - mean) < 5 * sem + 1e-4, (px, got, mean, sem)
