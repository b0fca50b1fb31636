"""Cancel, timeout and multi-device renders behind the C-ABI.

* b2vp_cancel from another thread stops a RUNNING render (Integrator::cancel, include/mitsuba/render/integrator.h:44-51;
  the reference polls should_stop() in its block loop, src/librender/integrator.cpp:121): the persistent kernel polls a
  device flag where it fetches new streams.  The flag is cleared when the next render starts (m_stop = false,
  integrator.cpp:59), so the handle stays usable.
* timeout_s (the integrator's `timeout`, integrator.h:143-146): the render returns what it has, normalised.
* n_devices > 1: interleaved stream blocks per device, one NCCL reduce of the films.  Needs two GPUs:
  run with `gpurun --gpus 2`; skipped on a single-GPU box."""
import threading
import time

import numpy as np
import pytest

import scenes

pytestmark = pytest.mark.gpu
KEYS = ("n_paths", "n_rng_draws", "n_medium_steps", "n_null", "n_real", "n_intersect", "n_surface", "n_invalid_samples")


def _long_scene(pkg):
    return pkg.load_dict(scenes.c2(spp=1 << 19, width=256))      # 134 M paths: about half a second on a B200


def test_cancel_stops_a_running_render(pkg):
    abi = pkg.abi
    sc = _long_scene(pkg)
    with pkg.DeviceScene(sc) as ds:
        ds.render(n_streams=4096)                                  # upload, warm up
        t0 = time.perf_counter()
        full = ds.render(math="fast")
        t_full = time.perf_counter() - t0
        threading.Timer(0.05, ds.cancel).start()
        t0 = time.perf_counter()
        with pytest.raises(pkg.B2vpError) as e:
            ds.render(math="fast")
        t_cancel = time.perf_counter() - t0
        assert e.value.code == abi.ERR_CANCELLED
        assert t_cancel < 0.5 * t_full, "cancel took %.3f s of a %.3f s render" % (t_cancel, t_full)
        # the handle is not dead: the next render runs to the end (the advisor's finding on round 1)
        again = ds.render(math="fast")
        for k in KEYS:
            assert getattr(again.totals, k) == getattr(full.totals, k), k
        assert again.totals.stopped_early == 0
        # a cancel BETWEEN renders has no effect on the next one (m_stop = false at the top of render)
        ds.cancel()
        small = ds.render(n_streams=8192)
        assert small.totals.n_paths == 8192


def test_timeout_returns_a_normalised_partial_film(pkg):
    sc = _long_scene(pkg)
    with pkg.DeviceScene(sc) as ds:
        ds.render(n_streams=4096)
        t0 = time.perf_counter()
        r = ds.render(math="fast", timeout=0.05)
        dt = time.perf_counter() - t0
        total = 256 * (1 << 19)
        assert r.totals.stopped_early == 1 and 0 < r.totals.n_paths < total
        assert dt < 0.3
        # every sample that was taken is in the film, with its weight: the developed image is a valid estimate
        assert float(r.film[..., 4].astype(np.float64).sum()) == pytest.approx(r.totals.n_paths, rel=1e-6)
        ref = ds.render(n_streams=1 << 22, math="fast")
        m_r, m_ref = r.film[..., 1].sum() / r.film[..., 4].sum(), ref.film[..., 1].sum() / ref.film[..., 4].sum()
        assert abs(m_r - m_ref) < 0.05 * m_ref
        # no timeout: runs to the end
        r = ds.render(n_streams=1 << 20, timeout=60.0)
        assert r.totals.stopped_early == 0 and r.totals.n_paths == 1 << 20


def _n_devices(pkg):
    return pkg.abi.lib().b2vp_device_count()


@pytest.mark.parametrize("variant", ["scalar_mono", "scalar_mono_double"])
@pytest.mark.parametrize("name", ["c2", "c3_small", "c5_small", "c2_blocks"])
def test_multi_device_equals_single_device(pkg, name, variant):
    n = _n_devices(pkg)
    if n < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    abi = pkg.abi
    make = {"c2": lambda: scenes.c2(spp=256, width=64, variant=variant),
            "c3_small": lambda: scenes.c3(spp=8, n=32, size=48, variant=variant),
            "c5_small": lambda: scenes.c5(spp=64, n_leaves=400, n_sensors=256, variant=variant),
            "c2_blocks": lambda: scenes.c2(spp=32, width=64, variant=variant, block_size=8)}[name]
    mode = abi.SEED_SCALAR_BLOCKS if name == "c2_blocks" else abi.SEED_WAVEFRONT
    sc = pkg.load_dict(make(), variant)
    devices = list(range(min(n, 8)))
    with pkg.DeviceScene(sc) as ds:
        one = ds.render(seed_mode=mode)
        for devs in (devices[:2], devices):
            many = ds.render(seed_mode=mode, devices=devs)
            for k in KEYS:     # the same streams, whoever renders them
                assert getattr(many.totals, k) == getattr(one.totals, k), (k, devs)
            tol = 2e-4 if one.film.dtype == np.float32 else 1e-12
            assert np.allclose(many.film, one.film, rtol=tol, atol=tol * 1e-2), devs
            assert many.totals.n_launches == len(devs)
        # accumulate into an existing film; a sub-range; the majorant-grid mode
        acc = ds.render(seed_mode=mode, devices=devices, film=one.film.copy())
        assert np.allclose(acc.film, 2 * one.film.astype(np.float64), rtol=4e-4, atol=1e-6)
        total, _ = ds.stream_count(mode)
        part = ds.render(seed_mode=mode, devices=devices, first_stream=total // 3, n_streams=total // 2)
        ref = ds.render(seed_mode=mode, first_stream=total // 3, n_streams=total // 2)
        for k in KEYS:
            assert getattr(part.totals, k) == getattr(ref.totals, k), k
        # a parameter update reaches every resident copy
        ds.update(medium=0, medium_scale=0.37)
        a, b = ds.render(seed_mode=mode), ds.render(seed_mode=mode, devices=devices)
        for k in KEYS:
            assert getattr(a.totals, k) == getattr(b.totals, k), k
        assert a.totals.n_medium_steps != one.totals.n_medium_steps


def test_multi_device_rejections(pkg):
    abi = pkg.abi
    sc = pkg.load_dict(scenes.c2(spp=4, width=16))
    with pkg.DeviceScene(sc) as ds:
        with pytest.raises(pkg.B2vpError, match="devices\\[0\\]"):
            ds.render(devices=[1, 0])
        with pytest.raises(pkg.B2vpError, match="twice"):
            ds.render(devices=[0, 0])
        with pytest.raises(pkg.B2vpError, match="single-device"):
            ds.render(devices=[0, 1], want_paths=True)
        if _n_devices(pkg) < 2:
            with pytest.raises(pkg.B2vpError, match="out of range"):
                ds.render(devices=[0, 1])
