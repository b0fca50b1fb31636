"""b2vp_render_noalloc (include/b2vp.h): the render entry point that performs no device allocation -- its kernels'
scratch lives in the library's static device array.  It has no counterpart in the reference,
so the check is against the existing entry point, b2vp_render, through the C-ABI: same
launch totals exactly, same film within the summation-order tolerance the other GPU tests use for two renders of
one film (tests/test_gpu_parity.py: rtol 2e-4, atol 1e-6), on two calls in a row (the scratch array is shared and
keeps the previous call's contents).  Collected last on purpose (pytest -x)."""
import ctypes as C

import numpy as np
import pytest

import scenes

pytestmark = pytest.mark.gpu

RTOL, ATOL = 2e-4, 1e-6
TOTALS = ("n_paths", "n_rng_draws", "n_medium_steps", "n_null", "n_real", "n_intersect", "n_surface",
          "n_invalid_samples")

CASES = {
    # queue kernel, film accumulated in shared memory, grid staged by TMA
    "c2": (lambda v: scenes.c2(spp=64, variant=v, width=64), "wavefront"),
    # queue kernel, film in HBM, two 3D grids
    "c3_small": (lambda v: scenes.c3(spp=4, variant=v, n=32, size=64), "wavefront"),
    # spiral block table in the scratch array (the reference's scalar seeding order)
    "c2_blocks": (lambda v: scenes.c2(spp=8, variant=v, width=64, block_size=8), "blocks"),
    # BVH scene
    "c5_small": (lambda v: scenes.c5(spp=8, variant=v, n_leaves=400, n_sensors=64), "wavefront"),
    # one-thread-per-stream kernel (volpathmis): no path records, totals and work counter only
    "c2_volpathmis": (lambda v: dict(scenes.c2(spp=8, variant=v, width=64),
                                     integrator={"type": "volpathmis", "max_depth": -1}), "wavefront"),
}


def _device_film(torch, sc, dtype):
    w, h = sc.film_size
    return torch.full((h, w, sc.film_channels), 7.0, dtype=dtype, device="cuda:0")   # non-zero: the call must clear it


@pytest.mark.parametrize("variant", ["scalar_mono", "scalar_mono_double"])
@pytest.mark.parametrize("name", sorted(CASES))
def test_noalloc_matches_render(pkg, name, variant):
    torch = pytest.importorskip("torch")
    make, seeding = CASES[name]
    sc = pkg.load_dict(make(variant))
    mode = pkg.abi.SEED_WAVEFRONT if seeding == "wavefront" else pkg.abi.SEED_SCALAR_BLOCKS
    tdtype = torch.float32 if sc.precision == pkg.abi.PREC_F32 else torch.float64
    with pkg.DeviceScene(sc, 0) as ds:
        ref = ds.render(seed_mode=mode)                    # existing entry point; also makes the scene resident
        film = _device_film(torch, sc, tdtype)
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info(0)[0]
        for call in range(2):                              # two calls in a row share the static scratch
            tot = ds.render_noalloc(film.data_ptr(), seed_mode=mode)
            for k in TOTALS:
                assert getattr(tot, k) == getattr(ref.totals, k), (name, variant, call, k)
            assert tot.n_launches == 1
            got = film.cpu().numpy()
            assert np.allclose(got, ref.film, rtol=RTOL, atol=ATOL), (name, variant, call)
            film.fill_(3.0)
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info(0)[0] >= free_before, "device memory was allocated during b2vp_render_noalloc"


def test_noalloc_accumulates_and_shards(pkg):
    """clear_film = 0 accumulates into the caller's film; two half ranges sum to the full render."""
    torch = pytest.importorskip("torch")
    sc = pkg.load_dict(scenes.c2(spp=32, variant="scalar_mono", width=64))
    with pkg.DeviceScene(sc, 0) as ds:
        full = ds.render()
        n, _ = ds.stream_count(pkg.abi.SEED_WAVEFRONT)
        film = torch.zeros((1, 64, 5), dtype=torch.float32, device="cuda:0")
        a = ds.render_noalloc(film.data_ptr(), first_stream=0, n_streams=n // 2, clear=True)
        b = ds.render_noalloc(film.data_ptr(), first_stream=n // 2, n_streams=n - n // 2, clear=False)
        assert a.n_paths + b.n_paths == full.totals.n_paths
        assert a.n_rng_draws + b.n_rng_draws == full.totals.n_rng_draws
        assert np.allclose(film.cpu().numpy(), full.film, rtol=RTOL, atol=ATOL)


def test_noalloc_rejections(pkg):
    """Clear errors: scene not resident yet, host film, per-stream outputs, scratch beyond the static array."""
    torch = pytest.importorskip("torch")
    abi = pkg.abi
    sc = pkg.load_dict(scenes.c2(spp=4, variant="scalar_mono", width=16))
    film = torch.zeros((1, 16, 5), dtype=torch.float32, device="cuda:0")
    with pkg.DeviceScene(sc, 0) as ds:
        with pytest.raises(abi.B2vpError, match="not resident"):
            ds.render_noalloc(film.data_ptr())
        ds.render()
        ds.render_noalloc(film.data_ptr())
        a = abi.RenderArgs()
        a.abi_version, a.struct_size = abi.ABI_VERSION, C.sizeof(abi.RenderArgs)
        a.precision, a.seed_mode, a.device, a.film_on_device = abi.PREC_F32, abi.SEED_WAVEFRONT, 0, 0
        assert abi.lib().b2vp_render_noalloc(ds._h, C.byref(a), film.data_ptr()) == abi.ERR_INVALID
        assert b"device buffer" in abi.lib().b2vp_last_error()
        a.film_on_device = 1
        state = np.zeros(16 * 4, dtype=np.uint64)
        a.out_state = state.ctypes.data_as(C.POINTER(C.c_uint64))
        assert abi.lib().b2vp_render_noalloc(ds._h, C.byref(a), film.data_ptr()) == abi.ERR_UNSUPPORTED
    # 2048^2 one-pixel spiral blocks x 16 B = 64 MiB: the block table alone fills the static array, so the launch
    # (table + header + path records) is over the limit whatever the path-record size is
    big = scenes.c2(spp=1, variant="scalar_mono", width=2048, block_size=1)
    big["sensor"]["film"]["height"] = 2048
    sc = pkg.load_dict(big)
    with pkg.DeviceScene(sc, 0) as ds:
        ds.render(n_streams=64)
        film = torch.zeros((2048, 2048, 5), dtype=torch.float32, device="cuda:0")
        with pytest.raises(abi.B2vpError, match="B2VP_NOALLOC_SCRATCH_BYTES"):
            ds.render_noalloc(film.data_ptr(), seed_mode=abi.SEED_SCALAR_BLOCKS)
        ds.render_noalloc(film.data_ptr())   # wavefront seeding of the same scene needs no block table: fits
