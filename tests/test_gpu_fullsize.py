"""Parity at the sizes that are quoted: BASELINE configs[2] (C3: two 256^3 grids, 512 x 512 film, 64 spp) and
configs[4] (C5: 10^6 triangles, 4096 radiancemeters, 256 spp) at FULL size on the device, against the CPU oracle on a
strided subsample of the stream space -- PCG32 end state, every event counter and the radiance bits of every path
must be equal, in both precisions.  What only exists at full size is covered: the 257^3 trilinear cell records
(1.09 GB, 64-bit cell indices), the BVH over 10^6 triangles (depth, the index-median splits the builder falls back
to from level 24 on, the 64-entry traversal stack).  The oracle's Scene::ray_intersect is brute force over all
triangles (oracle/volpath_oracle.cpp), so its pieces run on all host threads."""
import functools
import os
import threading

import numpy as np
import pytest

import scenes

pytestmark = pytest.mark.gpu

KEYS = ("n_paths", "n_rng_draws", "n_medium_steps", "n_null", "n_real", "n_intersect", "n_surface", "n_invalid_samples")


def strided_pieces(n_streams, n_pieces, piece):
    """n_pieces ranges of `piece` consecutive streams spread over [0, n_streams), the last one ending at the end."""
    firsts = np.linspace(0, n_streams - piece, n_pieces).astype(np.int64)
    return [(int(a), piece) for a in firsts]


def oracle_pieces(pkg, oracle, sc, mode, pieces):
    out, lock, todo = {}, threading.Lock(), list(enumerate(pieces))

    def work():
        while True:
            with lock:
                if not todo:
                    return
                i, (lo, n) = todo.pop()
            r = oracle.render(pkg, sc, mode, first_stream=lo, n_streams=n, want_paths=True)
            with lock:
                out[i] = r
    ts = [threading.Thread(target=work) for _ in range(min(len(pieces), os.cpu_count() or 4, 16))]   # each call holds its own copy of the scene
    [t.start() for t in ts]
    [t.join() for t in ts]
    return [out[i] for i in range(len(pieces))]


def assert_pieces_equal(pkg, oracle, ds, sc, mode, pieces):
    refs = oracle_pieces(pkg, oracle, sc, mode, pieces)
    n_paths = 0
    for (lo, n), ref in zip(pieces, refs):
        got = ds.render(seed_mode=mode, first_stream=lo, n_streams=n, want_paths=True)
        assert np.array_equal(got.state, ref.state), ("PCG32 end states differ", lo)
        assert np.array_equal(got.counters, ref.counters), ("event counters differ", lo)
        assert np.array_equal(got.radiance.view(np.uint8), ref.radiance.view(np.uint8)), ("radiance bits differ", lo)
        for k in KEYS:
            assert getattr(got.totals, k) == getattr(ref.totals, k), (k, lo)
        n_paths += got.totals.n_paths
    return n_paths


@functools.lru_cache(maxsize=2)
def _cloud(n):
    return scenes.cloud_field(n)


def _c3_full(variant):
    d = scenes.c3(spp=64, variant=variant, n=8, size=512)     # small stand-in grids, replaced below (cached: 20 s to make)
    sigma_t, albedo = _cloud(256)
    d["atm"]["sigma_t"]["data"] = sigma_t
    d["atm"]["albedo"]["data"] = albedo
    return d


@pytest.mark.parametrize("cells", ["1", "0"])
@pytest.mark.parametrize("variant", ["scalar_mono", "scalar_mono_double"])
def test_full_size_c3_parity(pkg, oracle, variant, cells):
    """C3 at full size, trilinear cell records on (the product layout) and off (plain gathers)."""
    abi = pkg.abi
    sc = pkg.load_dict(_c3_full(variant), variant)
    sc.desc.options = 0 if cells == "1" else abi.OPT_NO_CELL_RECORDS
    with pkg.DeviceScene(sc) as ds:
        n, per = ds.stream_count(abi.SEED_WAVEFRONT)
        assert (n, per) == (512 * 512 * 64, 1)
        pieces = strided_pieces(n, 24, 24)
        assert assert_pieces_equal(pkg, oracle, ds, sc, abi.SEED_WAVEFRONT, pieces) == 24 * 24
        assert ds.info().last_kernel.decode().startswith("queue/heterogeneous+albedo grid")
        if variant == "scalar_mono" and cells == "1":
            # the whole render at the size tools/bench_configs.py times: sample bookkeeping is exact
            full = ds.render(seed_mode=abi.SEED_WAVEFRONT)
            assert full.totals.n_paths == n and full.totals.n_invalid_samples == 0
            assert float(full.film[..., 4].astype(np.float64).sum()) == n
            assert np.all(np.abs(full.film[..., 4] - 64) <= 4)
            parts = [ds.render(seed_mode=abi.SEED_WAVEFRONT, first_stream=a, n_streams=b - a)
                     for a, b in ((0, n // 2 + 777), (n // 2 + 777, n))]
            for k in KEYS:
                assert sum(getattr(p.totals, k) for p in parts) == getattr(full.totals, k), k


@pytest.mark.parametrize("variant", ["scalar_mono", "scalar_mono_double"])
def test_full_size_c5_parity(pkg, oracle, variant):
    """C5 at full size: 500 000 leaves = 10^6 triangles through the BVH, 4096 radiancemeters, 256 spp."""
    abi = pkg.abi
    sc = pkg.load_dict(scenes.c5(spp=256, variant=variant, n_leaves=500000, n_sensors=4096), variant)
    with pkg.DeviceScene(sc) as ds:
        info = ds.info()
        assert info.n_triangles == 1000000 + 12          # leaves + the slab cube
        assert info.bvh_leaves >= 500006 and info.bvh_nodes == 2 * info.bvh_leaves - 1
        assert 19 <= info.bvh_max_depth < info.bvh_stack_capacity == 64
        n, per = ds.stream_count(abi.SEED_WAVEFRONT)
        assert (n, per) == (4096 * 256, 1)
        pieces = strided_pieces(n, 32, 8)
        assert assert_pieces_equal(pkg, oracle, ds, sc, abi.SEED_WAVEFRONT, pieces) == 256
        assert ds.info().last_kernel.decode().startswith("queue/homogeneous")
        print("C5 BVH: %d nodes, depth %d, %d forced-median splits, %d degenerate splits" % (
            info.bvh_nodes, info.bvh_max_depth, info.bvh_forced_median_splits, info.bvh_degenerate_splits))


def _graded_mesh(n=160, ratio=0.6, copies=24):
    """A mesh no SAH builder likes: triangle i has size and position ~ ratio^i, so every binned split peels only the
    few largest triangles off and the tree gets one level deeper per ~5 triangles (160 triangles: ~30 levels, beyond
    the level where the builder switches to index-median splits); then `copies` identical triangles, whose centroids
    coincide (no SAH split exists) and whose hits tie exactly -- the later face must win (kdtree.h:2078-2171)."""
    v, f = [], []
    for i in range(n):
        s = ratio ** i
        base = len(v)
        v += [(s, 0.1 * s, 0.0), (1.4 * s, 0.1 * s, 0.0), (s, 0.5 * s, 0.0)]
        f.append((base, base + 1, base + 2))
    for _ in range(copies):
        base = len(v)
        v += [(-1.0, -1.0, 0.05), (-0.5, -1.0, 0.05), (-1.0, -0.5, 0.05)]
        f.append((base, base + 1, base + 2))
    return np.asarray(v, np.float32), np.asarray(f, np.uint32)


@pytest.mark.parametrize("variant", ["scalar_mono", "scalar_mono_double"])
def test_deep_bvh_forced_median_and_ties(pkg, oracle, variant):
    """The builder's fallbacks are reached (index-median splits at depth >= 24 and for coincident centroids) and the
    traversal over that tree still returns the reference's hit: counters and radiance equal the oracle's brute force."""
    abi = pkg.abi
    v, f = _graded_mesh()
    d = scenes.c5(spp=8, variant=variant, n_leaves=50, n_sensors=256)
    d["leaves"]["vertex_positions"], d["leaves"]["faces"] = v, f
    # radiancemeters straight above the graded triangles (one per triangle, at every scale) and above the stack of copies
    n_s = 256
    tri = np.arange(n_s) % 160
    s = 0.6 ** tri
    origins = np.stack([1.1 * s, 0.2 * s, np.full(n_s, 0.5)], axis=1)
    origins[200:] = [-0.9, -0.9, 0.5]
    dirs = np.tile([0.0, 0.0, -1.0], (n_s, 1))
    d["sensor"]["origins"] = origins.reshape(-1).tolist()
    d["sensor"]["directions"] = dirs.reshape(-1).tolist()
    sc = pkg.load_dict(d, variant)
    with pkg.DeviceScene(sc) as ds:
        info = ds.info()
        assert info.bvh_forced_median_splits > 0, "the depth-24 fallback was not reached: make the mesh more lopsided"
        assert info.bvh_degenerate_splits > 0, "no split over coincident centroids"
        assert 24 < info.bvh_max_depth < info.bvh_stack_capacity
        got = ds.render(seed_mode=abi.SEED_WAVEFRONT, want_paths=True)
    ref = oracle.render(pkg, sc, abi.SEED_WAVEFRONT, want_paths=True)
    assert np.array_equal(got.state, ref.state)
    assert np.array_equal(got.counters, ref.counters)
    assert np.array_equal(got.radiance.view(np.uint8), ref.radiance.view(np.uint8))
    assert got.totals.n_surface > 0 and (got.radiance > 0).any()
