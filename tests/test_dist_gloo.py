"""world_size-2 gloo test of the sharding + film-reduce host logic (the N>1 path of bench.py), on CPU.
The per-rank renderer here is the oracle (this is a test); on GPUs it is the CUDA path."""
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys
sys.path.insert(0, %(root)r); sys.path.insert(0, os.path.join(%(root)r, "tests"))
import numpy as np, torch.distributed as dist
import scenes
from oracle import oracle
pkg = scenes.pkg
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
sc = pkg.load_dict(scenes.c2(spp=6, width=8, nz=50))
total = 8 * 6
first, count = pkg.shard_range(total, rank, world)
r = oracle.render(pkg, sc, pkg.abi.SEED_WAVEFRONT, first_stream=first, n_streams=count)
film = pkg.reduce_film(r.film.astype(np.float64))
if rank == 0:
    np.save(os.path.join(%(out)r, "film.npy"), film)
# the same render sharded by interleaved blocks: every rank renders 3 of the 6 samples of every pixel
r = oracle.render(pkg, sc, pkg.abi.SEED_WAVEFRONT, **pkg.shard_interleaved(6, 8, rank, world))
assert (r.film[..., 4] == 3).all()
film = pkg.reduce_film(r.film.astype(np.float64))
if rank == 0:
    np.save(os.path.join(%(out)r, "film_interleaved.npy"), film)
dist.destroy_process_group()
'''


def test_two_rank_sharded_render_matches_single(pkg, oracle, tmp_path):
    import scenes
    script = tmp_path / "worker.py"
    script.write_text(WORKER % {"root": ROOT, "out": str(tmp_path)})
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    subprocess.check_call([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                           "--master-addr", "127.0.0.1", "--master-port", "29533", str(script)], env=env, timeout=600)
    film = np.load(tmp_path / "film.npy")
    sc = pkg.load_dict(scenes.c2(spp=6, width=8, nz=50))
    full = oracle.render(pkg, sc, pkg.abi.SEED_WAVEFRONT)
    assert np.allclose(film, full.film, rtol=1e-5, atol=1e-7)
    assert film[..., 4].sum() == 48
    film = np.load(tmp_path / "film_interleaved.npy")
    assert np.allclose(film, full.film, rtol=1e-5, atol=1e-7)
    assert film[..., 4].sum() == 48


def test_shard_range_partitions(pkg):
    for n in (0, 1, 7, 48, 4194304):
        for w in (1, 2, 3, 8):
            r = [pkg.shard_range(n, k, w) for k in range(w)]
            assert r[0][0] == 0 and sum(c for _, c in r) == n
            assert all(r[i][0] + r[i][1] == r[i + 1][0] for i in range(w - 1))
            assert max(c for _, c in r) - min(c for _, c in r) <= 1


def test_shard_streams_partitions(pkg):
    """b2vp_shard_streams (the partition of a multi-device b2vp_render; host arithmetic, no GPU): the interleaved
    blocks of the n devices are disjoint, cover [0, count) exactly, are balanced to within one block, and the
    kernel's index map (block a power of two: shift, multiply, mask) lands every local stream in its own block."""
    for count in (1, 31, 32, 33, 1000, 4096, 65537, 4194304, 16777216 + 5):
        for n in (1, 2, 3, 4, 8):
            seen, sizes, total = None, [], 0
            if count <= 65537:
                seen = np.zeros(count, dtype=np.int32)
            for i in range(n):
                block, ci = pkg.shard_streams(count, n, i)
                assert block & (block - 1) == 0 and block >= 32
                sizes.append(ci)
                total += ci
                if seen is not None and ci:
                    k = np.arange(ci, dtype=np.uint64)
                    sh = int(block).bit_length() - 1
                    g = (((k >> np.uint64(sh)) * np.uint64(n) + np.uint64(i)) << np.uint64(sh)) | (k & np.uint64(block - 1))
                    assert g.max() < count
                    np.add.at(seen, g.astype(np.int64), 1)
            assert total == count, (count, n, sizes)
            assert max(sizes) - min(sizes) <= block
            if seen is not None:
                assert (seen == 1).all(), (count, n)
    with __import__("pytest").raises(pkg.B2vpError):
        pkg.shard_streams(100, 0, 0)
