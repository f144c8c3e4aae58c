"""N > 1 host logic on CPU: world_size-2 gloo processes render their sample-range shares (with the
oracle standing in for the GPU, which this container does not have), reduce the partial films with
the same call bench.py uses, and must reproduce the single-rank film."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest

REPO = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(REPO))
sys.path.insert(0, str(REPO / "tests"))

from first_raytracer_b200 import sharding  # noqa: E402


def test_shares_tile_the_range():
    for world in (1, 2, 3, 4, 8):
        for lo, hi in ((0, 1024), (0, 5), (7, 8), (3, 3), (100, 1123)):
            shares = [sharding.spp_share(lo, hi, r, world) for r in range(world)]
            assert shares[0][0] == lo and shares[-1][1] == hi
            assert all(a[1] == b[0] for a, b in zip(shares, shares[1:]))
            sizes = [e - b for b, e in shares]
            assert max(sizes) - min(sizes) <= 1
        blocks = [sharding.ray_share(1_000_003, r, world) for r in range(world)]
        assert blocks[0][0] == 0 and blocks[-1][1] == 1_000_003
        assert all(a[1] == b[0] for a, b in zip(blocks, blocks[1:]))
    with pytest.raises(ValueError):
        sharding.spp_share(0, 8, 2, 2)
    with pytest.raises(ValueError):
        sharding.ray_share(10, 0, 0)


def _worker(rank, world, port, nx, ny, spp, out_path):
    import torch
    import torch.distributed as dist
    from first_raytracer_b200 import api
    from oracle_c import Oracle
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        scene = api.HostScene("cornell", nx / ny, 2)
        s0, s1 = sharding.spp_share(0, spp, rank, world)
        part = np.zeros((ny, nx, 3))
        if s1 > s0:
            part = Oracle(scene).render(nx, ny, s1 - s0, seed=5, spp_begin=s0) * (s1 - s0)  # back to sums
        film = torch.from_numpy(np.ascontiguousarray(part))
        sharding.reduce_film(film, dst=0)
        if rank == 0:
            np.save(out_path, film.numpy() / spp)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,spp", [(2, 6), (2, 1)])
def test_two_ranks_reproduce_one(built, tmp_path, world, spp):
    import torch.multiprocessing as mp
    from first_raytracer_b200 import api
    from oracle_c import Oracle
    nx, ny = 24, 16
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = str(tmp_path / "film.npy")
    mp.spawn(_worker, args=(world, port, nx, ny, spp, out), nprocs=world, join=True)
    got = np.load(out)
    whole = Oracle(api.HostScene("cornell", nx / ny, 2)).render(nx, ny, spp, seed=5)
    assert got.shape == whole.shape
    np.testing.assert_allclose(got, whole, rtol=1e-12, atol=1e-15)
    assert whole.mean() > 0
