"""CPU tests of the multi-GPU host logic (no GPU): the sample-index split covers every index exactly once for any
world size / pass size, and a world_size-2 gloo run reproduces the single-process film sum and the NCCL-id exchange.
The device side (ncclReduce of the films) is covered by tests/test_multigpu_gpu.py."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_split_covers_every_sample_once(pkg):
    M = pkg.multigpu
    for spp in (0, 1, 2, 3, 7, 64, 255, 1024):
        for world in (1, 2, 3, 4, 8):
            for per_pass in (0, 1, 3, 16):
                seen = []
                for rank in range(world):
                    mine = []
                    for begin, count in M.passes(spp, world, rank, per_pass):
                        assert count > 0 and (per_pass <= 0 or count <= per_pass)
                        mine += M.pass_indices(begin, count, world, rank)
                    assert mine == M.sample_indices(spp, world, rank)
                    assert len(mine) == M.sample_share(spp, world, rank)
                    seen += mine
                assert sorted(seen) == list(range(spp))
                shares = [M.sample_share(spp, world, r) for r in range(world)]
                assert max(shares) - min(shares) <= 1          # balanced to within one sample per pixel
    with pytest.raises(ValueError):
        M.sample_share(4, 2, 2)


def _fake_film(indices, shape=(6, 5, 5)):
    """Stand-in for a rank's film: a deterministic function of the sample indices it rendered."""
    f = np.zeros(shape, np.float64)
    for s in indices:
        f += np.random.default_rng(1000 + s).random(shape)
    return f


def _worker(rank, world, port, spp, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    import __graft_entry__ as graft
    M = graft.load_package().multigpu
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        uid = M.exchange_unique_id(dist, lambda: bytes(range(128)), rank)
        idx = []
        for begin, count in M.passes(spp, world, rank, 3):
            idx += M.pass_indices(begin, count, world, rank)
        film = torch.from_numpy(_fake_film(idx))
        dist.reduce(film, dst=0, op=dist.ReduceOp.SUM)       # what mtsb_film_reduce does with ncclReduce on the device
        q.put((rank, uid, film.numpy() if rank == 0 else None, len(idx)))
    finally:
        dist.destroy_process_group()


def test_world2_gloo_film_sum(pkg):
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    spp, world = 13, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, spp, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = {}
    for _ in range(world):
        rank, uid, film, n = q.get(timeout=120)
        got[rank] = (uid, film, n)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert got[0][0] == got[1][0] == bytes(range(128))
    assert got[0][2] + got[1][2] == spp
    assert np.allclose(got[0][1], _fake_film(range(spp)), rtol=1e-12)
