"""Host-side logic of the N > 1 path on CPU (gloo, world_size 2): slab scatter / gather of the
global-with-ghost arrays, the object all-gather that carries the peer blobs, and max/sum
reductions used for timing.  The oracle stands in for the solver (each rank checks that the
slab it would hand to its GPU is the slab of the oracle's block decomposition)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from cice5_b200 import fixtures as fx


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = fx.synthetic_tripole(48, 20)
        case = fx.make_case(g, "full", ndte=4, dt=3600.0)
        mine = {k: fx.slab(v, rank, world) for k, v in case.inp.items()}
        assert all(v.shape == (22, 48 // world + 2) for v in mine.values())
        # ghost columns of my slab are my neighbours' edge columns (cyclic)
        nxl = 48 // world
        a = case.inp["uocn"]
        assert np.array_equal(mine["uocn"][:, 0], a[:, rank * nxl])
        assert np.array_equal(mine["uocn"][:, -1], a[:, rank * nxl + nxl + 1])
        blobs = [None] * world
        dist.all_gather_object(blobs, (rank, bytes([rank]) * 80))     # same call as the peer blobs
        assert [b[0] for b in blobs] == list(range(world))
        gathered = [None] * world
        dist.gather_object(mine, gathered if rank == 0 else None, dst=0)
        t = torch.tensor([float(rank + 1)], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        assert t.item() == world
        ok = True
        if rank == 0:
            for k in case.inp:
                ok &= np.array_equal(fx.unslab([p[k] for p in gathered]), case.inp[k])
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


def test_slab_scatter_gather_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert sorted(res) == [(0, True), (1, True)]


@pytest.mark.parametrize("n", [1, 2, 4, 8])
def test_slab_unslab_roundtrip(n):
    rng = np.random.default_rng(n)
    a = rng.standard_normal((12, 34))
    assert np.array_equal(fx.unslab([fx.slab(a, r, n) for r in range(n)]), a)


def test_unequal_slabs_roundtrip_and_bounds():
    rng = np.random.default_rng(7)
    a = rng.standard_normal((9, 42))                       # nx = 40
    starts = [0, 3, 17, 19, 33]
    parts = [fx.slab(a, r, 5, starts) for r in range(5)]
    assert [p.shape[1] - 2 for p in parts] == [3, 14, 2, 14, 7]
    assert np.array_equal(fx.unslab(parts), a)
    assert fx.slab_bounds(40, 4) == [0, 10, 20, 30, 40]


@pytest.mark.parametrize("nranks", [2, 4, 8])
def test_balanced_slab_starts_minimise_the_heaviest_rank(nranks):
    """Strip-aligned partition (fixtures.balanced_slab_starts): boundaries on multiples of the strip
    width, every rank non-empty, and the maximum work not worse than equal slabs of whole strips."""
    g = fx.make_grid("tp1")
    ocean = g.f["tmask"][1:-1, 1:-1] > 0
    strip = 31
    starts = fx.balanced_slab_starts(ocean, nranks, strip=strip)
    b = starts + [g.dom.nx]
    assert starts[0] == 0 and all(x % strip == 0 for x in starts) and all(np.diff(b) >= 2)

    def work(lo, hi):
        return sum(ocean[:, k:min(k + strip, hi)].any(axis=1).sum() for k in range(lo, hi, strip))
    heaviest = max(work(b[r], b[r + 1]) for r in range(nranks))
    nstrips = -(-g.dom.nx // strip)
    per = -(-nstrips // nranks)
    eq = [min(r * per * strip, g.dom.nx) for r in range(nranks)] + [g.dom.nx]
    assert heaviest <= max(work(eq[r], eq[r + 1]) for r in range(nranks) if eq[r] < eq[r + 1])
    # narrow grid: falls back to cutting the cumulative ocean-cell count
    s2 = fx.balanced_slab_starts(ocean[:, :40], 8, strip=124)
    assert len(s2) == 8 and s2[0] == 0 and all(np.diff(s2 + [40]) >= 2)
