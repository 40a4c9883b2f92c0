"""Host logic of the multi-GPU path on CPU: the z-slab split tiles the grid, slabs reassemble to the full array across a
world_size-2 gloo group (the exchange a scene does around mf_grid_upload_slab / mf_grid_download_slab), multigrid level ranges
cover every level exactly once."""
import os

import numpy as np
import pytest

from mantaflow_b200 import partition as P


@pytest.mark.parametrize("nz,n", [(4, 2), (20, 3), (256, 8), (1024, 8), (33, 4)])
def test_zsplit_tiles(nz, n):
    b = P.zsplit(nz, n)
    assert b[0] == 0 and b[-1] == nz and all(b[i] < b[i + 1] for i in range(n))


def test_zsplit_rejects_thin_grids():
    with pytest.raises(ValueError):
        P.zsplit(3, 2)


@pytest.mark.parametrize("size,n", [((256, 256, 256), 8), ((1024, 1024, 1024), 8), ((64, 96, 64), 2), ((24, 32, 20), 3)])
def test_mg_levels_are_covered_once(size, n):
    for l, (sx, sy, sz) in enumerate(P.mg_level_sizes(size)):
        if l == 0:
            continue
        bz = (sz + 1) // 2
        covered = []
        for r in range(n):
            a, b = P.mg_level_pairs(size, l, r, n)
            covered += list(range(a, b))
        assert covered == list(range(bz)), (l, covered)
    assert P.mg_level_sizes((256, 256, 256))[-1] == (9, 9, 9)


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    full = np.random.RandomState(7).rand(10, 4, 6, 3).astype(np.float32)      # same on every rank
    mine = np.ascontiguousarray(P.slab_of(full, rank, world)) * np.float32(2.0)   # "compute" on the own slab
    b = P.zsplit(full.shape[0], world)
    parts = [torch.empty((b[r + 1] - b[r],) + full.shape[1:], dtype=torch.float32) for r in range(world)]
    # ragged all-gather = one broadcast per rank
    for r in range(world):
        if r == rank:
            parts[r].copy_(torch.from_numpy(mine))
        dist.broadcast(parts[r], src=r)
    whole = P.assemble([p.numpy() for p in parts])
    ok = bool(np.array_equal(whole, full * np.float32(2.0)))
    t = torch.tensor([1.0 if ok else 0.0])
    dist.all_reduce(t)
    if rank == 0:
        q.put(float(t.item()))
    dist.destroy_process_group()


def test_slabs_reassemble_across_two_ranks():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 400)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=10) == 2.0
