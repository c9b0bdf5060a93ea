"""CPU-only tests of the multi-GPU host logic: slab partition, global numbering offsets, and the
count exchange over a real 2-rank process group (gloo)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from vacancy_b200 import sharded


def test_slab_ranges_cover_grid():
    for nz, world in ((1024, 8), (42, 4), (7, 3), (2048, 8), (5, 2)):
        r = sharded.slab_ranges(nz, world)
        assert r[0][0] == 0 and r[-1][1] == nz and len(r) == world
        assert all(a[1] == b[0] for a, b in zip(r[:-1], r[1:]))
        sizes = [b - a for a, b in r]
        assert min(sizes) >= 2 and max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sharded.slab_ranges(7, 4)


def test_rebalanced_ranges():
    ranges = sharded.slab_ranges(1024, 8)
    # equal costs: nothing moves
    assert sharded.rebalanced_ranges(ranges, [1.0] * 8) == ranges
    # heavier edge slabs get thinner, the partition stays contiguous and complete
    new = sharded.rebalanced_ranges(ranges, [118.8, 105.8, 102.5, 102.5, 101.9, 102.9, 106.5, 118.5])
    assert new[0][0] == 0 and new[-1][1] == 1024
    assert all(a[1] == b[0] for a, b in zip(new[:-1], new[1:]))
    sizes = [b - a for a, b in new]
    assert sizes[0] < 128 and sizes[-1] < 128 and max(sizes) > 128 and min(sizes) >= 2
    # degenerate costs still leave every slab its two halo planes
    tiny = sharded.rebalanced_ranges(sharded.slab_ranges(8, 4), [100.0, 0.0, 0.0, 0.0])
    assert all(b - a >= 2 for a, b in tiny) and tiny[-1][1] == 8
    # whole fusion tiles: every boundary a multiple of the quantum, still contiguous and complete
    q = sharded.FUSION_TILE_PLANES
    for costs in ([118.8, 105.8, 102.5, 102.5, 101.9, 102.9, 106.5, 118.5], [1.0] * 8,
                  [53.5, 51.0, 52.0, 51.0, 50.5, 51.4, 50.1, 49.4]):
        new = sharded.rebalanced_ranges(ranges, costs, quantum=q)
        assert new[0][0] == 0 and new[-1][1] == 1024
        assert all(a[1] == b[0] and a[1] % q == 0 for a, b in zip(new[:-1], new[1:]))
        assert all(b - a >= q for a, b in new)
    assert sharded.rebalanced_ranges(ranges, [1.0] * 8, quantum=q) == ranges
    # too few planes for whole tiles: falls back to single planes
    tiny = sharded.rebalanced_ranges(sharded.slab_ranges(8, 4), [100.0, 0.0, 0.0, 0.0], quantum=q)
    assert all(b - a >= 2 for a, b in tiny) and tiny[-1][1] == 8
    odd = sharded.rebalanced_ranges(sharded.slab_ranges(54, 3), [3.0, 2.0, 1.0], quantum=q)
    assert odd[0][0] == 0 and odd[-1][1] == 54 and all(b - a >= 2 for a, b in odd)


def test_exclusive_offsets():
    off, total = sharded.exclusive_offsets([(5, 9), (0, 0), (7, 2)])
    assert off.tolist() == [[0, 0], [5, 9], [5, 9]] and total.tolist() == [12, 11]


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    off, total = sharded.gather_offsets(100 * (rank + 1), 7 + rank, world)
    # the NCCL unique id the one-process-per-GPU engine is built from: made on rank 0 by the library,
    # carried to the other ranks by torch.distributed (the only job torch has on that path)
    uid = sharded.broadcast_unique_id(rank, world)
    # measured slab costs -> the same new partition on every rank
    new = sharded.agreed_ranges(sharded.slab_ranges(1024, world), [60.0, 40.0][rank], world)
    q.put((rank, off.tolist(), total.tolist(), uid, new))
    dist.destroy_process_group()


def test_count_exchange_world2_gloo():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, off, total, uid, new in res:
        assert off == [[0, 0], [100, 7]] and total == [300, 15]
    assert res[0][4] == res[1][4] and res[0][4][0][1] % 8 == 0 and res[0][4][0][1] < 512
    assert len(res[0][3]) == 128 and res[0][3] == res[1][3] and any(res[0][3])


def test_balance_bounds_host_only():
    """vcb_balance_bounds: the cut the engine's auto-balance makes from its per-layer costs (no GPU
    needed): equal costs give equal slabs, cheap ends give thick end slabs, every slab keeps a
    layer, boundaries are whole layers, a short last layer is allowed."""
    import ctypes as C
    from vacancy_b200 import capi
    lib = capi.load_library()

    def cut(cost, ppl, nz, world):
        arr = (C.c_uint64 * len(cost))(*[int(c) for c in cost])
        out = (C.c_int32 * (world + 1))()
        rc = lib.vcb_balance_bounds(arr, len(cost), ppl, nz, world, out)
        return rc, list(out)

    assert cut([10] * 128, 8, 1024, 8) == (0, [0, 128, 256, 384, 512, 640, 768, 896, 1024])
    # the shape measured on cfg4: the ends cost more per plane than the middle
    cost = [120] * 32 + [100] * 64 + [120] * 32
    rc, b = cut(cost, 8, 1024, 8)
    sizes = [b[i + 1] - b[i] for i in range(8)]
    assert rc == 0 and b[0] == 0 and b[-1] == 1024 and all(x % 8 == 0 for x in b)
    assert sizes[0] < sizes[3] and sizes[7] < sizes[4] and sizes == sizes[::-1]
    per_slab = [sum(cost[b[i] // 8:b[i + 1] // 8]) for i in range(8)]
    assert max(per_slab) - min(per_slab) <= 2 * 120  # within a layer or two of each other
    # all the cost in one layer: the others still get one layer each
    rc, b = cut([0, 0, 1000, 0, 0, 0], 8, 42, 3)
    assert rc == 0 and b[0] == 0 and b[-1] == 42 and b[1] >= 8 and b[2] - b[1] >= 8 and b[2] <= 40
    # short last layer (42 planes = 5 full layers + 2 planes), and refused arguments
    assert cut([1] * 6, 8, 42, 2)[1] == [0, 24, 42]
    assert cut([1] * 6, 8, 48, 7)[0] == -1     # fewer layers than slabs
    assert cut([1] * 5, 8, 42, 2)[0] == -1     # layer count does not match nz
