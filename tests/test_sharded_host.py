"""Host-side logic of the row-sharded build, on CPU: the partition used by the C library, and the
in-place all-gather exchange through torch.distributed (backend gloo, world_size 2 and 3)."""
import os
import socket

import numpy as np
import pytest


def test_partition_tiles_the_range():
    from gravomg_b200.sharded import partition
    for count in (0, 1, 31, 32, 33, 1000, 40962, 2621442):
        for world in (1, 2, 3, 4, 8):
            chunk = partition(count, world, 0)[0]
            assert chunk % 32 == 0 and chunk * world >= count
            prev = 0
            for r in range(world):
                c, lo, hi = partition(count, world, r)
                assert c == chunk and lo == prev and lo <= hi <= count
                prev = hi
            assert prev == count


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, count, q):
    import torch
    import torch.distributed as dist
    from gravomg_b200.sharded import TorchComm, partition
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        comm = TorchComm()
        ok = True
        # three exchange buffers with the element sizes the build uses (state u8, parents i32, dist u64)
        for elem in (1, 4, 8):
            chunk, lo, hi = partition(count, world, rank)
            full = torch.zeros(chunk * world * elem, dtype=torch.uint8)
            ref = (np.arange(chunk * world * elem, dtype=np.int64) * 7 + elem) % 251
            mine = torch.from_numpy(ref[lo * elem:hi * elem].astype(np.uint8))
            full[lo * elem:hi * elem] = mine                       # a rank writes only what it owns
            comm.all_gather_inplace(full, chunk * elem)
            got = full.numpy()
            for r in range(world):
                _, rlo, rhi = partition(count, world, r)
                ok = ok and np.array_equal(got[rlo * elem:rhi * elem], ref[rlo * elem:rhi * elem].astype(np.uint8))
        # slices of different sizes (packed rows): received in place around the rank's own slice
        sizes = [(7 * r + 3) % 11 * 8 for r in range(world)]          # bytes per rank, some empty
        off = np.concatenate([[0], np.cumsum(sizes)]).astype(int).tolist()
        ref = ((np.arange(off[-1], dtype=np.int64) * 13 + 5) % 241).astype(np.uint8)
        full = torch.zeros(off[-1], dtype=torch.uint8)
        full[off[rank]:off[rank + 1]] = torch.from_numpy(ref[off[rank]:off[rank + 1]])
        comm.all_gather_v(full, off)
        ok = ok and np.array_equal(full.numpy(), ref)
        ok = ok and comm.bytes_gathered > 0
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_all_gather_exchange_gloo(world):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, 1000, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(r, True) for r in range(world)]


def _halo_worker(rank, world, port, n, q):
    """A 1-D periodic stencil (i-2 .. i+2) row-sharded over `world` ranks: every rank lists the ghost entries its
    rows read, the plan is exchanged with all_to_all_single, and one halo exchange must fill exactly those ghosts
    with the owners' values -- for the whole pattern and for a "colour" sub-plan (i % 3)."""
    import torch
    import torch.distributed as dist
    from gravomg_b200.sharded import partition, plan_over_dist, exchange_over_dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        chunk, lo, hi = partition(n, world, rank)
        own = np.arange(lo, hi)
        reads = np.unique(np.concatenate([(own + d) % n for d in (-2, -1, 1, 2)])) if hi > lo else np.zeros(0, dtype=np.int64)
        ghost_np = reads[(reads < lo) | (reads >= hi)]
        ghost = torch.from_numpy(ghost_np.astype(np.int64))
        plan = plan_over_dist(dist, None, ghost, chunk, world)
        ok = sum(plan.recv_counts[0]) == len(ghost_np)
        # what the others read from me must be mine
        sidx = plan.send_idx[0].numpy()
        ok = ok and bool(np.all((sidx >= lo) & (sidx < hi)))
        truth = np.arange(n, dtype=np.float64) * 1.5 + 3.0
        vec = torch.full((chunk * world,), -1.0, dtype=torch.float64)
        vec[lo:hi] = torch.from_numpy(truth[lo:hi])
        exchange_over_dist(dist, None, plan, vec)
        got = vec.numpy()
        ok = ok and np.array_equal(got[ghost_np], truth[ghost_np])
        untouched = np.ones(chunk * world, dtype=bool)
        untouched[lo:hi] = False
        untouched[ghost_np] = False
        ok = ok and bool(np.all(got[untouched] == -1.0))
        # colour sub-plan: both sides filter with the same rule
        colour = torch.arange(chunk * world) % 3
        sub = plan.select([colour[plan.ghost[0]] == 1], [colour[plan.send_idx[0]] == 1])
        vec2 = torch.full((chunk * world,), -1.0, dtype=torch.float64)
        vec2[lo:hi] = torch.from_numpy(truth[lo:hi])
        exchange_over_dist(dist, None, sub, vec2)
        g1 = ghost_np[ghost_np % 3 == 1]
        g0 = ghost_np[ghost_np % 3 != 1]
        ok = ok and np.array_equal(vec2.numpy()[g1], truth[g1]) and bool(np.all(vec2.numpy()[g0] == -1.0))
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n", [(2, 200), (3, 1000), (4, 97)])
def test_halo_exchange_gloo(world, n):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_halo_worker, args=(r, world, port, n, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(r, True) for r in range(world)]
