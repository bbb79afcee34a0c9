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
