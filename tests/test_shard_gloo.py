"""CPU, world_size 2 over gloo: key-hash sharding covers every key exactly once and the
cross-rank plumbing bench.py uses (barrier, MAX all-reduce of timings) works."""
import importlib
import os
import socket

import __graft_entry__ as graft


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    graft.load_pkg()
    shard = importlib.import_module(graft.PKG_NAME + ".shard")
    keys = [("ns-%03d" % (i // 100), "nb-%06d" % i) for i in range(5000)]
    mine = shard.my_keys(keys, rank, world)
    count = torch.tensor([len(mine)], dtype=torch.int64)
    dist.all_reduce(count)
    t = torch.tensor([1.0 + rank], dtype=torch.float64)      # "my elapsed ms"
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    gathered = [None] * world
    dist.all_gather_object(gathered, set(mine))
    dist.barrier()
    if rank == 0:
        q.put((int(count.item()), float(t.item()), len(set.union(*gathered)),
               sum(len(g) for g in gathered), [len(g) for g in gathered]))
    dist.destroy_process_group()


def test_sharding_world_size_2():
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    total, tmax, union, summed, sizes = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert total == 5000 and union == 5000 and summed == 5000      # every key once, no overlap
    assert tmax == 2.0
    assert min(sizes) > 2000                                       # and the split is even enough
