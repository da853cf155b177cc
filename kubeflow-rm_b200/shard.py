"""Key-hash sharding of reconcile work across ranks (one process per GPU).

Notebooks are independent keys: rank r reconciles the keys with
shard_of(namespace, name, world) == r.  A key always lands on the same rank, so
controller-runtime's guarantee that a key is never reconciled concurrently with
itself still holds; no data-path collective is needed.  The only
cross-rank traffic is the barrier and the max-over-ranks of the timings."""
import zlib


def shard_of(namespace, name, world):
    return zlib.crc32(("%s/%s" % (namespace, name)).encode()) % max(world, 1)


def my_keys(keys, rank, world):
    return [k for k in keys if shard_of(k[0], k[1], world) == rank]
