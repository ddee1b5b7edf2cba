"""Multi-GPU aggregation plumbing (one process per GPU, torch.distributed).

The reference aggregates on one host, serially (server/src/aggregateEncryptedWeights.cpp:68-110).  Here every
client's ciphertexts re-encrypt independently, so clients are sharded across ranks with NO data-path collective
during ReEncrypt; the only exchange is ONE sum-reduce of the per-rank partial aggregates (NCCL over NVLink on the
GPU box, gloo in the CPU tests), followed by a mod-q canonicalisation on the root.

Residues are canonical (< q < 2^60) uint64 stored as int64 bit patterns; a sum over `world` ranks stays below
world * 2^60, so it cannot wrap in 64 bits for world <= 15 and the two's-complement int64 sum has the same bits as
the uint64 sum.
"""
MAX_WORLD = 15


def clients_of_rank(n_clients, world, rank, counts=None):
    """Client partition: rank r owns its clients' rekeys and nobody else's.  Without `counts`: round-robin,
    {c : c mod world == r}.  With `counts` (plan_shards): contiguous blocks of counts[r] clients."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    if counts is None:
        return [c for c in range(n_clients) if c % world == rank]
    if len(counts) != world or sum(counts) != n_clients or min(counts) < 0:
        raise ValueError("counts must be one non-negative entry per rank summing to n_clients")
    first = sum(counts[:rank])
    return list(range(first, first + counts[rank]))


def plan_shards(n_clients, cost):
    """Clients per rank when the ranks do not move data at the same rate (ppq_plan_shards, include/ppq_b200.h): cost[r] =
    seconds per client on rank r; minimises max_r counts[r] * cost[r].  Host only."""
    import ctypes as C
    from ._capi import check, lib
    world = len(cost)
    c = (C.c_double * world)(*[float(x) for x in cost])
    out = (C.c_size_t * world)()
    check(lib().ppq_plan_shards(n_clients, world, c, out))
    return [int(x) for x in out]


def reduce_partial_sums(acc, world, dst=0, group=None):
    """Sum the per-rank partial aggregates into rank `dst` (in place on dst).  acc: int64 tensor of uint64 bits.
    The caller canonicalises on dst afterwards (CryptoContext.finalize_sum)."""
    if world > MAX_WORLD:
        raise ValueError(f"a {world}-way uint64 sum of 60-bit residues could wrap; reduce hierarchically")
    if world == 1:
        return acc
    import torch.distributed as dist
    dist.reduce(acc, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return acc
