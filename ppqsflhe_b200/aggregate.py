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


def clients_of_rank(n_clients, world, rank):
    """Round-robin client partition: rank r owns clients {c : c mod world == r} (its rekeys live only there)."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    return [c for c in range(n_clients) if c % world == rank]


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
