"""world_size-2 gloo test (CPU) of the N>1 aggregation logic: client sharding + one int64 sum-reduce of canonical
partial sums + mod-q on the root must equal the single-process aggregate.  The compute on each rank is the
oracle here (no GPU in this container); the GPU finalize kernel is covered by the -m gpu tests."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    import oracle
    from ppqsflhe_b200.aggregate import clients_of_rank, plan_shards, reduce_partial_sums
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    o = oracle.Oracle.synthetic(12, 2, 60, 40, 20, 2)
    ell, n_clients, n_slots = o.L, 5, 2
    rng = np.random.default_rng(7)  # same stream on every rank: the full workload, then take my shard
    mods = o.moduli.astype(np.uint64)
    cts = rng.integers(0, 1 << 62, (n_clients, n_slots, 2, ell, o.N), dtype=np.uint64) % mods[:ell].reshape(1, 1, 1, ell, 1)
    keys = [(rng.integers(0, 1 << 62, (o.dnum, o.M, o.N), dtype=np.uint64) % mods.reshape(1, o.M, 1),
             rng.integers(0, 1 << 62, (o.dnum, o.M, o.N), dtype=np.uint64) % mods.reshape(1, o.M, 1))
            for _ in range(n_clients)]
    # ranks that move data at different rates get unequal shards (ppq_plan_shards); the sum must not care
    counts = plan_shards(n_clients, [1.0, 2.5, 1.7][:world])
    assert sum(counts) == n_clients and counts[0] > counts[1]
    mine = clients_of_rank(n_clients, world, rank, counts)
    part = o.reencrypt_accumulate_batch(ell, cts[mine], [keys[c][0] for c in mine], [keys[c][1] for c in mine])
    t = torch.from_numpy(part.view(np.int64).copy())
    reduce_partial_sums(t, world, dst=0)
    if rank == 0:
        lazy = t.numpy().view(np.uint64)
        got = lazy % mods[:ell].reshape(1, 1, ell, 1)
        want = o.reencrypt_accumulate_batch(ell, cts, [k[0] for k in keys], [k[1] for k in keys])
        q.put(bool(np.array_equal(got, want)) and bool((lazy >= want).all()))
    dist.destroy_process_group()


def test_two_rank_sharded_aggregate_matches_single_process():
    torch = pytest.importorskip("torch")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    ok = q.get(timeout=180)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ok


def test_partition_and_overflow_guard():
    from ppqsflhe_b200.aggregate import MAX_WORLD, clients_of_rank, reduce_partial_sums
    for world in (1, 2, 3, 8):
        seen = sorted(c for r in range(world) for c in clients_of_rank(64, world, r))
        assert seen == list(range(64))
        sizes = [len(clients_of_rank(64, world, r)) for r in range(world)]
        assert max(sizes) - min(sizes) <= 1
    assert MAX_WORLD * (1 << 60) < (1 << 64)
    with pytest.raises(ValueError):
        reduce_partial_sums(None, 16)
    with pytest.raises(ValueError):
        clients_of_rank(4, 2, 2)


def test_shard_plan_minimises_the_slowest_rank():
    import itertools
    from ppqsflhe_b200.aggregate import clients_of_rank, plan_shards
    # equal costs: balanced; the 8-GPU box's measured copy rates (23.1 GB/s on four ranks, 35.1 GB/s on the others)
    assert sorted(plan_shards(64, [1.0] * 8)) == [8] * 8
    c = plan_shards(64, [1 / 23.1] * 4 + [1 / 35.1] * 4)
    assert c == [6, 6, 6, 6, 10, 10, 10, 10]
    # optimal against brute force on small cases, and deterministic
    for cost in ([1.0, 2.0], [3.0, 1.0, 2.0], [1.0, 1.0, 5.0], [0.5, 0.7, 0.9, 1.1]):
        for n in (0, 1, 5, 9):
            got = plan_shards(n, cost)
            assert sum(got) == n and got == plan_shards(n, cost)
            best = min(max(k * w for k, w in zip(ks, cost))
                       for ks in itertools.product(range(n + 1), repeat=len(cost)) if sum(ks) == n)
            assert max(k * w for k, w in zip(got, cost)) == pytest.approx(best)
            seen = sorted(x for r in range(len(cost)) for x in clients_of_rank(n, len(cost), r, got))
            assert seen == list(range(n))
    from ppqsflhe_b200._capi import PPQError
    for bad in ([1.0, 0.0], [1.0, -2.0], [float("nan")], [float("inf"), 1.0]):
        with pytest.raises(PPQError):
            plan_shards(4, bad)
    with pytest.raises(ValueError):
        clients_of_rank(4, 2, 0, [3, 2])
