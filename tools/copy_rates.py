#!/usr/bin/env python
"""copy_rates.py -- host->device copy rate of every rank while ALL ranks copy at once (what bounds the e2e leg of
bench.py at N > 1), and of each rank copying alone:
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/copy_rates.py [--mib 1024]
Prints one line per rank: GB/s with everybody copying, GB/s alone.
"""
import argparse
import os
import time

import torch
import torch.distributed as dist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mib", type=int, default=1024)
    ap.add_argument("--reps", type=int, default=8)
    a = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n = a.mib << 17  # int64 words
    host = [torch.empty(n, dtype=torch.int64).pin_memory() for _ in range(2)]
    for h in host:
        h.fill_(rank)
    dev = [torch.empty(n, dtype=torch.int64, device="cuda") for _ in range(2)]

    def run():
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for r in range(a.reps):
            dev[r & 1].copy_(host[r & 1], non_blocking=True)
        e1.record()
        torch.cuda.synchronize()
        return a.reps * n * 8 / (e0.elapsed_time(e1) * 1e-3) / 1e9

    run()
    if world > 1:
        dist.barrier()
    together = run()
    alone = [0.0] * world
    for r in range(world):
        if world > 1:
            dist.barrier()
        if r == rank:
            alone[r] = run()
    if world > 1:
        dist.barrier()
    res = torch.tensor([together, alone[rank]], device="cuda", dtype=torch.float64)
    if world > 1:
        out = [torch.zeros_like(res) for _ in range(world)]
        dist.all_gather(out, res)
    else:
        out = [res]
    if rank == 0:
        tot = 0.0
        for r, o in enumerate(out):
            print(f"rank {r}: {o[0].item():6.1f} GB/s with all {world} ranks copying, {o[1].item():6.1f} GB/s alone")
            tot += o[0].item()
        print(f"sum over ranks while all copy: {tot:.1f} GB/s")
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
