#!/usr/bin/env python
"""C5 (BASELINE.json configs[4]): standalone RNS NTT / iNTT sweep, N = 2^12..2^17 x {1,2,4,8,16,32,60} limbs of
60-bit primes, batched through the C ABI (ppq_ntt_forward / ppq_ntt_inverse), next to the CPU oracle's NTT on all
host threads.  Algorithmic bytes per limb transform = 16 N.  Prints a markdown table (commit it under profiles/).

    python tools/ntt_sweep.py [--cpu] > profiles/rNN_ntt_sweep.md
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def chain(logN, n):
    """n primes below 2^60 congruent 1 mod 2N: PreviousPrime chain from FirstPrime(60, 2N)."""
    import oracle
    L = oracle.lib()
    m = 2 << logN
    q = L.ppqo_first_prime(60, m)
    out = []
    for _ in range(n):
        q = L.ppqo_previous_prime(q, m)
        out.append(q)
    return np.array(out, dtype=np.uint64)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cpu", action="store_true", help="also time the CPU oracle (bounded sample)")
    ap.add_argument("--logn", type=int, nargs="*", default=[12, 13, 14, 15, 16, 17])
    ap.add_argument("--limbs", type=int, nargs="*", default=[1, 2, 4, 8, 16, 32, 60])
    args = ap.parse_args()
    import torch
    import oracle
    from ppqsflhe_b200 import CryptoContext
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(
        os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    dev = torch.device("cuda", 0)
    oracle.lib().ppqo_first_prime.restype = oracle.C.c_uint64
    oracle.lib().ppqo_previous_prime.restype = oracle.C.c_uint64
    oracle.lib().ppqo_first_prime.argtypes = [oracle.C.c_int, oracle.C.c_uint64]
    oracle.lib().ppqo_previous_prime.argtypes = [oracle.C.c_uint64, oracle.C.c_uint64]
    print(f"| logN | limbs | polys | MiB | fwd GB/s | fwd frac | inv GB/s | inv frac | fwd limb-NTT/s | CPU limb-NTT/s ({oracle.max_threads()} thr) | GPU/CPU |")
    print("|---:|---:|---:|---:|---:|---:|---:|---:|---:|---:|---:|")
    for logN in args.logn:
        N = 1 << logN
        for nl in args.limbs:
            q = chain(logN, nl)
            # the aux basis is derived but unused here; limbs 0..nl-1 are the Q chain (a digit holds <= 32 limbs)
            dnum = 1 if nl <= 32 else 2
            cc = CryptoContext.from_moduli(logN, q, dnum=dnum, device=0)
            reps = max(1, (256 << 20) // (nl * N * 8))          # working set >= 256 MiB (> L2)
            ids = np.tile(np.arange(nl, dtype=np.int32), reps)
            qv = torch.from_numpy(q.astype(np.int64)).to(dev).repeat(reps).view(-1, 1)
            buf = torch.randint(0, 1 << 62, (len(ids), N), dtype=torch.int64, device=dev).remainder_(qv)
            res = {}
            for name, inv in (("fwd", False), ("inv", True)):
                for _ in range(3):
                    cc.ntt(buf, ids, inverse=inv)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(5):
                    cc.ntt(buf, ids, inverse=inv)
                e1.record()
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / 5
                res[name] = (16.0 * N * len(ids) / (ms * 1e-3) / 1e9, len(ids) / (ms * 1e-3))
            cpu = ""
            ratio = ""
            if args.cpu:
                o = oracle.Oracle.from_moduli(logN, q, dnum=dnum)
                n_cpu = min(len(ids), max(nl, (64 << 20) // (N * 8)))
                h = buf[:n_cpu].cpu().numpy().view(np.uint64).copy()
                o.ntt_batch(h, ids[:n_cpu])
                t0 = time.perf_counter()
                o.ntt_batch(h, ids[:n_cpu])
                dt = time.perf_counter() - t0
                cpu = f"{n_cpu / dt:.0f}"
                ratio = f"{res['fwd'][1] / (n_cpu / dt):.0f}x"
                o.close()
            print(f"| {logN} | {nl} | {len(ids)} | {buf.numel() * 8 >> 20} | {res['fwd'][0]:.0f} | {res['fwd'][0] / peak:.3f} | "
                  f"{res['inv'][0]:.0f} | {res['inv'][0] / peak:.3f} | {res['fwd'][1]:.3e} | {cpu} | {ratio} |", flush=True)
            del buf, qv
            cc.close()
            torch.cuda.empty_cache()
    print(f"\npeak = {peak} GB/s (MEASURED_PEAKS.json hbm_gbs); bytes per limb transform = 16 N (algorithmic; the 2-pass "
          "transform moves 32 N); inputs > L2; CUDA events, 5 iterations after 3 warm-ups.")


if __name__ == "__main__":
    main()
