#!/usr/bin/env python
"""stage_times.py -- per-stage microseconds per ciphertext of ReEncrypt+EvalAdd (CUDA events around every launch,
ppq_profile_*) for one build of libppq_b200.so; used to A/B kernel variants on a B200:
    PPQ_B200_LIB=/path/to/variant.so python tools/stage_times.py [--config c3] [--clients 2] [--reps 3]
"""
import argparse
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    from bench import CONFIGS
    from ppqsflhe_b200 import CiphertextBatch, CryptoContext, CtMeta, EvalKey
    from ppqsflhe_b200._capi import check, lib

    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="c3")
    ap.add_argument("--clients", type=int, default=2)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--wave", type=int, default=0)
    ap.add_argument("--legacy", action="store_true", help="per-client full ModDown (ppq_reencrypt_accumulate) instead of ppq_agg_*")
    ap.add_argument("--tag", default=os.environ.get("PPQ_B200_LIB", "default"))
    a = ap.parse_args()
    cfg = CONFIGS[a.config]
    dev = torch.device("cuda", 0)
    cc = CryptoContext.synthetic(cfg["logN"], cfg["depth"], cfg["first"], cfg["scale"], cfg["extra"], cfg["dnum"], 0)
    if a.wave:
        cc.set_wave(a.wave)
    ell, M, N, n = cc.L, cc.M, cc.N, cfg["slots"]
    qv = torch.from_numpy(cc.moduli.astype(np.int64)).to(dev)
    torch.manual_seed(1)

    def rr(shape, q):
        return torch.randint(0, 1 << 62, shape, dtype=torch.int64, device=dev).remainder_(q)

    cts = [CiphertextBatch(rr((n, 2, ell, N), qv[:ell].view(1, 1, ell, 1)), CtMeta()) for _ in range(a.clients)]
    keys = [EvalKey(cc, rr((cc.dnum, M, N), qv.view(1, M, 1)), rr((cc.dnum, M, N), qv.view(1, M, 1))) for _ in range(a.clients)]
    acc = CiphertextBatch(torch.zeros((n, 2, ell, N), dtype=torch.int64, device=dev), CtMeta())

    from ppqsflhe_b200 import Aggregator
    agg = None if a.legacy else Aggregator(cc, ell, n)

    def step():
        if a.legacy:
            for ct, rk in zip(cts, keys):
                cc.ReEncryptAccumulate(ct, rk, acc)
        else:
            agg.add_many(cts, keys)
            agg.finish(out=acc.data)

    step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.reps):
        step()
    e1.record()
    torch.cuda.synchronize()
    tot = e0.elapsed_time(e1) * 1e3 / (a.reps * a.clients * n)
    check(lib().ppq_profile_enable(cc.h, 1))
    for _ in range(a.reps):
        step()
    torch.cuda.synchronize()
    NST = 11
    pms, pl = (C.c_double * NST)(), (C.c_uint64 * NST)()
    check(lib().ppq_profile_read(cc.h, pms, pl, 1))
    check(lib().ppq_profile_enable(cc.h, 0))
    names = [lib().ppq_profile_stage_name(i).decode() for i in range(NST)]
    per = [m * 1e3 / (a.reps * a.clients * n) for m in pms]
    csum = int(acc.data.sum().item()) & 0xFFFFFFFFFFFF
    print(f"[{a.tag}] total {tot:.1f} us/ct ({1e6 / tot:.0f} ct/s)  checksum {csum:012x}")
    print("   " + "  ".join(f"{nm.split('(')[0][2:]}{'(' + nm.split('(')[1] if '(' in nm else ''}={v:.1f}" for nm, v in zip(names, per)))


if __name__ == "__main__":
    main()
