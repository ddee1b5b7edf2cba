#!/usr/bin/env python
"""fanout_times.py -- SURVEY 8 row f2 on a B200: the aggregate (cts at ell-1 limbs) re-encrypted to K target domains,
ppq_reencrypt_fanout (ModUp shared over the keys) against K separate ppq_reencrypt calls.  CUDA events, us per
(ciphertext, key).   python tools/fanout_times.py [--config c3] [--keys 8] [--reps 3]
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    from bench import CONFIGS
    from ppqsflhe_b200 import CiphertextBatch, CryptoContext, CtMeta, EvalKey

    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="c3")
    ap.add_argument("--keys", type=int, default=8)
    ap.add_argument("--reps", type=int, default=3)
    a = ap.parse_args()
    cfg = CONFIGS[a.config]
    dev = torch.device("cuda", 0)
    cc = CryptoContext.synthetic(cfg["logN"], cfg["depth"], cfg["first"], cfg["scale"], cfg["extra"], cfg["dnum"], 0)
    ell, M, N, n = cc.L - 1, cc.M, cc.N, cfg["slots"]  # the aggregate has lost one limb to the 1/C rescale
    qv = torch.from_numpy(cc.moduli.astype(np.int64)).to(dev)
    torch.manual_seed(1)

    def rr(shape, q):
        return torch.randint(0, 1 << 62, shape, dtype=torch.int64, device=dev).remainder_(q)

    ct = CiphertextBatch(rr((n, 2, ell, N), qv[:ell].view(1, 1, ell, 1)), CtMeta())
    keys = [EvalKey(cc, rr((cc.dnum, M, N), qv.view(1, M, 1)), rr((cc.dnum, M, N), qv.view(1, M, 1))) for _ in range(a.keys)]

    def timed(fn):
        fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.reps):
            out = fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e3 / (a.reps * a.keys * n), out

    t_sep, o_sep = timed(lambda: [cc.ReEncrypt(ct, rk) for rk in keys])
    t_fan, o_fan = timed(lambda: cc.ReEncryptFanOut(ct, keys))
    same = all(torch.equal(x.data, y.data) for x, y in zip(o_sep, o_fan))
    print(f"{a.config} ell={ell} cts={n} keys={a.keys}: separate {t_sep:.1f} us/(ct,key)  fan-out {t_fan:.1f} us/(ct,key)  "
          f"x{t_sep / t_fan:.2f}  identical={same}")


if __name__ == "__main__":
    main()
