#!/usr/bin/env python
"""client_times.py -- where the client half goes (encode / sample / Encrypt / Decrypt core / decode), CUDA events, per ct.
    python tools/client_times.py [--config c3] [--cts 54]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    from bench import CONFIGS
    from ppqsflhe_b200 import CryptoContext
    from ppqsflhe_b200._capi import check, lib, vp
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="c3")
    ap.add_argument("--cts", type=int, default=54)
    a = ap.parse_args()
    cfg = CONFIGS[a.config]
    cc = CryptoContext.synthetic(cfg["logN"], cfg["depth"], cfg["first"], cfg["scale"], cfg["extra"], cfg["dnum"], 0)
    n = a.cts
    sk, pk = cc.KeyGen(seed=1)
    vals = torch.empty((n, cc.N // 2), dtype=torch.float64, device=cc.device).normal_(0, 0.05).clamp_(-0.3, 0.3)
    small = torch.empty((n, 3, cc.N), dtype=torch.int64, device=cc.device)

    def t(fn, reps=5):
        fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            r = fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e3 / reps / n, r

    us_enc, pt = t(lambda: cc.MakeCKKSPackedPlaintext(vals))
    us_smp, _ = t(lambda: check(lib().ppq_sample_encrypt_noise(cc.h, None, vp(small.data_ptr()), n, None)))
    us_ect, ct = t(lambda: cc.Encrypt(pk, pt, small=small))
    us_fus, ct2 = t(lambda: cc.EncodeEncrypt(pk, vals, small=small))
    assert torch.equal(ct2.data, ct.data)
    us_fus2, _ = t(lambda: cc.EncodeEncrypt(pk, vals))
    us_dec, co = t(lambda: cc.DecryptCore(sk, ct))
    us_dcd, _ = t(lambda: cc.Decrypt(sk, ct))
    print(f"[{a.config}] per ct (us): encode {us_enc:.1f}  sample {us_smp:.1f}  encrypt {us_ect:.1f}  encode+encrypt fused {us_fus:.1f} (with sampling {us_fus2:.1f})  decrypt_core {us_dec:.1f}  "
          f"decrypt+decode {us_dcd:.1f}   ({cc.L} limbs, N=2^{cc.logN}, {n} cts)")


if __name__ == "__main__":
    main()
