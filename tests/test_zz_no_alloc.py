"""ppq_encrypt_noalloc (include/ppq_b200.h): ppq_encrypt with its scratch in a file-scope device array instead of
cudaMalloc'd buffers.  No counterpart in the reference, so the check is
against the existing entry point ppq_encrypt, which the oracle pins (tests/test_gpu_parity.py): bit-exact, the tolerance of
every integer test here.  Named test_zz_* so that pytest collects it last."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

MAX_BYTES = 64 << 20  # PPQ_NOALLOC_SCRATCH_BYTES (csrc/capi.cu)


def _inputs(torch, cc, ell, n, seed):
    g = torch.Generator(device=cc.device)
    g.manual_seed(seed)
    qv = torch.from_numpy(cc.moduli.astype(np.int64)).to(cc.device)

    def rr(shape, q):
        return torch.randint(0, 1 << 62, shape, dtype=torch.int64, device=cc.device, generator=g).remainder_(q)

    pkb = rr((cc.M, cc.N), qv.view(-1, 1))
    pka = rr((cc.M, cc.N), qv.view(-1, 1))
    pt = rr((n, ell, cc.N), qv[:ell].view(1, ell, 1))
    # (v, e0, e1): ternary and small signed noise, as the samplers produce them
    small = torch.randint(-1, 2, (n, 3, cc.N), dtype=torch.int64, device=cc.device, generator=g)
    small[:, 1:, :] = torch.randint(-19, 20, (n, 2, cc.N), dtype=torch.int64, device=cc.device, generator=g)
    return pkb, pka, pt, small


@pytest.mark.gpu
def test_encrypt_noalloc_matches_encrypt_on_two_calls_in_a_row():
    import torch
    from ppqsflhe_b200 import CryptoContext
    from ppqsflhe_b200._capi import check, lib, vp
    PPQ_EINVAL = -1  # include/ppq_b200.h
    text = open(os.path.join(ROOT, "tests", "golden", "CC.json")).read()
    cc = CryptoContext.from_ccjson(text, 0)
    try:
        L = lib()
        # two calls in a row on the shared array, the second with another shape (fewer limbs, more ciphertexts), so stale
        # contents of the first call sit exactly where the second one works
        for call, (ell, n, seed) in enumerate([(4, 5, 11), (3, 7, 12), (4, 5, 11)]):
            pkb, pka, pt, small = _inputs(torch, cc, ell, n, seed)
            want = torch.zeros((n, 2, ell, cc.N), dtype=torch.int64, device=cc.device)
            got = torch.full_like(want, -1)
            check(L.ppq_encrypt(cc.h, ell, vp(pkb.data_ptr()), vp(pka.data_ptr()), vp(pt.data_ptr()), vp(small.data_ptr()),
                                vp(want.data_ptr()), n, None))
            torch.cuda.synchronize()
            before = torch.cuda.memory_allocated()
            check(L.ppq_encrypt_noalloc(cc.h, ell, vp(pkb.data_ptr()), vp(pka.data_ptr()), vp(pt.data_ptr()),
                                        vp(small.data_ptr()), vp(got.data_ptr()), n, None))
            torch.cuda.synchronize()
            assert torch.cuda.memory_allocated() == before
            assert torch.equal(got, want), f"call {call}: ppq_encrypt_noalloc differs from ppq_encrypt"
            assert int((want != 0).sum().item()) > want.numel() // 2  # the comparison is not over zeros

        # the largest batch that fits, and the first one that does not (C1: 3 * 4 * (8 N + 4) bytes per ciphertext)
        ell = 4
        per_ct = 3 * ell * (8 * cc.N + 4)
        n_max = MAX_BYTES // per_ct
        pkb, pka, pt, small = _inputs(torch, cc, ell, n_max + 1, 13)
        want = torch.zeros((n_max + 1, 2, ell, cc.N), dtype=torch.int64, device=cc.device)
        got = torch.zeros_like(want)
        args = (vp(pkb.data_ptr()), vp(pka.data_ptr()), vp(pt.data_ptr()), vp(small.data_ptr()))
        check(L.ppq_encrypt(cc.h, ell, *args, vp(want.data_ptr()), n_max, None))
        check(L.ppq_encrypt_noalloc(cc.h, ell, *args, vp(got.data_ptr()), n_max, None))
        torch.cuda.synchronize()
        assert torch.equal(got[:n_max], want[:n_max])
        rc = L.ppq_encrypt_noalloc(cc.h, ell, *args, vp(got.data_ptr()), n_max + 1, None)
        assert rc == PPQ_EINVAL
        msg = L.ppq_last_error().decode()
        assert "maximum" in msg and str(MAX_BYTES) in msg
        assert int(got[n_max].abs().sum().item()) == 0  # a refused call launches nothing
        # and the library is still usable afterwards
        check(L.ppq_encrypt_noalloc(cc.h, ell, *args, vp(got.data_ptr()), 1, None))
        torch.cuda.synchronize()
        assert torch.equal(got[0], want[0])
    finally:
        cc.close()
