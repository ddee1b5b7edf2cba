"""Parity tests proper: the sm_100a path (through the C ABI) against the CPU oracle on the same seeded inputs.
Bit-exact for every integer stage; CKKS encode/decode within the stated tolerance.  Run with -m gpu on a B200."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

import oracle  # noqa: E402  (the checker)

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _cc_pair(kind):
    from ppqsflhe_b200 import CryptoContext
    if kind == "c1":
        text = open(os.path.join(ROOT, "tests", "golden", "CC.json")).read()
        return oracle.Oracle.from_ccjson(text), CryptoContext.from_ccjson(text, 0)
    logN, depth, dnum, scale = kind
    o = oracle.Oracle.synthetic(logN, depth, 60, scale, 20, dnum)
    cc = CryptoContext.synthetic(logN, depth, 60, scale, 20, dnum, 0)
    return o, cc


def _rand_poly(rng, o, limbs):
    return np.stack([rng.integers(0, int(o.moduli[r]), o.N, dtype=np.uint64) for r in limbs])


def _rand_ct(rng, o, ell, n):
    return np.stack([np.stack([_rand_poly(rng, o, range(ell)) for _ in range(2)]) for _ in range(n)])


def _rand_key(rng, o):
    kb = np.stack([_rand_poly(rng, o, range(o.M)) for _ in range(o.dnum)])
    ka = np.stack([_rand_poly(rng, o, range(o.M)) for _ in range(o.dnum)])
    return kb, ka


def test_context_matches_oracle_and_ccjson():
    o, cc = _cc_pair("c1")
    assert (cc.logN, cc.L, cc.K, cc.dnum, cc.alpha) == (o.logN, o.L, o.K, o.dnum, o.alpha) == (14, 4, 2, 2, 2)
    assert np.array_equal(cc.moduli, o.moduli) and np.array_equal(cc.roots, o.roots)
    assert [cc.sf(i) for i in range(4)] == [o.sf(i) for i in range(4)]
    assert cc.sf_big(0) == o.sf_big(0)
    for kind in [(12, 3, 2, 40), (16, 18, 3, 59)]:
        o2, cc2 = _cc_pair(kind)
        assert np.array_equal(cc2.moduli, o2.moduli) and np.array_equal(cc2.roots, o2.roots)
        assert (cc2.L, cc2.K, cc2.alpha) == (o2.L, o2.K, o2.alpha)


@pytest.mark.parametrize("logN", [12, 13, 14, 15, 16, 17])
def test_ntt_bit_exact(logN):
    from ppqsflhe_b200 import to_device_u64, to_host_u64
    o, cc = _cc_pair((logN, 3, 2, 45))
    rng = np.random.default_rng(100 + logN)
    limbs = list(range(o.M)) + [0, o.M - 1]
    x = _rand_poly(rng, o, limbs)
    d = to_device_u64(x, cc.device)
    cc.ntt(d, limbs)
    got = to_host_u64(d)
    for i, r in enumerate(limbs):
        assert np.array_equal(got[i], o.ntt(r, x[i])), f"forward NTT differs, limb {r}"
    cc.ntt(d, limbs, inverse=True)
    assert np.array_equal(to_host_u64(d), x), "inverse(forward(x)) != x"
    # inverse alone against the oracle
    d = to_device_u64(x, cc.device)
    cc.ntt(d, limbs, inverse=True)
    got = to_host_u64(d)
    for i, r in enumerate(limbs):
        assert np.array_equal(got[i], o.ntt(r, x[i], inverse=True)), f"inverse NTT differs, limb {r}"


def test_ntt_edge_values():
    from ppqsflhe_b200 import to_device_u64, to_host_u64
    o, cc = _cc_pair("c1")
    limbs = list(range(o.M))
    for fill in ("zero", "max", "delta"):
        x = np.zeros((o.M, o.N), dtype=np.uint64)
        for r in limbs:
            if fill == "max":
                x[r, :] = o.moduli[r] - np.uint64(1)
            elif fill == "delta":
                x[r, 1] = 1
        d = to_device_u64(x, cc.device)
        cc.ntt(d, limbs)
        got = to_host_u64(d)
        for r in limbs:
            assert np.array_equal(got[r], o.ntt(r, x[r]))


def test_secret_key_golden_through_gpu():
    """The real OpenFHE secret keys: NTT(ternary) on the GPU must hash to the digests of the serialized limbs."""
    import hashlib
    from ppqsflhe_b200 import to_device_u64, to_host_u64
    o, cc = _cc_pair("c1")
    g = np.load(os.path.join(ROOT, "tests", "golden", "openfhe_sk_golden.npz"))
    for c in (1, 2):
        tern = g[f"c{c}_ternary"].astype(np.int64)
        x = np.stack([np.where(tern < 0, tern + int(o.moduli[i]), tern).astype(np.uint64) for i in range(4)])
        d = to_device_u64(x, cc.device)
        cc.ntt(d, [0, 1, 2, 3])
        got = to_host_u64(d)
        for i in range(4):
            assert hashlib.sha256(got[i].astype("<u8").tobytes()).hexdigest() == str(g[f"c{c}_sha256"][i])
            assert np.array_equal(got[i][:8], g[f"c{c}_head"][i])


@pytest.mark.parametrize("kind,ell", [("c1", 4), ("c1", 3), ("c1", 2), ("c1", 1), ((12, 5, 3, 45), 7), ((12, 5, 3, 45), 4),
                                      ((13, 4, 2, 50), 6), ((15, 2, 2, 40), 4)])
def test_modup_bit_exact(kind, ell):
    from ppqsflhe_b200 import to_device_u64, to_host_u64
    o, cc = _cc_pair(kind)
    rng = np.random.default_rng(7)
    n = 3
    c1 = np.stack([_rand_poly(rng, o, range(ell)) for _ in range(n)])
    got = to_host_u64(cc.KeySwitchModUp(to_device_u64(c1, cc.device)))
    for g in range(n):
        assert np.array_equal(got[g], o.modup(ell, c1[g])), f"ModUp differs (ct {g})"


@pytest.mark.parametrize("kind,ell,n", [("c1", 4, 5), ("c1", 3, 2), ("c1", 1, 1), ((12, 5, 3, 45), 7, 3), ((12, 12, 7, 45), 14, 2),
                                        ((13, 14, 8, 45), 16, 2),
                                        ((12, 5, 3, 45), 5, 2), ((13, 4, 2, 50), 6, 2), ((15, 2, 2, 40), 4, 2),
                                        ((16, 6, 3, 59), 8, 2), ((17, 2, 2, 40), 4, 1),
                                        # BASELINE.json full sizes: C3 (N=2^16, 20 limbs, K=7, digits 7/7/6), its fan-out
                                        # level ell-1 (digits 7/7/5), C4 (N=2^17, 40 limbs, K=14, digits 14/14/12)
                                        ((16, 18, 3, 59), 20, 2), ((16, 18, 3, 59), 19, 1), ((17, 38, 3, 59), 40, 1)])
def test_reencrypt_bit_exact(kind, ell, n):
    from ppqsflhe_b200 import CiphertextBatch, CtMeta, EvalKey, to_device_u64
    o, cc = _cc_pair(kind)
    rng = np.random.default_rng(11)
    ct = _rand_ct(rng, o, ell, n)
    kb, ka = _rand_key(rng, o)
    rk = EvalKey(cc, kb, ka)
    out = cc.ReEncrypt(CiphertextBatch(to_device_u64(ct, cc.device), CtMeta()), rk).numpy()
    for g in range(n):
        want = o.reencrypt(ell, ct[g], kb, ka)
        assert np.array_equal(out[g][1], want[1]), f"c1' differs (ct {g})"
        assert np.array_equal(out[g][0], want[0]), f"c0' differs (ct {g})"


def test_rekey_upload_and_borrowed_key_agree():
    """ppq_rekey_upload (host buffers, the handle owns its copy) and ppq_rekey_from_device (borrowed device buffers,
    the handle keeps a private P^-1-scaled copy of the Q limbs) must give the same ciphertexts, and the borrowed
    buffers must stay untouched (reference: the rekey is only read by cc->ReEncrypt, server/src/changeCipherDomain.cpp:40-44,74,89,105)."""
    from ppqsflhe_b200 import CiphertextBatch, CtMeta, EvalKey, to_device_u64
    o, cc = _cc_pair((13, 4, 2, 50))
    rng = np.random.default_rng(29)
    ct = _rand_ct(rng, o, o.L, 2)
    kb, ka = _rand_key(rng, o)
    kb, ka = np.ascontiguousarray(kb), np.ascontiguousarray(ka)
    rk = EvalKey(cc, kb, ka)
    batch = CiphertextBatch(to_device_u64(ct, cc.device), CtMeta())
    out_borrowed = cc.ReEncrypt(batch, rk).numpy()
    assert np.array_equal(rk.b.cpu().numpy().view(np.uint64), kb), "borrowed key B was modified"
    assert np.array_equal(rk.a.cpu().numpy().view(np.uint64), ka), "borrowed key A was modified"
    kb0, ka0 = kb.copy(), ka.copy()
    up = EvalKey.upload(cc, kb, ka)
    assert np.array_equal(kb, kb0) and np.array_equal(ka, ka0), "host key buffers were modified"
    out_upload = cc.ReEncrypt(batch, up).numpy()
    assert np.array_equal(out_upload, out_borrowed)
    assert np.array_equal(out_upload[0], o.reencrypt(o.L, ct[0], kb, ka))


def test_full_size_round_trip_c3():
    """Size-independent property at BASELINE.json's C3 size (N=2^16, 20 limbs, dnum 3): Decrypt_{sk2}(ReEncrypt_{1->2}(
    Enc_{pk1}(m))) = m within the CKKS tolerance, and ReEncrypt+EvalAdd of two clients decrypts to the sum."""
    from ppqsflhe_b200 import CiphertextBatch, CtMeta, EvalKey, to_device_u64
    o, cc = _cc_pair((16, 18, 3, 59))
    dev = cc.device
    rng = np.random.default_rng(3)
    keys = {c: o.keygen(70 + c) for c in (1, 2, 3)}          # (s, sk_eval, pk_b, pk_a)
    rk = {c: EvalKey(cc, *o.rekeygen(80 + c, keys[c][1], keys[3][2], keys[3][3])) for c in (1, 2)}
    m = {c: np.clip(rng.normal(0, 0.05, (2, o.N // 2)), -0.3, 0.3) for c in (1, 2)}
    acc = CiphertextBatch(torch.zeros((2, 2, o.L, o.N), dtype=torch.int64, device=dev), CtMeta())
    for c in (1, 2):
        pt = cc.MakeCKKSPackedPlaintext(m[c])
        ct = cc.Encrypt((to_device_u64(keys[c][2], dev), to_device_u64(keys[c][3], dev)), pt, seed=5 + c)
        if c == 1:
            one = cc.ReEncrypt(ct, rk[1])
            dec = cc.Decrypt(to_device_u64(keys[3][1], dev), one).GetRealPackedValue()
            assert np.max(np.abs(dec - m[1])) < 2.0 ** -20
            acc.meta = ct.meta
        cc.ReEncryptAccumulate(ct, rk[c], acc)
    dec = cc.Decrypt(to_device_u64(keys[3][1], dev), acc).GetRealPackedValue()
    assert np.max(np.abs(dec - (m[1] + m[2]))) < 2.0 ** -19


def test_reencrypt_wave_size_independent():
    from ppqsflhe_b200 import CiphertextBatch, CtMeta, EvalKey, to_device_u64
    o, cc = _cc_pair((12, 3, 2, 45))
    rng = np.random.default_rng(5)
    ell, n = o.L, 7
    ct = _rand_ct(rng, o, ell, n)
    kb, ka = _rand_key(rng, o)
    rk = EvalKey(cc, kb, ka)
    b = CiphertextBatch(to_device_u64(ct, cc.device), CtMeta())
    ref = cc.ReEncrypt(b, rk).numpy()
    for wave in (1, 2, 3, 7):
        cc.set_wave(wave)
        assert np.array_equal(cc.ReEncrypt(b, rk).numpy(), ref)
    cc.set_wave(0)


def test_reencrypt_accumulate_matches_oracle_batch():
    from ppqsflhe_b200 import CiphertextBatch, CtMeta, EvalKey, to_device_u64
    o, cc = _cc_pair("c1")
    rng = np.random.default_rng(13)
    ell, n_clients, n_slots = 4, 3, 4
    cts = np.stack([_rand_ct(rng, o, ell, n_slots) for _ in range(n_clients)])
    keys = [_rand_key(rng, o) for _ in range(n_clients)]
    acc = CiphertextBatch(torch.zeros((n_slots, 2, ell, cc.N), dtype=torch.int64, device=cc.device), CtMeta())
    for c in range(n_clients):
        rk = EvalKey(cc, *keys[c])
        cc.ReEncryptAccumulate(CiphertextBatch(to_device_u64(cts[c], cc.device), CtMeta()), rk, acc)
    want = o.reencrypt_accumulate_batch(ell, cts, [k[0] for k in keys], [k[1] for k in keys])
    assert np.array_equal(acc.numpy(), want)


@pytest.mark.parametrize("kind,ell,n_clients,n_slots", [("c1", 4, 2, 3), ("c1", 3, 9, 2), ("c1", 1, 3, 1),
                                                        ((12, 5, 3, 45), 7, 17, 2), ((13, 4, 2, 50), 5, 8, 3),
                                                        ((13, 38, 3, 59), 40, 9, 1), ((15, 6, 3, 50), 8, 3, 5), ((12, 12, 7, 45), 14, 5, 2),
                                                        ((16, 18, 3, 59), 20, 9, 2), ((16, 18, 3, 59), 19, 2, 3)])
def test_aggregator_bit_exact(kind, ell, n_clients, n_slots):
    """ppq_agg_* (deferred ModDown: the linear tail of ApproxModDown applied to sums over clients) against the oracle's
    per-client ReEncrypt followed by EvalAdd, client by client -- the reference's order of operations
    (server/src/changeCipherDomain.cpp:105, server/src/aggregateEncryptedWeights.cpp:106).  Client counts cross the
    group-of-8 boundary of the wide conversion; the handle is reused for a second aggregation."""
    from ppqsflhe_b200 import Aggregator, CiphertextBatch, CtMeta, EvalKey, to_device_u64
    o, cc = _cc_pair(kind)
    rng = np.random.default_rng(1000 + n_clients * 7 + ell)
    cts = np.stack([_rand_ct(rng, o, ell, n_slots) for _ in range(n_clients)])
    keys = [_rand_key(rng, o) for _ in range(n_clients)]
    rks = [EvalKey(cc, *k) for k in keys]
    agg = Aggregator(cc, ell, n_slots)
    for c in range(n_clients):
        agg.add(CiphertextBatch(to_device_u64(cts[c], cc.device), CtMeta()), rks[c])
    assert agg.count == n_clients
    got = agg.finish().numpy()
    want = o.reencrypt_accumulate_batch(ell, cts, [k[0] for k in keys], [k[1] for k in keys])
    assert np.array_equal(got, want), "aggregate differs from per-client ReEncrypt + EvalAdd"
    # second use of the same handle, fewer clients, reversed order (the sum commutes; state must have been reset)
    assert agg.count == 0
    sub = list(range(min(n_clients, 3)))[::-1]
    for c in sub:
        agg.add(CiphertextBatch(to_device_u64(cts[c], cc.device), CtMeta()), rks[c])
    got2 = agg.finish().numpy()
    want2 = o.reencrypt_accumulate_batch(ell, cts[sub], [keys[c][0] for c in sub], [keys[c][1] for c in sub])
    assert np.array_equal(got2, want2)
    agg.close()


@pytest.mark.parametrize("group,wave", [(1, 0), (3, 1), (8, 2)])
def test_aggregator_group_and_wave_independent(group, wave, monkeypatch):
    from ppqsflhe_b200 import Aggregator, CiphertextBatch, CtMeta, EvalKey, to_device_u64
    monkeypatch.setenv("PPQ_AGG_GROUP", str(group))
    o, cc = _cc_pair((13, 4, 2, 50))
    cc.set_wave(wave)
    rng = np.random.default_rng(77)
    ell, n_clients, n_slots = o.L, 7, 3
    cts = np.stack([_rand_ct(rng, o, ell, n_slots) for _ in range(n_clients)])
    keys = [_rand_key(rng, o) for _ in range(n_clients)]
    agg = Aggregator(cc, ell, n_slots)
    for c in range(n_clients):
        agg.add(CiphertextBatch(to_device_u64(cts[c], cc.device), CtMeta()), EvalKey(cc, *keys[c]))
    want = o.reencrypt_accumulate_batch(ell, cts, [k[0] for k in keys], [k[1] for k in keys])
    assert np.array_equal(agg.finish().numpy(), want)
    cc.set_wave(0)


@pytest.mark.parametrize("kind,ell,n_clients,n_slots,nb", [("c1", 4, 5, 3, 2), ("c1", 3, 9, 2, 4), ((12, 5, 3, 45), 7, 11, 2, 4),
                                                           ((13, 38, 3, 59), 40, 6, 1, 2), ((16, 18, 3, 59), 20, 7, 2, 2),
                                                           ((16, 18, 3, 59), 20, 9, 2, 4)])
def test_aggregator_add_many_bit_exact(kind, ell, n_clients, n_slots, nb, monkeypatch):
    """ppq_agg_add_many (k_ip_agg: nb clients per inner-product pass, slot sums touched once per pass) against the oracle's
    per-client ReEncrypt + EvalAdd, with client counts that leave odd remainders and cross the group-of-8 boundary."""
    from ppqsflhe_b200 import Aggregator, CiphertextBatch, CtMeta, EvalKey, to_device_u64
    monkeypatch.setenv("PPQ_AGG_NB", str(nb))
    o, cc = _cc_pair(kind)
    rng = np.random.default_rng(2000 + n_clients * 7 + ell)
    cts = np.stack([_rand_ct(rng, o, ell, n_slots) for _ in range(n_clients)])
    keys = [_rand_key(rng, o) for _ in range(n_clients)]
    rks = [EvalKey(cc, *k) if i % 2 else EvalKey.upload(cc, *k) for i, k in enumerate(keys)]  # both key layouts
    agg = Aggregator(cc, ell, n_slots)
    batches = [CiphertextBatch(to_device_u64(cts[c], cc.device), CtMeta()) for c in range(n_clients)]
    agg.add_many(batches[:1], rks[:1])
    agg.add_many(batches[1:], rks[1:])
    assert agg.count == n_clients
    want = o.reencrypt_accumulate_batch(ell, cts, [k[0] for k in keys], [k[1] for k in keys])
    assert np.array_equal(agg.finish().numpy(), want)


def test_aggregator_extreme_residues():
    """All-(q-1) and all-zero inputs and keys: the largest operands every lazy bound has to hold."""
    from ppqsflhe_b200 import Aggregator, CiphertextBatch, CtMeta, EvalKey, to_device_u64
    for kind in ("c1", (12, 5, 3, 45)):
        o, cc = _cc_pair(kind)
        ell, n_clients, n_slots = o.L, 9, 1
        q = o.moduli
        top_ct = np.stack([np.stack([np.stack([np.full(o.N, q[i] - np.uint64(1), dtype=np.uint64) for i in range(ell)])
                                     for _ in range(2)]) for _ in range(n_slots)])
        top_k = np.stack([np.stack([np.full(o.N, q[r] - np.uint64(1), dtype=np.uint64) for r in range(o.M)])
                          for _ in range(o.dnum)])
        rng = np.random.default_rng(5)
        cts, keys = [], []
        for c in range(n_clients):
            if c % 3 == 0:
                cts.append(top_ct)
                keys.append((top_k, top_k))
            elif c % 3 == 1:
                cts.append(np.zeros_like(top_ct))
                keys.append(_rand_key(rng, o))
            else:
                cts.append(_rand_ct(rng, o, ell, n_slots))
                keys.append((top_k, np.zeros_like(top_k)))
        cts = np.stack(cts)
        agg = Aggregator(cc, ell, n_slots)
        for c in range(n_clients):
            agg.add(CiphertextBatch(to_device_u64(cts[c], cc.device), CtMeta()), EvalKey(cc, *keys[c]))
        want = o.reencrypt_accumulate_batch(ell, cts, [k[0] for k in keys], [k[1] for k in keys])
        assert np.array_equal(agg.finish().numpy(), want)
        rks = [EvalKey(cc, *keys[c]) for c in range(n_clients)]  # the several-clients-per-pass form on the same extremes
        agg.add_many([CiphertextBatch(to_device_u64(cts[c], cc.device), CtMeta()) for c in range(n_clients)], rks)
        assert np.array_equal(agg.finish().numpy(), want)


def test_aggregator_equals_sequential_gpu_path_full_c3_size():
    """At BASELINE.json's C3 shape (N=2^16, 20 limbs, 54 slots) the oracle is too slow for a whole client, so the
    size-independent property is checked instead: ppq_agg_* == the per-client ppq_reencrypt_accumulate chain (itself
    oracle-checked at small sizes), bit for bit, over 10 clients (one full group of 8 + a partial one)."""
    from ppqsflhe_b200 import Aggregator, CiphertextBatch, CtMeta, EvalKey
    _, cc = _cc_pair((16, 18, 3, 59))
    dev, ell, n_slots, n_clients = cc.device, cc.L, 54, 10
    qv = torch.from_numpy(cc.moduli.astype(np.int64)).to(dev)
    torch.manual_seed(9)

    def rr(shape, q):
        return torch.randint(0, 1 << 62, shape, dtype=torch.int64, device=dev).remainder_(q)

    acc = CiphertextBatch(torch.zeros((n_slots, 2, ell, cc.N), dtype=torch.int64, device=dev), CtMeta())
    agg = Aggregator(cc, ell, n_slots)
    for _ in range(n_clients):
        ct = CiphertextBatch(rr((n_slots, 2, ell, cc.N), qv[:ell].view(1, 1, ell, 1)), CtMeta())
        rk = EvalKey(cc, rr((cc.dnum, cc.M, cc.N), qv.view(1, cc.M, 1)), rr((cc.dnum, cc.M, cc.N), qv.view(1, cc.M, 1)))
        cc.ReEncryptAccumulate(ct, rk, acc)
        agg.add(ct, rk)
        del rk
    out = agg.finish()
    assert torch.equal(out.data, acc.data)


def test_host_buffer_aggregate_entry_point():
    import ctypes as C
    from ppqsflhe_b200 import EvalKey
    from ppqsflhe_b200._capi import check, lib, vp
    o, cc = _cc_pair("c1")
    rng = np.random.default_rng(17)
    ell, n_clients, n_slots = 3, 4, 2
    cts = np.ascontiguousarray(np.stack([_rand_ct(rng, o, ell, n_slots) for _ in range(n_clients)]))
    keys = [_rand_key(rng, o) for _ in range(n_clients)]
    rks = [EvalKey(cc, *k) for k in keys]
    arr = (vp * n_clients)(*[r.h for r in rks])
    ptrs = (vp * n_clients)(*[cts[c].ctypes.data for c in range(n_clients)])
    out = np.zeros((n_slots, 2, ell, o.N), dtype=np.uint64)
    check(lib().ppq_reencrypt_aggregate_host(cc.h, arr, ell, ptrs, n_clients, n_slots, out.ctypes.data_as(vp)))
    want = o.reencrypt_accumulate_batch(ell, cts, [k[0] for k in keys], [k[1] for k in keys])
    assert np.array_equal(out, want)


def _pack_ref(ct, widths):
    """packed residues in Python integers: limb i = little-endian bit stream of N residues of widths[i] bits"""
    words = []
    for poly in ct:
        for i, limb in enumerate(poly):
            w = widths[i]
            big = 0
            for k, r in enumerate(limb.tolist()):
                big |= int(r) << (k * w)
            nw = len(limb) * w // 64
            words += [(big >> (64 * m)) & (2**64 - 1) for m in range(nw)]
    return np.array(words, dtype=np.uint64)


@pytest.mark.parametrize("kind,ell", [("c1", 4), ("c1", 3), ((12, 5, 3, 45), 7), ((13, 3, 2, 30), 4)])
def test_packed_residues_round_trip(kind, ell):
    # the wire format of the *_host_packed calls (include/ppq_b200.h): device and host packers against a Python-integer
    # restatement, and both unpackers back to the input; extreme residues (0 and q-1) included
    import torch
    from ppqsflhe_b200._capi import check, lib, vp
    o, cc = _cc_pair(kind)
    rng = np.random.default_rng(101)
    n = 3
    ct = _rand_ct(rng, o, ell, n)
    for i in range(ell):
        ct[0, 0, i, :] = int(o.moduli[i]) - 1
        ct[0, 1, i, ::2] = 0
        ct[1, 0, i, -1] = int(o.moduli[i]) - 1
    widths = [int(o.moduli[i]).bit_length() for i in range(ell)]
    pw = lib().ppq_packed_words(cc.h, ell)
    assert pw == 2 * sum(o.N * w // 64 for w in widths)
    want = np.stack([_pack_ref(ct[g], widths) for g in range(n)])
    hp = np.zeros((n, pw), dtype=np.uint64)
    check(lib().ppq_pack_host(cc.h, ell, ct.ctypes.data_as(vp), hp.ctypes.data_as(vp), n))
    assert np.array_equal(hp, want)
    d_ct = torch.from_numpy(ct.view(np.int64)).to("cuda:0")
    d_pk = torch.zeros((n, pw), dtype=torch.int64, device="cuda:0")
    check(lib().ppq_pack(cc.h, ell, vp(d_ct.data_ptr()), vp(d_pk.data_ptr()), n, None))
    torch.cuda.synchronize()
    assert np.array_equal(d_pk.cpu().numpy().view(np.uint64), want)
    back = np.zeros_like(ct)
    check(lib().ppq_unpack_host(cc.h, ell, hp.ctypes.data_as(vp), back.ctypes.data_as(vp), n))
    assert np.array_equal(back, ct)
    d_back = torch.zeros_like(d_ct)
    check(lib().ppq_unpack(cc.h, ell, vp(d_pk.data_ptr()), vp(d_back.data_ptr()), n, None))
    torch.cuda.synchronize()
    assert np.array_equal(d_back.cpu().numpy().view(np.uint64), ct)


def test_host_buffer_aggregate_packed_entry_point():
    from ppqsflhe_b200 import EvalKey
    from ppqsflhe_b200._capi import check, lib, vp
    o, cc = _cc_pair("c1")
    rng = np.random.default_rng(18)
    ell, n_clients, n_slots = 3, 5, 2
    cts = np.ascontiguousarray(np.stack([_rand_ct(rng, o, ell, n_slots) for _ in range(n_clients)]))
    keys = [_rand_key(rng, o) for _ in range(n_clients)]
    rks = [EvalKey(cc, *k) for k in keys]
    pw = lib().ppq_packed_words(cc.h, ell)
    packed = np.zeros((n_clients, n_slots, pw), dtype=np.uint64)
    for c in range(n_clients):
        check(lib().ppq_pack_host(cc.h, ell, cts[c].ctypes.data_as(vp), packed[c].ctypes.data_as(vp), n_slots))
    arr = (vp * n_clients)(*[r.h for r in rks])
    ptrs = (vp * n_clients)(*[packed[c].ctypes.data for c in range(n_clients)])
    out_p = np.zeros((n_slots, pw), dtype=np.uint64)
    check(lib().ppq_reencrypt_aggregate_host_packed(cc.h, arr, ell, ptrs, n_clients, n_slots, out_p.ctypes.data_as(vp)))
    out = np.zeros((n_slots, 2, ell, o.N), dtype=np.uint64)
    check(lib().ppq_unpack_host(cc.h, ell, out_p.ctypes.data_as(vp), out.ctypes.data_as(vp), n_slots))
    want = o.reencrypt_accumulate_batch(ell, cts, [k[0] for k in keys], [k[1] for k in keys])
    assert np.array_equal(out, want)
    # the copy-only leg moves the same bytes and must leave the library usable
    check(lib().ppq_copy_ceiling_host_packed(cc.h, ell, ptrs, n_clients, n_slots, out_p.ctypes.data_as(vp)))
    check(lib().ppq_reencrypt_aggregate_host_packed(cc.h, arr, ell, ptrs, n_clients, n_slots, out_p.ctypes.data_as(vp)))
    check(lib().ppq_unpack_host(cc.h, ell, out_p.ctypes.data_as(vp), out.ctypes.data_as(vp), n_slots))
    assert np.array_equal(out, want)


def test_eval_add_and_finalize_bit_exact():
    from ppqsflhe_b200 import CiphertextBatch, CtMeta, to_device_u64
    o, cc = _cc_pair("c1")
    rng = np.random.default_rng(19)
    ell, n = 4, 3
    a, b = _rand_ct(rng, o, ell, n), _rand_ct(rng, o, ell, n)
    A = CiphertextBatch(to_device_u64(a, cc.device), CtMeta())
    B = CiphertextBatch(to_device_u64(b, cc.device), CtMeta())
    got = cc.EvalAdd(A, B).numpy()
    for g in range(n):
        assert np.array_equal(got[g], o.eval_add(ell, a[g], b[g]))
    # lazy u64 sum of 8 canonical operands (what the NCCL reduce produces) then finalize
    parts = [_rand_ct(rng, o, ell, n) for _ in range(8)]
    lazy = np.zeros_like(parts[0])
    want = np.zeros_like(parts[0])
    for p in parts:
        lazy = lazy + p
        for g in range(n):
            want[g] = o.eval_add(ell, want[g], p[g])
    L = cc.finalize_sum(CiphertextBatch(to_device_u64(lazy, cc.device), CtMeta()))
    assert np.array_equal(L.numpy(), want)


@pytest.mark.parametrize("kind,ell,c", [("c1", 4, 0.5), ("c1", 3, 0.25), ((12, 5, 3, 45), 7, 1.0 / 3.0),
                                        ((16, 6, 3, 59), 8, 1.0 / 64.0), ((16, 18, 3, 59), 20, 1.0 / 64.0),
                                        ((16, 18, 3, 59), 19, 0.5), ((17, 38, 3, 59), 40, 1.0 / 16.0)])
def test_eval_mult_rescale_bit_exact(kind, ell, c):
    from ppqsflhe_b200 import CiphertextBatch, CtMeta, to_device_u64
    o, cc = _cc_pair(kind)
    rng = np.random.default_rng(23)
    n = 2
    ct = _rand_ct(rng, o, ell, n)
    level = o.L - ell
    meta = CtMeta(noise_scale_deg=2, level=level, scaling_factor=o.sf_big(level) if level < o.L - 1 else 1.0)
    got = cc.EvalMult(CiphertextBatch(to_device_u64(ct, cc.device), meta), c)
    for g in range(n):
        want, wmeta = o.eval_mult_const(ell, ct[g], dict(noise_scale_deg=2, level=level,
                                                         scaling_factor=meta.scaling_factor), c)
        assert np.array_equal(got.numpy()[g], want)
    assert got.meta.level == wmeta["level"] and got.meta.noise_scale_deg == wmeta["noise_scale_deg"]
    assert got.meta.scaling_factor == wmeta["scaling_factor"]
    # multiply only (noiseScaleDeg 1)
    meta1 = CtMeta(noise_scale_deg=1, level=level, scaling_factor=o.sf(level))
    got1 = cc.EvalMult(CiphertextBatch(to_device_u64(ct, cc.device), meta1), c)
    for g in range(n):
        want1, _ = o.eval_mult_const(ell, ct[g], dict(noise_scale_deg=1, level=level, scaling_factor=o.sf(level)), c)
        assert np.array_equal(got1.numpy()[g], want1)


def test_encrypt_bit_exact_with_injected_samples_and_decrypt_core():
    from ppqsflhe_b200 import PlaintextBatch, CtMeta, to_device_u64, to_host_u64
    o, cc = _cc_pair("c1")
    ell, n = 4, 3
    s, sk, pkb, pka = o.keygen(31)
    rng = np.random.default_rng(29)
    pt = np.stack([_rand_poly(rng, o, range(ell)) for _ in range(n)])
    small = np.stack([np.stack([oracle.sample_ternary(100 + g, o.N), oracle.sample_gauss(200 + g, o.N),
                                oracle.sample_gauss(300 + g, o.N)]) for g in range(n)])
    P = PlaintextBatch(data=to_device_u64(pt, cc.device), meta=CtMeta())
    pk = (to_device_u64(pkb, cc.device), to_device_u64(pka, cc.device))
    ct = cc.Encrypt(pk, P, small=torch.from_numpy(small).to(cc.device))
    got = ct.numpy()
    for g in range(n):
        want = o.encrypt(ell, pkb, pka, pt[g], small[g][0], small[g][1], small[g][2])
        assert np.array_equal(got[g], want)
    co = to_host_u64(cc.DecryptCore(to_device_u64(sk, cc.device), ct))
    for g in range(n):
        assert np.array_equal(co[g], o.decrypt_core(ell, sk, got[g]))


@pytest.mark.parametrize("kind,ell", [("c1", 4), ("c1", 2), ((13, 3, 2, 40), 4), ((16, 18, 3, 59), 20)])
def test_encode_encrypt_fused_equals_the_two_calls(kind, ell):
    # ppq_encode_encrypt adds the encoded coefficients to e0 before the transform (3*ell limb transforms instead of
    # 4*ell): bit-identical to MakeCKKSPackedPlaintext followed by Encrypt with the same samples, which is the pair the
    # oracle pins (test_encrypt_bit_exact_with_injected_samples_and_decrypt_core, test_encode_decode_tolerance)
    o, cc = _cc_pair(kind)
    rng = np.random.default_rng(41)
    n = 3
    vals = np.clip(rng.normal(0, 0.5, (n, o.N // 2)), -3.0, 3.0)
    vals[1, 7:] = 0.0
    small = np.stack([np.stack([oracle.sample_ternary(100 + g, o.N), oracle.sample_gauss(200 + g, o.N),
                                oracle.sample_gauss(300 + g, o.N)]) for g in range(n)])
    d_small = torch.from_numpy(small).to(cc.device)
    qv = torch.from_numpy(cc.moduli.astype(np.int64)).to(cc.device)
    pk = tuple(torch.randint(0, 1 << 62, (cc.M, cc.N), dtype=torch.int64, device=cc.device).remainder_(qv.view(-1, 1))
               for _ in range(2))
    two = cc.Encrypt(pk, cc.MakeCKKSPackedPlaintext(vals, ell=ell), small=d_small)
    one = cc.EncodeEncrypt(pk, vals, ell=ell, small=d_small)
    assert torch.equal(one.data, two.data)
    assert one.meta.scaling_factor == two.meta.scaling_factor and one.meta.noise_scale_deg == 2


def test_encode_decode_tolerance():
    """CKKS encode is not bit-reproducible: compare through decode.  Tolerance 2^-20 (north_star);
    observed ~1e-9."""
    from ppqsflhe_b200 import to_host_u64, to_device_u64
    o, cc = _cc_pair("c1")
    rng = np.random.default_rng(37)
    n = 3
    vals = np.clip(rng.normal(0, 0.05, (n, o.N // 2)), -0.3, 0.3)
    vals[1, 5:] = 0.0  # a 5-value plaintext zero-padded to all slots, like `mean` / `std_dev`
    P = cc.MakeCKKSPackedPlaintext(vals)
    pt = to_host_u64(P.data)
    sf = o.sf_big(0)
    for g in range(n):
        # (a) the GPU-encoded polynomial decodes (oracle decode) to the input
        co = np.stack([o.ntt(i, pt[g][i], inverse=True) for i in range(o.L)])
        back = o.decode(o.L, co, sf)
        assert np.max(np.abs(back - vals[g])) < 2.0 ** -20
        # (b) coefficients agree with the oracle's encoding up to FFT rounding (|diff| tiny relative to sf)
        want = o.encode(o.L, vals[g], sf)
        wco = o.ntt(0, want[0], inverse=True).astype(np.int64)
        gco = co[0].astype(np.int64)
        q0 = int(o.moduli[0])
        diff = (gco - wco + q0 // 2) % q0 - q0 // 2
        assert np.max(np.abs(diff)) <= 2 ** 12
    # GPU decode of oracle-encoded data (4 limbs, scale sfBig[0] ~ 2^59; a 2^80 scale would overflow the
    # int64 coefficients of Encode, in OpenFHE as in the oracle)
    enc = np.stack([o.encode(4, vals[g], sf) for g in range(n)])
    co = np.stack([[o.ntt(i, enc[g][i], inverse=True) for i in range(4)] for g in range(n)])
    out = torch.empty((n, o.N // 2), dtype=torch.float64, device=cc.device)
    from ppqsflhe_b200._capi import check, lib, vp
    d = to_device_u64(co, cc.device)
    check(lib().ppq_decode(cc.h, 4, vp(d.data_ptr()), sf, vp(out.data_ptr()), n, None))
    torch.cuda.synchronize()
    assert np.max(np.abs(out.cpu().numpy() - vals)) < 2.0 ** -20
    for g in range(n):
        assert np.max(np.abs(out.cpu().numpy()[g] - o.decode(4, co[g], sf))) < 1e-12


def test_end_to_end_golden_round():
    """The reference's recorded round: Enc(w1) under pk1 -> ReEncrypt to pk2 -> EvalAdd Enc(w2) -> EvalMult 0.5
    -> Decrypt with sk2 must reproduce decrypted_weights_c2.json ((w1+w2)/2) within the CKKS tolerance 2^-20;
    and the fan-out ReEncrypt(2->1) must decrypt under sk1 (decrypted_weights_c1.json)."""
    from ppqsflhe_b200 import EvalKey, to_device_u64
    o, cc = _cc_pair("c1")
    g = np.load(os.path.join(ROOT, "tests", "golden", "weights_golden.npz"))
    w1, w2 = g["c1_sample_values"].astype(np.float64), g["c2_sample_values"].astype(np.float64)
    slots = o.N // 2
    nct = (len(w1) + slots - 1) // slots

    def chunks(w):
        p = np.zeros(nct * slots)
        p[: len(w)] = w
        return p.reshape(nct, slots)

    keys = {}
    for c in (1, 2):
        tern = np.load(os.path.join(ROOT, "tests", "golden", "openfhe_sk_golden.npz"))[f"c{c}_ternary"].astype(np.int64)
        # scaffolding keygen from the oracle, but with the REAL OpenFHE secret of client c
        s, sk, pkb, pka = o.keygen(40 + c)
        sk = np.stack([o.small_to_eval(r, tern) for r in range(o.M)])
        e = oracle.sample_gauss(900 + c, o.N)
        pkb = np.stack([(o.small_to_eval(r, e).astype(object) - pka[r].astype(object) * sk[r].astype(object)) % int(o.moduli[r])
                        for r in range(o.M)]).astype(np.uint64)
        keys[c] = (sk, pkb, pka)
    rk12 = EvalKey(cc, *o.rekeygen(51, keys[1][0], keys[2][1], keys[2][2]))
    rk21 = EvalKey(cc, *o.rekeygen(52, keys[2][0], keys[1][1], keys[1][2]))
    dev = cc.device
    cts = {}
    for c, w in ((1, w1), (2, w2)):
        pt = cc.MakeCKKSPackedPlaintext(chunks(w))
        cts[c] = cc.Encrypt((to_device_u64(keys[c][1], dev), to_device_u64(keys[c][2], dev)), pt, seed=1000 + c)
    c12 = cc.ReEncrypt(cts[1], rk12)
    agg = cc.EvalMult(cc.EvalAdd(cts[2], c12), 0.5)
    assert agg.ell == 3 and agg.meta.level == 1 and agg.meta.noise_scale_deg == 2
    assert abs(agg.meta.scaling_factor / 1.2089e24 - 1) < 1e-3
    dec2 = cc.Decrypt(to_device_u64(keys[2][0], dev), agg).GetRealPackedValue().reshape(-1)[: len(w1)]
    want = (w1 + w2) / 2
    assert np.max(np.abs(dec2 - want)) < 2.0 ** -20
    assert np.max(np.abs(dec2 - g["c2_decrypted_values"])) < 2.0 ** -20
    back = cc.ReEncrypt(agg, rk21)
    dec1 = cc.Decrypt(to_device_u64(keys[1][0], dev), back).GetRealPackedValue().reshape(-1)[: len(w1)]
    assert np.max(np.abs(dec1 - g["c1_decrypted_values"])) < 2.0 ** -20
    print("e2e max abs err:", np.max(np.abs(dec2 - want)), np.max(np.abs(dec1 - want)))


# ---- KeyGen / ReKeyGen on the device
def _keygen_samples(o, seed):
    """the oracle's own draws for keygen(seed) (oracle/ppq_oracle.c ppqo_keygen), to inject into the device path"""
    e = oracle.sample_gauss(seed * 1000003 + 2, o.N)
    a = np.stack([oracle.sample_uniform(seed * 1000003 + 100 + r, int(o.moduli[r]), o.N) for r in range(o.M)])
    return e, a


def _rekeygen_samples(o, seed):
    return np.stack([np.stack([oracle.sample_ternary(seed * 7919 + 10 * j + 1, o.N),
                               oracle.sample_gauss(seed * 7919 + 10 * j + 2, o.N),
                               oracle.sample_gauss(seed * 7919 + 10 * j + 3, o.N)]) for j in range(o.dnum)])


@pytest.mark.parametrize("kind", ["c1", (12, 5, 3, 45), (13, 4, 2, 50), (16, 18, 3, 59)])
def test_keygen_and_rekeygen_bit_exact_with_injected_samples(kind):
    """cc->KeyGen() (client/src/keyGen.cpp:33) and cc->ReKeyGen(sk, pk) (client/src/REkeyGen.cpp:52) through the C ABI
    against the oracle, the oracle's own sampled polynomials injected."""
    from ppqsflhe_b200 import to_device_u64, to_host_u64
    o, cc = _cc_pair(kind)
    dev = cc.device
    got = {}
    want = {}
    for seed in (61, 62):
        s, sk, pkb, pka = o.keygen(seed)
        e, a = _keygen_samples(o, seed)
        assert np.array_equal(a, pka)
        gsk, (gb, ga) = cc.KeyGen(samples=(torch.from_numpy(s).to(dev), torch.from_numpy(e).to(dev), to_device_u64(a, dev)))
        assert np.array_equal(to_host_u64(gsk), sk[: o.L]), "secret key differs"
        assert np.array_equal(to_host_u64(gb), pkb), "public key b differs"
        got[seed] = (gsk, (gb, ga))
        want[seed] = (sk, pkb, pka)
    # re-encryption key 61 -> 62
    kb, ka = o.rekeygen(71, want[61][0], want[62][1], want[62][2])
    small = _rekeygen_samples(o, 71)
    rk = cc.ReKeyGen(got[61][0], got[62][1], small=torch.from_numpy(small).to(dev))
    assert np.array_equal(to_host_u64(rk.b), kb), "rekey B differs"
    assert np.array_equal(to_host_u64(rk.a), ka), "rekey A differs"


@pytest.mark.parametrize("kind", ["c1", (15, 6, 3, 50)])
def test_device_generated_keys_round_trip(kind):
    """SURVEY 8(c) self-consistency KAT with nothing from the oracle on the data path: keys, re-encryption key and
    encryption noise all drawn on the device; Decrypt_{sk2}(ReEncrypt_{1->2}(Enc_{pk1}(m))) ~ m within 2^-20."""
    o, cc = _cc_pair(kind)
    sk1, pk1 = cc.KeyGen(seed=101)
    sk2, pk2 = cc.KeyGen(seed=202)
    # the draws look like what they should be
    s = cc.KeyGen(seed=101)[0]
    assert torch.equal(s, sk1), "KeyGen is not deterministic in its seed"
    q = [int(m) for m in o.moduli]
    a = pk1[1].cpu().numpy().view(np.uint64)
    for r in range(o.M):
        assert a[r].max() < q[r]
        assert abs(float(a[r].astype(np.float64).mean()) / q[r] - 0.5) < 0.02
    assert not torch.equal(pk1[1], pk2[1])
    rk = cc.ReKeyGen(sk1, pk2, seed=303)
    rng = np.random.default_rng(5)
    vals = np.clip(rng.normal(0, 0.05, (2, o.N // 2)), -0.3, 0.3)
    pt = cc.MakeCKKSPackedPlaintext(vals)
    ct1 = cc.Encrypt(pk1, pt, seed=404)
    back1 = cc.Decrypt(sk1, ct1).GetRealPackedValue()
    assert np.max(np.abs(back1 - vals)) < 2.0 ** -20
    ct2 = cc.ReEncrypt(ct1, rk)
    back2 = cc.Decrypt(sk2, ct2).GetRealPackedValue()
    assert np.max(np.abs(back2 - vals)) < 2.0 ** -20
    wrong = cc.Decrypt(sk1, ct2).GetRealPackedValue()
    assert np.max(np.abs(wrong - vals)) > 1.0, "the re-encrypted ciphertext still decrypts under the old key"


# ---- fan-out of one batch to several target domains
@pytest.mark.parametrize("kind,ell,n,n_keys", [("c1", 3, 3, 3), ("c1", 4, 2, 2), ((13, 4, 2, 50), 5, 2, 4),
                                               ((16, 18, 3, 59), 19, 2, 3)])
def test_reencrypt_fanout_bit_exact(kind, ell, n, n_keys, monkeypatch):
    """ppq_reencrypt_fanout (shared ModUp, per-key inner product + ModDown) against the oracle's ReEncrypt per key and
    against separate ppq_reencrypt calls; C1 at ell = 3 is the reference's own fan-out level (uneven digits {q0,q1},{q2})."""
    from ppqsflhe_b200 import CiphertextBatch, CtMeta, EvalKey, to_device_u64
    for _ in range(1):
        o, cc = _cc_pair(kind)
        rng = np.random.default_rng(41)
        ct = _rand_ct(rng, o, ell, n)
        keys = [_rand_key(rng, o) for _ in range(n_keys)]
        rks = [EvalKey(cc, kb, ka) for kb, ka in keys]
        batch = CiphertextBatch(to_device_u64(ct, cc.device), CtMeta())
        outs = cc.ReEncryptFanOut(batch, rks)
        assert len(outs) == n_keys
        for k in range(n_keys):
            got = outs[k].numpy()
            assert np.array_equal(got, cc.ReEncrypt(batch, rks[k]).numpy()), f"key {k}: differs from ppq_reencrypt"
            want = o.reencrypt(ell, ct[0], keys[k][0], keys[k][1])
            assert np.array_equal(got[0], want), f"key {k}: differs from the oracle"
        assert np.array_equal(batch.numpy(), ct), "input modified"
        # small waves exercise the per-wave reuse of the digit buffers
        cc.set_wave(1)
        outs1 = cc.ReEncryptFanOut(batch, rks)
        cc.set_wave(0)
        for k in range(n_keys):
            assert np.array_equal(outs1[k].numpy(), outs[k].numpy())
        cc.close()  # the keys outlive their context here: ppq_rekey_destroy must not need it
        del rks


@pytest.mark.parametrize("kind", [(15, 3, 2, 50), (16, 6, 3, 59)])
def test_paired_row_kernels_odd_waves(kind):
    """From N = 2^15 on the ROW passes take ciphertexts in pairs (k_fwd_row_plain2, k_fwd_row_moddown2, k_inv_row2), an odd
    one at the end of a CTA's range paired with itself: waves of 1, 2, 3, 5 and 7 ciphertexts must give the same
    ciphertexts, with and without the fused EvalAdd, and ciphertext 6 (the self-paired tail) must match the oracle."""
    from ppqsflhe_b200 import CiphertextBatch, CtMeta, EvalKey, to_device_u64
    o, cc = _cc_pair(kind)
    rng = np.random.default_rng(77)
    ell, n = o.L, 7
    ct = _rand_ct(rng, o, ell, n)
    kb, ka = _rand_key(rng, o)
    rk = EvalKey(cc, kb, ka)
    b = CiphertextBatch(to_device_u64(ct, cc.device), CtMeta())
    start = _rand_ct(rng, o, ell, n)
    ref = cc.ReEncrypt(b, rk).numpy()
    assert np.array_equal(ref[6], o.reencrypt(ell, ct[6], kb, ka))
    q = o.moduli[:ell].astype(object).reshape(1, 1, ell, 1)
    want_acc = ((ref.astype(object) + start.astype(object)) % q).astype(np.uint64)
    for wave in (1, 2, 3, 5, 7):
        cc.set_wave(wave)
        assert np.array_equal(cc.ReEncrypt(b, rk).numpy(), ref), f"wave {wave}"
        acc = CiphertextBatch(to_device_u64(start, cc.device), CtMeta())
        cc.ReEncryptAccumulate(b, rk, acc)
        assert np.array_equal(acc.numpy(), want_acc), f"wave {wave} (accumulate)"
    cc.set_wave(0)


def test_sampler_statistics_and_keying():
    """Device samplers (csrc/sampler.cu: ChaCha20 streams, ternary by rejection, discrete Gaussian sigma 3.19 by table
    inversion, uniform residues by rejection) -- distribution checks on 3 x 2^20 draws and keying behaviour."""
    from ppqsflhe_b200._capi import check, lib, vp
    _, cc = _cc_pair((16, 2, 2, 50))
    n = 16
    small = torch.empty((n, 3, cc.N), dtype=torch.int64, device=cc.device)
    seed = (12345).to_bytes(32, "little")
    check(lib().ppq_sample_encrypt_noise(cc.h, seed, vp(small.data_ptr()), n, None))
    torch.cuda.synchronize()
    x = small.cpu().numpy()
    tern, g = x[:, 0, :].ravel(), x[:, 1:, :].ravel()
    cnt = np.array([(tern == v).sum() for v in (-1, 0, 1)], dtype=np.float64)
    assert cnt.sum() == tern.size
    chi2 = ((cnt - tern.size / 3) ** 2 / (tern.size / 3)).sum()
    assert chi2 < 20.0, f"ternary not uniform: {cnt}"                      # 2 degrees of freedom
    sigma = 3.19
    ks = np.arange(-60, 61)
    pmf = np.exp(-ks.astype(np.float64) ** 2 / (2 * sigma * sigma))
    pmf /= pmf.sum()
    assert np.abs(g).max() <= 40 and abs(g.mean()) < 0.02
    assert abs(g.var() - (pmf * ks ** 2).sum()) < 0.08
    sel = np.abs(ks) <= 9
    obs = np.array([(g == k).sum() for k in ks[sel]], dtype=np.float64)
    exp = pmf[sel] * g.size
    assert ((obs - exp) ** 2 / exp).sum() < 60.0, "Gaussian histogram off"  # 18 degrees of freedom
    # every polynomial has its own stream; a fixed seed repeats; no seed = fresh OS entropy per call
    assert len({x[i, j, :64].tobytes() for i in range(n) for j in range(3)}) == 3 * n
    again = torch.empty_like(small)
    check(lib().ppq_sample_encrypt_noise(cc.h, seed, vp(again.data_ptr()), n, None))
    assert torch.equal(again, small)
    a1, a2 = torch.empty_like(small), torch.empty_like(small)
    check(lib().ppq_sample_encrypt_noise(cc.h, None, vp(a1.data_ptr()), n, None))
    check(lib().ppq_sample_encrypt_noise(cc.h, None, vp(a2.data_ptr()), n, None))
    assert not torch.equal(a1, a2) and not torch.equal(a1, small)
    # keygen: s ternary, e Gaussian, a uniform in [0, q_r) on every limb of QP; two unseeded key pairs differ
    sk1, (b1, pa1) = cc.KeyGen()
    sk2, (b2, pa2) = cc.KeyGen()
    assert not torch.equal(sk1, sk2) and not torch.equal(pa1, pa2)
    a = pa1.cpu().numpy().view(np.uint64)
    for r in range(cc.M):
        q = int(cc.moduli[r])
        assert int(a[r].max()) < q
        assert abs(float(a[r].astype(np.float64).mean()) / q - 0.5) < 0.01
        assert abs(float((a[r] > np.uint64(q // 2)).mean()) - 0.5) < 0.01


# ---------------------------------------------------------------------------------------------------------------------
# Stage views and big-integer properties: checks that do NOT go through the C oracle's arithmetic (Python integers only)
# ---------------------------------------------------------------------------------------------------------------------
def _crt_lift(res, mods):
    """exact lift in [0, prod mods) of per-limb residue arrays (object arrays of Python ints)"""
    import math
    Q = math.prod(mods)
    x = np.zeros(len(res[0]), dtype=object)
    for r, q in zip(res, mods):
        h = Q // q
        x = (x + np.array([int(v) for v in r], dtype=object) * (h * pow(h % q, -1, q))) % Q
    return x, Q


@pytest.mark.parametrize("kind,ell", [("c1", 4), ("c1", 3), ((12, 5, 3, 45), 7), ((13, 4, 2, 50), 5), ((16, 18, 3, 59), 20)])
def test_inner_product_stage_bit_exact(kind, ell):
    """k_ip_tma alone (EvalFastKeySwitchCoreExt) through ppq_ks_inner_product against oracle.inner_product."""
    from ppqsflhe_b200 import EvalKey, to_device_u64, to_host_u64
    from ppqsflhe_b200._capi import check, lib, vp
    o, cc = _cc_pair(kind)
    rng = np.random.default_rng(31)
    n = 2
    c1 = np.stack([_rand_poly(rng, o, range(ell)) for _ in range(n)])
    digits = cc.KeySwitchModUp(to_device_u64(c1, cc.device))
    kb, ka = _rand_key(rng, o)
    rk = EvalKey(cc, kb, ka)
    M = ell + o.K
    ext = torch.empty((n, 2, M, cc.N), dtype=torch.int64, device=cc.device)
    check(lib().ppq_ks_inner_product(cc.h, rk.h, ell, vp(digits.data_ptr()), vp(ext.data_ptr()), n, None))
    got = to_host_u64(ext)
    dh = to_host_u64(digits)
    for g in range(n):
        w0, w1 = o.inner_product(ell, dh[g], kb, ka)
        assert np.array_equal(got[g][0], w0) and np.array_equal(got[g][1], w1)
    if o.N <= 8192:  # the same in Python integers on a few coefficients of every limb
        idx = rng.integers(0, o.N, 16)
        lim = o.ext_limbs(ell)
        for r, gl in enumerate(lim):
            q = int(o.moduli[gl])
            for i in idx:
                acc = sum(int(dh[0][j][r][i]) * int(kb[j][gl][i]) for j in range(dh.shape[1])) % q
                assert int(got[0][0][r][i]) == acc


@pytest.mark.parametrize("kind,ell", [("c1", 4), ("c1", 3), ((12, 5, 3, 45), 7), ((12, 5, 3, 45), 3), ((13, 38, 3, 59), 40)])
def test_moddown_stage_bit_exact_and_big_integer_property(kind, ell):
    """ppq_ks_moddown against oracle.moddown, and against the DEFINITION in Python integers: with X the exact lift of
    the extended polynomial over Q_ell x P, ApproxModDown returns floor(X / P) - u (mod q_i) with one 0 <= u < K per
    coefficient, the same u on every limb (the uncorrected overflow of the approximate conversion)."""
    import math
    from ppqsflhe_b200 import to_device_u64, to_host_u64
    from ppqsflhe_b200._capi import check, lib, vp
    o, cc = _cc_pair(kind)
    rng = np.random.default_rng(37)
    n = 2
    lim = o.ext_limbs(ell)
    M = len(lim)
    ext = np.stack([np.stack([_rand_poly(rng, o, lim) for _ in range(2)]) for _ in range(n)])
    d_ext = to_device_u64(ext, cc.device)
    out = torch.empty((n, 2, ell, cc.N), dtype=torch.int64, device=cc.device)
    check(lib().ppq_ks_moddown(cc.h, ell, vp(d_ext.data_ptr()), vp(out.data_ptr()), n, None))
    got = to_host_u64(out)
    for g in range(n):
        for poly in range(2):
            assert np.array_equal(got[g][poly], o.moddown(ell, ext[g][poly]))
    if o.N > 4096 or M > 12:
        return
    # big-integer definition, polynomial (0, 0): bring every limb to COEFF with the GPU's own inverse NTT (pinned to
    # OpenFHE by the secret-key golden), lift, divide
    mods = [int(o.moduli[r]) for r in lim]
    P = math.prod(mods[ell:])
    co = to_device_u64(ext[0][0], cc.device)
    cc.ntt(co, lim, inverse=True)
    X, _ = _crt_lift(to_host_u64(co), mods)
    oc = out[0][0].clone()
    cc.ntt(oc, list(range(ell)), inverse=True)
    res = to_host_u64(oc)
    F = X // P
    K = o.K
    u0 = None
    for i in range(ell):
        q = mods[i]
        u = (F - np.array([int(v) for v in res[i]], dtype=object)) % q
        assert all(0 <= int(v) < K for v in u), "ModDown is not floor(X/P) - u with 0 <= u < K"
        if u0 is None:
            u0 = u
        assert all(int(a) == int(b) for a, b in zip(u, u0)), "u differs between limbs"


@pytest.mark.parametrize("kind,ell", [("c1", 4), ("c1", 3), ((12, 5, 3, 45), 7), ((12, 5, 3, 45), 5)])
def test_modup_big_integer_property(kind, ell):
    """ApproxSwitchCRTBasis in Python integers: limb t of digit j (COEFF) = (x + u D_j) mod t, x the exact lift of the
    digit's own limbs in [0, D_j), one 0 <= u < |digit j| per coefficient for every target limb t."""
    import math
    from ppqsflhe_b200 import to_device_u64, to_host_u64
    o, cc = _cc_pair(kind)
    rng = np.random.default_rng(41)
    c1 = np.stack([_rand_poly(rng, o, range(ell))])
    dig = cc.KeySwitchModUp(to_device_u64(c1, cc.device))[0]          # [parts][M][N] EVAL
    lim = o.ext_limbs(ell)
    M = len(lim)
    parts = dig.shape[0]
    co = dig.clone().view(parts * M, cc.N)
    cc.ntt(co, lim * parts, inverse=True)
    co = to_host_u64(co).reshape(parts, M, o.N)
    for j in range(parts):
        own = list(range(j * o.alpha, min((j + 1) * o.alpha, ell)))
        mods = [int(o.moduli[i]) for i in own]
        x, D = _crt_lift([co[j][i] for i in own], mods)
        y = [np.array([int(v) for v in co[j][i]], dtype=object) * pow(D // q % q, -1, q) % q for i, q in zip(own, mods)]
        v = sum(yi * (D // q) for yi, q in zip(y, mods))
        u = (v - x) // D
        assert all(int(a) * D + int(b) == int(c) for a, b, c in zip(u, x, v)) and all(0 <= int(a) < len(own) for a in u)
        for r, gl in enumerate(lim):
            if r in own:
                continue
            t = int(o.moduli[gl])
            assert all(int(a) == int(b) for a, b in zip(co[j][r], v % t)), f"digit {j} limb {r}"


@pytest.mark.parametrize("kind,ell", [("c1", 4), ("c1", 3), ((12, 5, 3, 45), 7)])
def test_rescale_is_exact_centred_division_big_integers(kind, ell):
    """DropLastElementAndScale in Python integers: with X the exact lift over q_0..q_{ell-1} and r the CENTRED residue of
    X mod q_last, the result is (X - r) / q_last reduced modulo every remaining limb.  (c = 1/sf makes the EvalMult
    constant 1, so ppq_eval_mult_const with rescale = 1 is the bare rescale.)"""
    import math
    from ppqsflhe_b200 import to_device_u64, to_host_u64
    from ppqsflhe_b200._capi import check, lib, vp
    o, cc = _cc_pair(kind)
    rng = np.random.default_rng(43)
    ct = _rand_ct(rng, o, ell, 1)
    d_in = to_device_u64(ct, cc.device)
    out = torch.empty((1, 2, ell - 1, cc.N), dtype=torch.int64, device=cc.device)
    sf = cc.sf(o.L - ell + 1)
    check(lib().ppq_eval_mult_const(cc.h, ell, vp(d_in.data_ptr()), vp(out.data_ptr()), 1, 1.0 / sf, sf, 1, None))
    mods = [int(o.moduli[i]) for i in range(ell)]
    ql = mods[-1]
    co = d_in[0][0].clone()
    cc.ntt(co, list(range(ell)), inverse=True)
    X, _ = _crt_lift(to_host_u64(co), mods)
    r = X % ql
    r = np.array([int(v) - ql if int(v) > ql // 2 else int(v) for v in r], dtype=object)
    Y = (X - r) // ql
    assert all((int(a) - int(b)) % ql == 0 for a, b in zip(X, r))
    oc = out[0][0].clone()
    cc.ntt(oc, list(range(ell - 1)), inverse=True)
    res = to_host_u64(oc)
    for i in range(ell - 1):
        q = mods[i]
        assert all(int(a) == int(b) % q for a, b in zip(res[i], Y)), f"limb {i}"


def test_empty_batches_and_bad_arguments_through_the_c_abi():
    # the reference's loops simply do not run on an empty file (server/src/changeCipherDomain.cpp:57-124 over an empty
    # weights_summary); here n = 0 is PPQ_OK and writes nothing, and a bad limb count / null pointer / empty aggregate is
    # PPQ_EINVAL with a message, never a launch
    import ctypes as C
    from ppqsflhe_b200 import EvalKey
    from ppqsflhe_b200._capi import lib, vp
    o, cc = _cc_pair("c1")
    L = lib()
    rng = np.random.default_rng(5)
    rk = EvalKey(cc, *_rand_key(rng, o))
    ell = 4
    buf = torch.full((2, 2, ell, cc.N), 7, dtype=torch.int64, device=cc.device)
    out = torch.full_like(buf, 9)
    p, q = vp(buf.data_ptr()), vp(out.data_ptr())
    assert L.ppq_reencrypt(cc.h, rk.h, ell, p, q, 0, None) == 0
    assert L.ppq_reencrypt_accumulate(cc.h, rk.h, ell, p, q, 0, None) == 0
    assert L.ppq_eval_add(cc.h, ell, p, p, q, 0, None) == 0
    assert L.ppq_finalize_sum(cc.h, ell, q, 0, None) == 0
    assert L.ppq_eval_mult_const(cc.h, ell, p, q, 0, 0.5, cc.sf(1), 1, None) == 0
    assert L.ppq_decrypt_core(cc.h, ell, p, p, q, 0, None) == 0
    assert L.ppq_encrypt(cc.h, ell, p, p, p, p, q, 0, None) == 0
    assert L.ppq_encode_encrypt(cc.h, ell, p, p, p, 0, 1.0, p, q, 0, None) == 0
    assert L.ppq_pack(cc.h, ell, p, q, 0, None) == 0 and L.ppq_unpack(cc.h, ell, p, q, 0, None) == 0
    assert L.ppq_ntt_forward(cc.h, q, None, 0, None) in (0, -1)  # a null id list may be refused; nothing is launched
    torch.cuda.synchronize()
    assert int((out != 9).sum().item()) == 0 and int((buf != 7).sum().item()) == 0
    for bad_ell in (0, -1, cc.L + 1):
        assert L.ppq_reencrypt(cc.h, rk.h, bad_ell, p, q, 1, None) == -1
        assert L.ppq_eval_add(cc.h, bad_ell, p, p, q, 1, None) == -1
        assert L.ppq_encrypt(cc.h, bad_ell, p, p, p, p, q, 1, None) == -1
        assert L.ppq_packed_words(cc.h, bad_ell) == 0
        assert b"limb" in L.ppq_last_error() or b"bad" in L.ppq_last_error()
    assert L.ppq_eval_mult_const(cc.h, 1, p, q, 1, 0.5, 1.0, 1, None) == -1  # nothing to rescale into
    assert L.ppq_reencrypt(cc.h, None, ell, p, q, 1, None) == -1
    assert L.ppq_reencrypt(cc.h, rk.h, ell, None, q, 1, None) == -1
    agg = C.c_void_p()
    assert L.ppq_agg_create(cc.h, ell, 2, C.byref(agg)) == 0
    assert L.ppq_agg_count(agg) == 0
    assert L.ppq_agg_finish(agg, q, None) == -1  # nothing was added
    assert L.ppq_agg_add_many(agg, None, None, 0, None) in (0, -1)
    L.ppq_agg_destroy(agg)
    torch.cuda.synchronize()
    assert int((out != 9).sum().item()) == 0


def test_short_value_vectors_are_zero_padded():
    # `mean` / `std_dev` are one-value plaintexts and the last chunk of a layer is ragged
    # (client/src/encryptModelWeights.cpp:82,90,100-109): n_vals < N/2 must equal the zero-padded full vector, bit for bit
    o, cc = _cc_pair("c1")
    rng = np.random.default_rng(6)
    for nv in (1, 5, 1000, cc.N // 2 - 1):
        vals = rng.normal(0, 0.1, (2, nv))
        full = np.zeros((2, cc.N // 2))
        full[:, :nv] = vals
        a = cc.MakeCKKSPackedPlaintext(vals)
        b = cc.MakeCKKSPackedPlaintext(full)
        assert torch.equal(a.data, b.data)


@pytest.mark.parametrize("kind,ell", [("c1", 4), ((12, 5, 3, 45), 7), ((13, 3, 2, 30), 3)])
def test_decode_short_interpolation_and_its_fallback(kind, ell):
    # ppq_decode interpolates from three limbs and checks the result against every other residue (ckks.cu); values too
    # large for that (here: values anywhere in (-Q/2, Q/2]) must take the full Garner path.  Residues are made from
    # Python integers; the checker is the oracle's decode, which always interpolates over all limbs.
    from ppqsflhe_b200._capi import check, lib, vp
    from ppqsflhe_b200 import to_device_u64
    o, cc = _cc_pair(kind)
    rng = np.random.default_rng(77)
    n = 2
    mods = [int(q) for q in o.moduli[:ell]]
    Q = 1
    for q in mods:
        Q *= q
    small = rng.integers(-(1 << 62), 1 << 62, (n, o.N), dtype=np.int64)
    vals = [[int(v) for v in small[g]] for g in range(n)]
    for k in range(0, o.N, 3):           # every third coefficient of ciphertext 1: anywhere in (-Q/2, Q/2]
        vals[1][k] = int(rng.integers(0, 1 << 62)) * (Q >> 63) - Q // 2 + 1
    vals[0][5] = -1
    vals[0][6] = (mods[0] * mods[1] * mods[2]) // 2 if ell > 3 else 7   # the edge of the short path's range
    vals[0][7] = -vals[0][6]
    co = np.zeros((n, ell, o.N), dtype=np.uint64)
    for g in range(n):
        for i, q in enumerate(mods):
            co[g, i] = np.array([v % q for v in vals[g]], dtype=np.uint64)
    sf = float(2 ** 40)
    out = torch.empty((n, o.N // 2), dtype=torch.float64, device=cc.device)
    d = to_device_u64(co, cc.device)
    check(lib().ppq_decode(cc.h, ell, vp(d.data_ptr()), sf, vp(out.data_ptr()), n, None))
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    for g in range(n):
        want = o.decode(ell, co[g], sf)
        scale = max(1.0, float(np.max(np.abs(want))))
        assert np.max(np.abs(got[g] - want)) / scale < 1e-9
