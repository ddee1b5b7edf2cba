"""The CPU oracle against every pin the reference tree offers, plus self-consistency KATs.
CPU only.  The oracle is TEST INFRASTRUCTURE (oracle/ppq_oracle.h)."""
import hashlib
import os

import numpy as np
import pytest

import oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = os.path.join(ROOT, "tests", "golden")


def test_ccjson_primes_roots_reproduced(ccjson_text, oracle_c1):
    info = oracle.parse_ccjson(ccjson_text)
    o = oracle_c1
    assert info["N"] == 16384 and info["dnum"] == 2 and info["ks"] == 2 and info["rs"] == 3
    # roots serialized by OpenFHE == minimal primitive 2N-th roots derived by the oracle
    assert [int(x) for x in o.roots[:4]] == info["roots"] == [62213374832584, 42618759, 36692422, 19]
    # the prime-selection rules regenerate the four Q primes (depth 2, scaling 40, first 60, extra 20)
    syn = oracle.Oracle.synthetic(14, 2, 60, 40, 20, 2)
    assert [int(x) for x in syn.moduli] == [int(x) for x in o.moduli]
    assert (o.K, o.alpha) == (2, 2)
    assert [int(x) for x in o.moduli[4:]] == [1152921504606683137, 1152921504606584833]
    # composite modulus of CC.json (three little-endian words) is the product of the four primes
    Q = 1
    for q in info["moduli"]:
        Q *= q
    assert Q == 613792465231773697 + (18415746740090024801 << 64) + (2281702819 << 128)


def test_real_openfhe_secret_keys(oracle_c1):
    """client_{1,2}-private.key: NTT(ternary s) must reproduce every serialized EVAL limb bit for bit."""
    o = oracle_c1
    g = np.load(os.path.join(G, "openfhe_sk_golden.npz"))
    for c in (1, 2):
        tern = g[f"c{c}_ternary"].astype(np.int64)
        assert set(np.unique(tern)) <= {-1, 0, 1}
        for i in range(4):
            limb = o.small_to_eval(i, tern)
            assert hashlib.sha256(limb.astype("<u8").tobytes()).hexdigest() == str(g[f"c{c}_sha256"][i])
            assert np.array_equal(limb[:8], g[f"c{c}_head"][i])


def test_ntt_definition_and_convolution():
    o = oracle.Oracle.synthetic(12, 2, 60, 40, 20, 2)
    rng = np.random.default_rng(0)
    for limb in (0, 1, o.M - 1):
        q = int(o.moduli[limb])
        a = rng.integers(0, q, o.N, dtype=np.uint64)
        assert np.array_equal(o.ntt(limb, a), o.ntt_naive(limb, a))
        assert np.array_equal(o.ntt(limb, o.ntt(limb, a), inverse=True), a)
    # NTT(a) .* NTT(b) == NTT(negacyclic schoolbook product), sparse b to keep it cheap
    limb, q = 1, int(o.moduli[1])
    a = rng.integers(0, q, o.N, dtype=np.uint64)
    b = np.zeros(o.N, dtype=np.uint64)
    idx = [0, 3, 77, o.N - 1]
    for i in idx:
        b[i] = rng.integers(1, q)
    prod = np.zeros(o.N, dtype=object)
    for i in idx:
        rolled = np.roll(a.astype(object), i)
        rolled[:i] = -rolled[:i]
        prod = (prod + rolled * int(b[i])) % q
    ea, eb = o.ntt(limb, a), o.ntt(limb, b)
    pw = np.array([(int(x) * int(y)) % q for x, y in zip(ea, eb)], dtype=np.uint64)
    assert np.array_equal(o.ntt(limb, pw, inverse=True), prod.astype(np.uint64))


@pytest.mark.parametrize("ell", [4, 3])
def test_reencrypt_is_a_key_switch(oracle_c1, ell):
    """Decrypt_{sk2}(ReEncrypt_{1->2}(ct)) == Decrypt_{sk1}(ct) up to key-switch noise (stage composition too)."""
    o = oracle_c1
    s1, sk1, pkb1, pka1 = o.keygen(1)
    s2, sk2, pkb2, pka2 = o.keygen(2)
    kb, ka = o.rekeygen(3, sk1, pkb2, pka2)
    rng = np.random.default_rng(1)
    ct = np.stack([np.stack([rng.integers(0, int(o.moduli[i]), o.N, dtype=np.uint64) for i in range(ell)])
                   for _ in range(2)])
    out = o.reencrypt(ell, ct, kb, ka)
    m1 = o.decrypt_core(ell, sk1, ct)
    m2 = o.decrypt_core(ell, sk2, out)
    for i in range(ell):
        q = int(o.moduli[i])
        d = (m2[i].astype(np.int64) - m1[i].astype(np.int64)) % q
        d = np.where(d > q // 2, d - q, d)
        assert np.max(np.abs(d)) < 2 ** 16, f"limb {i}: key-switch noise too large"
    # stages compose to the whole
    dig = o.modup(ell, ct[1])
    e0, e1 = o.inner_product(ell, dig, kb, ka)
    ks0, ks1 = o.moddown(ell, e0), o.moddown(ell, e1)
    assert np.array_equal(out[1], ks1)
    assert np.array_equal(out[0], o.eval_add(ell, np.stack([ct[0], ct[0]]), np.stack([ks0, ks0]))[0])


def test_end_to_end_round_on_the_golden_weights(oracle_c1):
    """Oracle alone on the reference's recorded round (first chunk of values): enc, re-enc, add, x0.5, dec."""
    o = oracle_c1
    g = np.load(os.path.join(G, "weights_golden.npz"))
    slots = o.N // 2
    w1 = g["c1_sample_values"][:slots].astype(np.float64)
    w2 = g["c2_sample_values"][:slots].astype(np.float64)
    s1, sk1, pkb1, pka1 = o.keygen(11)
    s2, sk2, pkb2, pka2 = o.keygen(12)
    kb, ka = o.rekeygen(13, sk1, pkb2, pka2)
    sf0 = o.sf_big(0)
    cts = []
    for c, (w, pkb, pka) in enumerate(((w1, pkb1, pka1), (w2, pkb2, pka2))):
        pt = o.encode(4, w, sf0)
        cts.append(o.encrypt(4, pkb, pka, pt, oracle.sample_ternary(50 + c, o.N), oracle.sample_gauss(60 + c, o.N),
                             oracle.sample_gauss(70 + c, o.N)))
    c12 = o.reencrypt(4, cts[0], kb, ka)
    s = o.eval_add(4, cts[1], c12)
    avg, meta = o.eval_mult_const(4, s, dict(noise_scale_deg=2, level=0, scaling_factor=sf0), 0.5)
    assert avg.shape[1] == 3 and meta["level"] == 1 and meta["noise_scale_deg"] == 2
    assert abs(meta["scaling_factor"] / 1.2089e24 - 1) < 1e-3
    dec = o.decrypt(3, sk2, avg, meta["scaling_factor"])
    assert np.max(np.abs(dec - (w1 + w2) / 2)) < 2.0 ** -20
    assert np.max(np.abs(dec - g["c2_decrypted_values"][:slots])) < 2.0 ** -20


def test_golden_weight_pair_is_the_fedavg_mean():
    g = np.load(os.path.join(G, "weights_golden.npz"))
    want = (g["c1_sample_values"].astype(np.float64) + g["c2_sample_values"].astype(np.float64)) / 2
    assert np.max(np.abs(g["c1_decrypted_values"] - want)) < 2e-8
    assert np.max(np.abs(g["c2_decrypted_values"] - want)) < 2e-8
