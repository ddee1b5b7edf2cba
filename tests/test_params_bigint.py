"""Independent big-integer derivation of every constant the library precomputes (no GPU, no C oracle in the loop).

`csrc/params.cpp` (product) and `oracle/ppq_oracle.c` (checker) are two C restatements of OpenFHE's
PrecomputeCRTTables by the same author, so comparing them with each other shares their blind spots.  Here the same
quantities are re-derived in plain Python integers with sympy's primality test, straight from the definitions of
SURVEY.md Appendix A (A.1 primes/roots, A.3 conversion tables, A.5 scaling factors, A.6 rescale), and compared with
what the product library reports through the host-only `ppq_params_dump` AND with the oracle.  The reference-held
facts used: `server/storage/CC.json` (N, four Q primes and their roots, dnum, auxBits).
"""
import ctypes as C
import json
import math
import os

import numpy as np
import pytest

sympy = pytest.importorskip("sympy")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def first_prime(bits, m):
    q = (1 << bits) + 1
    q += (-(q - 1)) % m
    while not sympy.isprime(q):
        q += m
    return q


def previous_prime(q, m):
    q -= m
    while not sympy.isprime(q):
        q -= m
    return q


def min_primitive_root(m, q):
    """minimum over ALL primitive m-th roots of unity mod q (m a power of two), by exhaustion"""
    assert (q - 1) % m == 0
    g = 2
    while True:
        r = pow(g, (q - 1) // m, q)
        if pow(r, m // 2, q) == q - 1:
            break
        g += 1
    best, x, r2 = r, r, r * r % q
    for _ in range(1, m, 2):
        best = min(best, x)
        x = x * r2 % q
    return best


def dump(logN, q, dnum, aux_bits, extra_bits, ell):
    from ppqsflhe_b200 import _capi
    L = _capi.lib()
    arr = (C.c_uint64 * len(q))(*q)
    need = C.c_size_t(0)
    assert L.ppq_params_dump(logN, len(q), arr, dnum, aux_bits, extra_bits, ell, None, 0, C.byref(need)) == 0
    buf = C.create_string_buffer(need.value)
    assert L.ppq_params_dump(logN, len(q), arr, dnum, aux_bits, extra_bits, ell, buf, need.value, C.byref(need)) == 0
    d = json.loads(buf.value.decode())
    for k, v in d.items():
        if isinstance(v, list) and v and isinstance(v[0], str):
            d[k] = [int(x) for x in v]
    return d


def derive(logN, q, dnum, aux_bits):
    N, m, L = 1 << logN, 2 << logN, len(q)
    alpha = -(-L // dnum)
    digits = [list(range(j * alpha, min(L, (j + 1) * alpha))) for j in range(dnum)]
    K = max(-(-math.prod(q[i] for i in d).bit_length() // aux_bits) for d in digits)
    P, c = [], first_prime(aux_bits, m)
    while len(P) < K:
        c = previous_prime(c, m)
        if c not in q:
            P.append(c)
    return N, m, alpha, K, P


CASES = [("cc", None), ("syn", (12, 5, 3, 45)), ("syn", (13, 18, 3, 59)), ("syn", (12, 6, 2, 50))]


def chain_of(case):
    kind, prm = case
    if kind == "cc":
        import oracle
        cc = oracle.parse_ccjson(open(os.path.join(ROOT, "tests", "golden", "CC.json")).read())
        return cc["logN"], [int(x) for x in cc["moduli"]], cc["dnum"], 60, 20
    logN, depth, dnum, scale = prm
    # the synthetic chain itself is product code; it is checked separately below -- here it only supplies a Q basis
    import oracle
    o = oracle.Oracle.synthetic(logN, depth, 60, scale, 20, dnum)
    q = [int(x) for x in o.moduli[:o.L]]
    o.close()
    return logN, q, dnum, 60, 20


@pytest.mark.parametrize("case", CASES)
def test_basis_roots_and_scaling_factors(case):
    logN, q, dnum, aux, extra = chain_of(case)
    N, m, alpha, K, P = derive(logN, q, dnum, aux)
    L = len(q)
    d = dump(logN, q, dnum, aux, extra, L)
    assert (d["K"], d["alpha"], d["L"]) == (K, alpha, L)
    assert d["moduli"] == q + P, "auxiliary primes differ from the PreviousPrime chain below FirstPrime(auxBits, 2N)"
    if logN <= 13 or case[0] == "cc":
        for r, mod in enumerate(q + P):
            assert d["roots"][r] == min_primitive_root(m, mod), f"root of limb {r} is not the minimal primitive 2N-th root"
    for r, mod in enumerate(q + P):
        assert d["ninv"][r] * N % mod == 1
        assert pow(d["roots"][r], N, mod) == mod - 1
    # A.5 (FLEXIBLEAUTOEXT): SF[0] = q_{L-1}, SF[1] = q_{L-2}, SF[k] = SF[k-1]^2 / q_{L-k}; SFBig[0] = SF[0] SF[1]
    sf = [float(q[L - 1]), float(q[L - 2])]
    for k in range(2, L):
        sf.append(sf[k - 1] * sf[k - 1] / float(q[L - k]))
    assert d["sf"] == sf[:L]
    assert d["sfbig"][0] == sf[0] * sf[1]
    for k in range(1, L - 1):
        assert d["sfbig"][k] == sf[k] * sf[k]
    # the oracle must agree with the same independent numbers
    import oracle
    o = oracle.Oracle.from_moduli(logN, np.array(q, dtype=np.uint64), dnum, aux, extra)
    assert [int(x) for x in o.moduli] == q + P and [int(x) for x in o.roots] == d["roots"]
    assert [o.sf(i) for i in range(L)] == sf[:L] and o.sf_big(0) == sf[0] * sf[1]
    o.close()


def test_cc_json_pins_the_predicted_aux_primes():
    """for CC.json the rule predicts p_0 = 1152921504606683137, p_1 = 1152921504606584833."""
    logN, q, dnum, aux, extra = chain_of(("cc", None))
    _, _, alpha, K, P = derive(logN, q, dnum, aux)
    assert (alpha, K) == (2, 2) and P == [1152921504606683137, 1152921504606584833]


@pytest.mark.parametrize("case", CASES)
def test_key_switch_tables_at_every_level(case):
    """ModUp / ModDown / rescale constants at every ell from L down to 1 (uneven last digit included)."""
    logN, q, dnum, aux, extra = chain_of(case)
    N, m, alpha, K, P = derive(logN, q, dnum, aux)
    L = len(q)
    Pprod = math.prod(P)
    for ell in range(L, 0, -1):
        d = dump(logN, q, dnum, aux, extra, ell)
        parts = min(-(-ell // alpha), dnum)
        assert d["parts"] == parts
        M = ell + K
        ext = q[:ell] + P  # moduli of the extended basis at this level, slot r -> modulus
        seen = 0
        for j in range(parts):
            own = list(range(j * alpha, min((j + 1) * alpha, ell)))
            D = math.prod(q[i] for i in own)
            for i in own:
                hat = D // q[i]
                want = pow(hat % q[i], -1, q[i]) * pow(N, -1, q[i]) % q[i]
                assert d["c1_scale"][i] == want, f"ell={ell}: (D_j/q_i)^-1 N^-1 of limb {i}"
            for r in range(M):
                if r in own:
                    continue
                src_first, ns, dst_slot, limb, off = d["up_jobs"][seen]
                seen += 1
                assert (src_first, ns, dst_slot) == (own[0], len(own), j * M + r)
                assert limb == (r if r < ell else L + (r - ell)), "P-limb index mapping when ell < L"
                t = ext[r]
                assert d["up_chat"][off:off + ns] == [(D // q[i]) % t for i in own], f"ell={ell} digit {j} -> slot {r}"
        assert seen == len(d["up_jobs"])
        for k in range(K):
            want = pow((Pprod // P[k]) % P[k], -1, P[k]) * pow(N, -1, P[k]) % P[k]
            assert d["p_scale"][k] == want and d["p_scale"][K + k] == want
        for i in range(ell):
            pinv = pow(Pprod % q[i], -1, q[i])
            assert d["pinv"][i] == pinv
            for poly in range(2):
                off = (poly * ell + i) * K
                assert d["down_chat"][off:off + K] == [(Pprod // P[k]) % q[i] * pinv % q[i] for k in range(K)]
        if ell >= 2:
            assert d["rescale_qlinv"] == [pow(q[ell - 1] % q[i], -1, q[i]) for i in range(ell - 1)]


def test_approximate_conversion_identity_in_integers():
    """What 'approximate' means: sum_i [x_i (D/q_i)^-1]_{q_i} (D/q_i) = x + u D with 0 <= u < |digit|,
    for the exact lift x in [0, D): the unreduced sum the kernels reduce modulo every target."""
    logN, q, dnum, aux, extra = chain_of(("syn", (12, 5, 3, 45)))
    rng = np.random.default_rng(3)
    own = q[:3]
    D = math.prod(own)
    for _ in range(200):
        x = int(rng.integers(0, 1 << 62)) * int(rng.integers(0, 1 << 62)) * int(rng.integers(0, 1 << 62)) % D
        v = sum((x % qi) * pow(D // qi % qi, -1, qi) % qi * (D // qi) for qi in own)
        assert v % D == x and 0 <= (v - x) // D < len(own)
