"""ctypes binding of the CPU ORACLE (oracle/ppq_oracle.c).

TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this module; the product package (ppqsflhe_b200) never does.
Pin status: see oracle/ppq_oracle.h ("partially pinned").
"""
import ctypes as C
import json
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "build", "libppq_oracle.so")

u64p = C.POINTER(C.c_uint64)
i64p = C.POINTER(C.c_int64)
f64p = C.POINTER(C.c_double)


def build(force=False):
    src = [os.path.join(_HERE, f) for f in ("ppq_oracle.c", "ppq_oracle.h", "Makefile")]
    stale = (not os.path.exists(_LIB_PATH)) or any(
        os.path.exists(s) and os.path.getmtime(s) > os.path.getmtime(_LIB_PATH) for s in src)
    if force or stale:
        subprocess.check_call(["make", "-C", _HERE], stdout=subprocess.DEVNULL)
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        L.ppqo_ctx_create.restype = C.c_void_p
        L.ppqo_ctx_create.argtypes = [C.c_int, C.c_int, u64p, C.c_int, C.c_int, C.c_int]
        L.ppqo_ctx_create_synthetic.restype = C.c_void_p
        L.ppqo_ctx_create_synthetic.argtypes = [C.c_int] * 6
        L.ppqo_ctx_free.argtypes = [C.c_void_p]
        for f in ("ppqo_logN", "ppqo_L", "ppqo_K", "ppqo_dnum", "ppqo_alpha"):
            getattr(L, f).argtypes = [C.c_void_p]
            getattr(L, f).restype = C.c_int
        L.ppqo_num_parts.argtypes = [C.c_void_p, C.c_int]
        L.ppqo_num_parts.restype = C.c_int
        L.ppqo_scaling_factor.restype = C.c_double
        L.ppqo_scaling_factor.argtypes = [C.c_void_p, C.c_int]
        L.ppqo_scaling_factor_big.restype = C.c_double
        L.ppqo_scaling_factor_big.argtypes = [C.c_void_p, C.c_int]
        for f in ("ppqo_mulmod", "ppqo_powmod"):
            getattr(L, f).restype = C.c_uint64
            getattr(L, f).argtypes = [C.c_uint64] * 3
        L.ppqo_invmod.restype = C.c_uint64
        L.ppqo_invmod.argtypes = [C.c_uint64] * 2
        L.ppqo_is_prime.argtypes = [C.c_uint64]
        L.ppqo_first_prime.restype = C.c_uint64
        L.ppqo_first_prime.argtypes = [C.c_int, C.c_uint64]
        for f in ("ppqo_previous_prime", "ppqo_next_prime", "ppqo_min_root_of_unity"):
            getattr(L, f).restype = C.c_uint64
            getattr(L, f).argtypes = [C.c_uint64] * 2
        L.ppqo_max_threads.restype = C.c_int
        _lib = L
    return _lib


def _p(a, t=u64p):
    return a.ctypes.data_as(t)


def _u64(a):
    a = np.ascontiguousarray(a, dtype=np.uint64)
    return a


class Oracle:
    """One CKKS-RNS context (N, Q chain, derived P primes, tables)."""

    def __init__(self, handle):
        self.h = C.c_void_p(handle)
        L = lib()
        self.logN = L.ppqo_logN(self.h)
        self.N = 1 << self.logN
        self.L = L.ppqo_L(self.h)
        self.K = L.ppqo_K(self.h)
        self.dnum = L.ppqo_dnum(self.h)
        self.alpha = L.ppqo_alpha(self.h)
        self.M = self.L + self.K
        m = np.zeros(self.M, dtype=np.uint64)
        L.ppqo_moduli(self.h, _p(m))
        self.moduli = m
        r = np.zeros(self.M, dtype=np.uint64)
        L.ppqo_roots(self.h, _p(r))
        self.roots = r

    # ---- constructors
    @classmethod
    def from_moduli(cls, logN, q_moduli, dnum, aux_bits=60, extra_bits=20):
        q = _u64(q_moduli)
        return cls(lib().ppqo_ctx_create(logN, len(q), _p(q), dnum, aux_bits, extra_bits))

    @classmethod
    def from_ccjson(cls, text):
        """Parse the cereal-JSON CryptoContext the reference ships (server/storage/CC.json)."""
        info = parse_ccjson(text)
        return cls.from_moduli(info["logN"], info["moduli"], info["dnum"], info["aux_bits"], info["extra_bits"])

    @classmethod
    def synthetic(cls, logN, depth, first_bits=60, scale_bits=59, extra_bits=20, dnum=3):
        return cls(lib().ppqo_ctx_create_synthetic(logN, depth, first_bits, scale_bits, extra_bits, dnum))

    def close(self):
        if self.h:
            lib().ppqo_ctx_free(self.h)
            self.h = None

    def sf(self, level):
        return lib().ppqo_scaling_factor(self.h, level)

    def sf_big(self, level):
        return lib().ppqo_scaling_factor_big(self.h, level)

    def num_parts(self, ell):
        return lib().ppqo_num_parts(self.h, ell)

    def ext_limbs(self, ell):
        return list(range(ell)) + [self.L + k for k in range(self.K)]

    # ---- polynomial layer
    def ntt(self, limb, a, inverse=False):
        a = _u64(a).copy()
        (lib().ppqo_ntt_inv if inverse else lib().ppqo_ntt_fwd)(self.h, int(limb), _p(a))
        return a

    def ntt_naive(self, limb, a):
        a = _u64(a)
        o = np.zeros_like(a)
        lib().ppqo_ntt_fwd_naive(self.h, int(limb), _p(a), _p(o))
        return o

    def ntt_batch(self, data, limb_ids, inverse=False, nthreads=0):
        """data: [n_polys][N] in place; limb_ids: [n_polys]."""
        ids = np.ascontiguousarray(limb_ids, dtype=np.int32)
        assert data.dtype == np.uint64 and data.flags.c_contiguous
        lib().ppqo_ntt_batch(self.h, int(inverse), ids.ctypes.data_as(C.POINTER(C.c_int)),
                             C.c_size_t(len(ids)), _p(data), int(nthreads))
        return data

    # ---- key switch
    def modup(self, ell, c1):
        c1 = _u64(c1)
        out = np.zeros((self.num_parts(ell), ell + self.K, self.N), dtype=np.uint64)
        lib().ppqo_modup(self.h, ell, _p(c1), _p(out))
        return out

    def inner_product(self, ell, digits, kb, ka):
        o0 = np.zeros((ell + self.K, self.N), dtype=np.uint64)
        o1 = np.zeros_like(o0)
        lib().ppqo_inner_product(self.h, ell, _p(_u64(digits)), _p(_u64(kb)), _p(_u64(ka)), _p(o0), _p(o1))
        return o0, o1

    def moddown(self, ell, ext):
        out = np.zeros((ell, self.N), dtype=np.uint64)
        lib().ppqo_moddown(self.h, ell, _p(_u64(ext)), _p(out))
        return out

    def reencrypt(self, ell, ct, kb, ka):
        ct = _u64(ct)
        out = np.zeros((2, ell, self.N), dtype=np.uint64)
        lib().ppqo_reencrypt(self.h, ell, _p(ct), _p(_u64(kb)), _p(_u64(ka)), _p(out))
        return out

    def eval_add(self, ell, a, b):
        out = np.zeros((2, ell, self.N), dtype=np.uint64)
        lib().ppqo_eval_add(self.h, ell, _p(_u64(a)), _p(_u64(b)), _p(out))
        return out

    def rescale(self, ell, ct):
        out = np.zeros((2, ell - 1, self.N), dtype=np.uint64)
        lib().ppqo_rescale(self.h, ell, _p(_u64(ct)), _p(out))
        return out

    def mult_const_factors(self, ell, sf, c):
        f = np.zeros(ell, dtype=np.uint64)
        lib().ppqo_mult_const_factors(self.h, ell, C.c_double(sf), C.c_double(c), _p(f))
        return f

    def mult_scalar(self, ell, factors, ct):
        out = np.zeros((2, ell, self.N), dtype=np.uint64)
        lib().ppqo_mult_scalar(self.h, ell, _p(_u64(factors)), _p(_u64(ct)), _p(out))
        return out

    def eval_mult_const(self, ell, ct, meta, c):
        """EvalMult(ct, double) under FLEXIBLEAUTOEXT: rescale first when noiseScaleDeg == 2."""
        meta = dict(meta)
        if meta["noise_scale_deg"] == 2:
            ct = self.rescale(ell, ct)
            meta["scaling_factor"] /= float(self.moduli[ell - 1])
            ell -= 1
            meta["level"] += 1
            meta["noise_scale_deg"] -= 1
        sf = self.sf(meta["level"])
        f = self.mult_const_factors(ell, sf, c)
        out = self.mult_scalar(ell, f, ct)
        meta["noise_scale_deg"] += 1
        meta["scaling_factor"] *= sf
        return out, meta

    # ---- CKKS
    def encode(self, ell, vals, sf):
        vals = np.ascontiguousarray(vals, dtype=np.float64)
        pt = np.zeros((ell, self.N), dtype=np.uint64)
        lib().ppqo_encode(self.h, ell, _p(vals, f64p), C.c_size_t(len(vals)), C.c_double(sf), _p(pt))
        return pt

    def encrypt(self, ell, pkb, pka, pt, v, e0, e1):
        ct = np.zeros((2, ell, self.N), dtype=np.uint64)
        v, e0, e1 = (np.ascontiguousarray(x, dtype=np.int64) for x in (v, e0, e1))
        lib().ppqo_encrypt(self.h, ell, _p(_u64(pkb)), _p(_u64(pka)), _p(_u64(pt)),
                           _p(v, i64p), _p(e0, i64p), _p(e1, i64p), _p(ct))
        return ct

    def decrypt_core(self, ell, sk, ct):
        co = np.zeros((ell, self.N), dtype=np.uint64)
        lib().ppqo_decrypt_core(self.h, ell, _p(_u64(sk)), _p(_u64(ct)), _p(co))
        return co

    def decode(self, ell, coeff, sf):
        out = np.zeros(self.N // 2, dtype=np.float64)
        lib().ppqo_decode(self.h, ell, _p(_u64(coeff)), C.c_double(sf), _p(out, f64p))
        return out

    def decrypt(self, ell, sk, ct, sf):
        out = np.zeros(self.N // 2, dtype=np.float64)
        lib().ppqo_decrypt(self.h, ell, _p(_u64(sk)), _p(_u64(ct)), C.c_double(sf), _p(out, f64p))
        return out

    # ---- scaffolding
    def small_to_eval(self, limb, small):
        small = np.ascontiguousarray(small, dtype=np.int64)
        out = np.zeros(self.N, dtype=np.uint64)
        lib().ppqo_small_to_eval(self.h, int(limb), _p(small, i64p), _p(out))
        return out

    def keygen(self, seed):
        s = np.zeros(self.N, dtype=np.int64)
        sk = np.zeros((self.M, self.N), dtype=np.uint64)
        pkb = np.zeros_like(sk)
        pka = np.zeros_like(sk)
        lib().ppqo_keygen(self.h, C.c_uint64(seed), _p(s, i64p), _p(sk), _p(pkb), _p(pka))
        return s, sk, pkb, pka

    def rekeygen(self, seed, sk_old, pkb_new, pka_new):
        kb = np.zeros((self.dnum, self.M, self.N), dtype=np.uint64)
        ka = np.zeros_like(kb)
        lib().ppqo_rekeygen(self.h, C.c_uint64(seed), _p(_u64(sk_old)), _p(_u64(pkb_new)), _p(_u64(pka_new)),
                            _p(kb), _p(ka))
        return kb, ka

    def reencrypt_accumulate_batch(self, ell, cts, keys_b, keys_a, nthreads=0):
        """cts: [n_clients][n_slots][2][ell][N]; keys_*: list of [dnum][M][N]. Returns acc [n_slots][2][ell][N]."""
        cts = _u64(cts)
        nc, ns = cts.shape[0], cts.shape[1]
        acc = np.zeros((ns, 2, ell, self.N), dtype=np.uint64)
        kb = (u64p * nc)(*[_p(k) for k in keys_b])
        ka = (u64p * nc)(*[_p(k) for k in keys_a])
        lib().ppqo_reencrypt_accumulate_batch(self.h, ell, nc, ns, _p(cts), kb, ka, _p(acc), int(nthreads))
        return acc


def sample_ternary(seed, n):
    out = np.zeros(n, dtype=np.int64)
    lib().ppqo_sample_ternary(C.c_uint64(seed), C.c_size_t(n), _p(out, i64p))
    return out


def sample_gauss(seed, n, sigma=3.19):
    out = np.zeros(n, dtype=np.int64)
    lib().ppqo_sample_gauss(C.c_uint64(seed), C.c_size_t(n), C.c_double(sigma), _p(out, i64p))
    return out


def sample_uniform(seed, q, n):
    out = np.zeros(n, dtype=np.uint64)
    lib().ppqo_sample_uniform(C.c_uint64(seed), C.c_uint64(int(q)), C.c_size_t(n), _p(out))
    return out


def max_threads():
    """Host threads the CPU baseline may use: the cores this process may run on.  (torchrun exports
    OMP_NUM_THREADS=1, so omp_get_max_threads() alone would silently time a single thread.)"""
    import os
    try:
        n = len(os.sched_getaffinity(0))
    except AttributeError:
        n = os.cpu_count() or 1
    return max(n, lib().ppqo_max_threads())


def parse_ccjson(text):
    """Extract (logN, Q moduli, roots, dnum, aux/extra bits, ...) from a cereal-JSON CryptoContext
    (fields per SURVEY 0.3: elp.rd, elp.p[*].cm/ru, dnum, ab, eb, ks, rs)."""
    d = json.loads(text)

    found = {}

    def walk(o):
        if isinstance(o, dict):
            for k, v in o.items():
                if k in ("dnum", "ab", "eb", "ks", "rs") and not isinstance(v, (dict, list)):
                    found.setdefault(k, int(v))
                if k == "elp" and "elp" not in found:
                    found["elp"] = v
                if k == "enp" and "enp" not in found:
                    found["enp"] = v
                walk(v)
        elif isinstance(o, list):
            for v in o:
                walk(v)

    walk(d)
    elp = found["elp"]["ptr_wrapper"]["data"]
    top = elp["value0"] if "value0" in elp else elp
    N = int(top["rd"])
    limbs = elp["p"]
    moduli, roots = [], []
    for p in limbs:
        v = p["ptr_wrapper"]["data"]
        v = v["value0"] if "value0" in v else v
        moduli.append(int(v["cm"]["v"]))
        roots.append(int(v["ru"]["v"]))
    return {
        "logN": N.bit_length() - 1, "N": N, "moduli": moduli, "roots": roots,
        "dnum": found.get("dnum", 0), "aux_bits": found.get("ab", 60), "extra_bits": found.get("eb", 0),
        "ks": found.get("ks"), "rs": found.get("rs"),
    }
