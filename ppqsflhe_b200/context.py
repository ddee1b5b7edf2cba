"""Host-side mirror of the OpenFHE CryptoContext<DCRTPoly> surface that PPQSFLHE uses on its hot path, over the
C ABI of libppq_b200.so.  Names, argument meaning and error behaviour follow the reference's call sites:

    cc->ReEncrypt(ct, reKey)            server/src/changeCipherDomain.cpp:74,89,105
    cc->EvalAdd(ct1, ct2)               server/src/aggregateEncryptedWeights.cpp:82,91,106
    cc->EvalMult(ct, 0.5)               server/src/aggregateEncryptedWeights.cpp:83,92,107
    cc->MakeCKKSPackedPlaintext(vals)   client/src/encryptModelWeights.cpp:82,90,109
    cc->Encrypt(pk, pt)                 client/src/encryptModelWeights.cpp:83,91,110
    cc->Decrypt(sk, ct, &pt)            client/src/decryptModelWeights.cpp:81,90,108
    pt->SetLength / GetRealPackedValue  client/src/decryptModelWeights.cpp:82-83,91-92,109

Every object here is a BATCH (the reference loops over ciphertexts one at a time; the B200 path processes all of
them per launch).  torch is used only for device memory and streams.  No CPU fallback: constructing a context
without a CUDA device raises.
"""
import ctypes as C
from dataclasses import dataclass, replace

import numpy as np
import torch

from . import _capi
from ._capi import check, lib, vp


def _ptr(t):
    return vp(t.data_ptr())


def _stream():
    return vp(torch.cuda.current_stream().cuda_stream)


def _seed32(seed):
    """None -> NULL (the library draws a fresh 256-bit key from the OS for this call); an int or bytes -> 32 bytes of
    key material for tests and reproducible runs."""
    if seed is None:
        return None
    if isinstance(seed, (bytes, bytearray)):
        if len(seed) != 32:
            raise ValueError("a byte seed must be 32 bytes")
        return bytes(seed)
    return int(seed).to_bytes(32, "little", signed=False)


def to_device_u64(a, device):
    """numpy uint64 array -> torch int64 CUDA tensor with the same bits."""
    a = np.ascontiguousarray(a, dtype=np.uint64)
    return torch.from_numpy(a.view(np.int64)).to(device)


def to_host_u64(t):
    return t.detach().cpu().numpy().view(np.uint64)


@dataclass
class CtMeta:
    """The scalar fields of an OpenFHE CiphertextImpl (d, l, h, e, s, sl)."""
    noise_scale_deg: int = 2
    level: int = 0
    hops_level: int = 0
    scaling_factor: float = 1.0
    slots: int = 0


class CiphertextBatch:
    """n ciphertexts of ell limbs: data is a CUDA int64 tensor [n][2][ell][N] (uint64 bit patterns)."""

    def __init__(self, data, meta):
        assert data.dim() == 4 and data.shape[1] == 2 and data.dtype == torch.int64 and data.is_cuda
        self.data = data.contiguous()
        self.meta = meta

    @property
    def n(self):
        return self.data.shape[0]

    @property
    def ell(self):
        return self.data.shape[2]

    def numpy(self):
        return to_host_u64(self.data)


class PlaintextBatch:
    """CKKS packed plaintexts, either encoded (EVAL residues) or decoded (real values)."""

    def __init__(self, data=None, meta=None, values=None):
        self.data = data      # CUDA int64 [n][ell][N] or None
        self.meta = meta
        self.values = values  # CUDA float64 [n][N/2] or None
        self.length = None

    def SetLength(self, n):
        self.length = int(n)

    def GetRealPackedValue(self):
        v = self.values.detach().cpu().numpy()
        return v if self.length is None else v[:, : self.length]


class EvalKey:
    """Re-encryption key (EvalKeyRelin B/A vectors, [dnum][L+K][N] EVAL) resident on the device."""

    def __init__(self, cc, b, a):
        self.cc = cc
        self.b = to_device_u64(b, cc.device) if isinstance(b, np.ndarray) else b
        self.a = to_device_u64(a, cc.device) if isinstance(a, np.ndarray) else a
        exp = (cc.dnum, cc.L + cc.K, cc.N)
        if tuple(self.b.shape) != exp or tuple(self.a.shape) != exp:
            raise ValueError(f"rekey must have shape {exp}")
        h = vp()
        check(lib().ppq_rekey_from_device(cc.h, _ptr(self.b), _ptr(self.a), C.byref(h)))
        self.h = h

    @classmethod
    def upload(cls, cc, b, a):
        """Host arrays -> a handle that owns its device copy (ppq_rekey_upload); nothing is borrowed."""
        b = np.ascontiguousarray(b, dtype=np.uint64)
        a = np.ascontiguousarray(a, dtype=np.uint64)
        exp = (cc.dnum, cc.L + cc.K, cc.N)
        if b.shape != exp or a.shape != exp:
            raise ValueError(f"rekey must have shape {exp}")
        self = cls.__new__(cls)
        self.cc, self.b, self.a = cc, None, None
        h = vp()
        check(lib().ppq_rekey_upload(cc.h, b.ctypes.data_as(_capi.u64p), a.ctypes.data_as(_capi.u64p), C.byref(h)))
        self.h = h
        return self

    def __del__(self):
        if getattr(self, "h", None):
            lib().ppq_rekey_destroy(self.h)
            self.h = None


class Aggregator:
    """Running sum over clients of ReEncrypt(ct_c, rk_c) per ciphertext slot (ppq_agg_*): the reference's
    changeCipherDomain (server/src/changeCipherDomain.cpp:105) on every client's file followed by the EvalAdd chain of
    aggregateEncryptedWeights (server/src/aggregateEncryptedWeights.cpp:106), generalised from 2 to C clients.
    Bit-identical to cc.ReEncrypt + cc.EvalAdd per client; the linear tail of ModDown runs on sums."""

    def __init__(self, cc, ell, n_slots):
        self.cc, self.ell, self.n_slots = cc, int(ell), int(n_slots)
        h = vp()
        check(lib().ppq_agg_create(cc.h, self.ell, self.n_slots, C.byref(h)))
        self.h = h

    def add(self, ct, rekey):
        if ct.n != self.n_slots or ct.ell != self.ell:
            raise ValueError("Aggregator.add: batch shape differs from the aggregator's")
        check(lib().ppq_agg_add(self.h, rekey.h, _ptr(ct.data), _stream()))
        return self

    def add_many(self, cts, rekeys):
        """Several clients in one call (ppq_agg_add_many): same result as add() per client."""
        cts, rekeys = list(cts), list(rekeys)
        if len(cts) != len(rekeys):
            raise ValueError("Aggregator.add_many: one rekey per client batch")
        for ct in cts:
            if ct.n != self.n_slots or ct.ell != self.ell:
                raise ValueError("Aggregator.add_many: batch shape differs from the aggregator's")
        if cts:
            hk = (vp * len(cts))(*[r.h for r in rekeys])
            hp = (vp * len(cts))(*[_ptr(ct.data) for ct in cts])
            check(lib().ppq_agg_add_many(self.h, hk, hp, len(cts), _stream()))
        return self

    @property
    def count(self):
        return int(lib().ppq_agg_count(self.h))

    def finish(self, out=None, meta=None):
        if out is None:
            out = self.cc.empty_cts(self.n_slots, self.ell)
        check(lib().ppq_agg_finish(self.h, _ptr(out), _stream()))
        return CiphertextBatch(out, meta if meta is not None else CtMeta())

    def close(self):
        if getattr(self, "h", None):
            lib().ppq_agg_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()


class Communicator:
    """One NCCL communicator per rank for the cross-GPU sum (ppq_comm_*, ppq_reduce_across_ranks).  The 128-byte id
    made on rank 0 travels through whatever channel the host has (bench.py: torch.distributed broadcast)."""

    def __init__(self, cc, world, rank, unique_id):
        self.cc, self.world, self.rank = cc, int(world), int(rank)
        h = vp()
        check(lib().ppq_comm_init(cc.h, self.world, self.rank, bytes(unique_id), C.byref(h)))
        self.h = h

    @staticmethod
    def unique_id():
        buf = C.create_string_buffer(128)
        check(lib().ppq_comm_unique_id(buf))
        return buf.raw

    def reduce(self, acc, root=0):
        """acc: CiphertextBatch of canonical partial sums, summed in place onto `root` (root < 0: every rank)."""
        check(lib().ppq_reduce_across_ranks(self.cc.h, self.h, acc.ell, _ptr(acc.data), acc.n, int(root), _stream()))
        return acc

    def close(self):
        if getattr(self, "h", None):
            lib().ppq_comm_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()


class CryptoContext:
    def __init__(self, handle, device):
        self.h = handle
        self.device = torch.device("cuda", device)
        info = _capi.CtxInfo()
        check(lib().ppq_ctx_info(self.h, C.byref(info)))
        self.logN, self.L, self.K, self.dnum, self.alpha = info.logN, info.L, info.K, info.dnum, info.alpha
        self.N = 1 << self.logN
        self.M = self.L + self.K
        m = np.zeros(self.M, dtype=np.uint64)
        check(lib().ppq_ctx_moduli(self.h, m.ctypes.data_as(_capi.u64p)))
        self.moduli = m
        r = np.zeros(self.M, dtype=np.uint64)
        check(lib().ppq_ctx_roots(self.h, r.ctypes.data_as(_capi.u64p)))
        self.roots = r

    # ---- construction (Serial::DeserializeFromFile(cc_path, cc, SerType::JSON))
    @classmethod
    def from_ccjson(cls, text, device=0):
        raw = text.encode() if isinstance(text, str) else text
        h = vp()
        check(lib().ppq_ctx_create_from_ccjson(raw, len(raw), device, C.byref(h)))
        return cls(h, device)

    @classmethod
    def from_moduli(cls, logN, q_moduli, dnum, aux_bits=60, extra_bits=20, device=0):
        q = np.ascontiguousarray(q_moduli, dtype=np.uint64)
        h = vp()
        check(lib().ppq_ctx_create(logN, len(q), q.ctypes.data_as(_capi.u64p), dnum, aux_bits, extra_bits, device,
                                   C.byref(h)))
        return cls(h, device)

    @classmethod
    def synthetic(cls, logN, depth, first_bits=60, scale_bits=59, extra_bits=20, dnum=3, device=0):
        h = vp()
        check(lib().ppq_ctx_create_synthetic(logN, depth, first_bits, scale_bits, extra_bits, dnum, device, C.byref(h)))
        return cls(h, device)

    def close(self):
        if self.h:
            lib().ppq_ctx_destroy(self.h)
            self.h = None

    def sf(self, level):
        return lib().ppq_ctx_scaling_factor(self.h, level)

    def sf_big(self, level):
        return lib().ppq_ctx_scaling_factor_big(self.h, level)

    def num_parts(self, ell):
        return lib().ppq_ctx_num_parts(self.h, ell)

    def GetBatchSize(self):
        return self.N // 2

    def set_wave(self, n):
        check(lib().ppq_ctx_set_wave(self.h, int(n)))

    def sync(self):
        check(lib().ppq_sync(self.h, _stream()))

    def empty_cts(self, n, ell):
        return torch.empty((n, 2, ell, self.N), dtype=torch.int64, device=self.device)

    # ---- RNS polynomial layer
    def ntt(self, data, limb_ids, inverse=False):
        """data: CUDA int64 [n_polys][N], in place."""
        ids = np.ascontiguousarray(limb_ids, dtype=np.int32)
        assert data.is_cuda and data.is_contiguous() and data.numel() == len(ids) * self.N
        fn = lib().ppq_ntt_inverse if inverse else lib().ppq_ntt_forward
        for s in range(0, len(ids), 32768):
            chunk = ids[s:s + 32768]
            check(fn(self.h, vp(data.data_ptr() + 8 * s * self.N), chunk.ctypes.data_as(_capi.i32p), len(chunk),
                     _stream()))
        return data

    # ---- server side
    def ReEncrypt(self, ct, rekey, out=None):
        if out is None:
            out = torch.empty_like(ct.data)
        check(lib().ppq_reencrypt(self.h, rekey.h, ct.ell, _ptr(ct.data), _ptr(out), ct.n, _stream()))
        return CiphertextBatch(out, replace(ct.meta))

    def ReEncryptFanOut(self, ct, rekeys):
        """[ReEncrypt(ct, rk) for rk in rekeys] with the key-independent ModUp shared (ppq_reencrypt_fanout): the
        second changeCipherDomain pass of orchestration/server_fns.sh:76-80 for C-1 target domains."""
        rekeys = list(rekeys)
        if not rekeys:
            return []
        outs = [torch.empty_like(ct.data) for _ in rekeys]
        hk = (vp * len(rekeys))(*[r.h for r in rekeys])
        ho = (vp * len(rekeys))(*[_ptr(o) for o in outs])
        check(lib().ppq_reencrypt_fanout(self.h, hk, len(rekeys), ct.ell, _ptr(ct.data), ho, ct.n, _stream()))
        return [CiphertextBatch(o, replace(ct.meta)) for o in outs]

    def ReEncryptAccumulate(self, ct, rekey, acc):
        """acc += ReEncrypt(ct) (fused ReEncrypt + EvalAdd); acc: CiphertextBatch of the same shape."""
        check(lib().ppq_reencrypt_accumulate(self.h, rekey.h, ct.ell, _ptr(ct.data), _ptr(acc.data), ct.n, _stream()))
        return acc

    def KeySwitchModUp(self, c1):
        """EvalKeySwitchPrecomputeCore: c1 [n][ell][N] -> digits [n][parts][ell+K][N] (for parity tests)."""
        n, ell = c1.shape[0], c1.shape[1]
        out = torch.zeros((n, self.num_parts(ell), ell + self.K, self.N), dtype=torch.int64, device=self.device)
        check(lib().ppq_ks_modup(self.h, ell, _ptr(c1), _ptr(out), n, _stream()))
        return out

    def EvalAdd(self, a, b):
        if a.ell != b.ell or a.n != b.n:
            raise ValueError("EvalAdd: ciphertext batches differ in shape")
        if (a.meta.level, a.meta.noise_scale_deg) != (b.meta.level, b.meta.noise_scale_deg):
            raise ValueError("EvalAdd: operands at different level / noise scale degree (no automatic adjustment)")
        out = torch.empty_like(a.data)
        step = max(1, 65535 // (2 * a.ell))
        for s in range(0, a.n, step):
            e = min(a.n, s + step)
            check(lib().ppq_eval_add(self.h, a.ell, _ptr(a.data[s:e]), _ptr(b.data[s:e]), _ptr(out[s:e]), e - s,
                                     _stream()))
        return CiphertextBatch(out, replace(a.meta))

    def EvalMult(self, ct, c):
        """cc->EvalMult(ct, double) under FLEXIBLEAUTOEXT: rescale first when noiseScaleDeg == 2."""
        meta = replace(ct.meta)
        ell = ct.ell
        rescale = meta.noise_scale_deg == 2
        if rescale:
            meta.scaling_factor /= float(self.moduli[ell - 1])
            meta.level += 1
            meta.noise_scale_deg -= 1
        sf = self.sf(meta.level)
        lo = ell - 1 if rescale else ell
        out = torch.empty((ct.n, 2, lo, self.N), dtype=torch.int64, device=self.device)
        step = 8192
        for s in range(0, ct.n, step):
            e = min(ct.n, s + step)
            check(lib().ppq_eval_mult_const(self.h, ell, _ptr(ct.data[s:e]), _ptr(out[s:e]), e - s, float(c), sf,
                                            int(rescale), _stream()))
        meta.noise_scale_deg += 1
        meta.scaling_factor *= sf
        return CiphertextBatch(out, meta)

    def finalize_sum(self, acc):
        step = max(1, 65535 // (2 * acc.ell))
        for s in range(0, acc.n, step):
            e = min(acc.n, s + step)
            check(lib().ppq_finalize_sum(self.h, acc.ell, _ptr(acc.data[s:e]), e - s, _stream()))
        return acc

    # ---- client side
    def MakeCKKSPackedPlaintext(self, values, ell=None):
        """values: [n][<= N/2] real numbers (numpy or CUDA float64 tensor)."""
        ell = self.L if ell is None else ell
        v = values if torch.is_tensor(values) else torch.from_numpy(np.ascontiguousarray(values, dtype=np.float64))
        v = v.to(self.device, dtype=torch.float64).contiguous()
        if v.dim() == 1:
            v = v[None, :]
        n, nv = v.shape
        sf = self.sf_big(0)
        pt = torch.empty((n, ell, self.N), dtype=torch.int64, device=self.device)
        step = max(1, 65535 // ell)
        for s in range(0, n, step):
            e = min(n, s + step)
            check(lib().ppq_encode(self.h, ell, _ptr(v[s:e]), nv, sf, _ptr(pt[s:e]), e - s, _stream()))
        meta = CtMeta(noise_scale_deg=2, level=0, scaling_factor=sf, slots=self.N // 2)
        return PlaintextBatch(data=pt, meta=meta)

    def Encrypt(self, pk, pt, small=None, seed=None):
        """pk = (b, a) CUDA int64 [>=ell][N] each (QP public key; only the first ell limbs are used).
        small: optional CUDA int64 [n][3][N] = injected (v, e0, e1) samples; sampled on the device otherwise, under a
        fresh OS-entropy key per call unless `seed` (tests only: a fixed seed repeats the same noise) is given."""
        b, a = pk
        n, ell = pt.data.shape[0], pt.data.shape[1]
        if small is None:
            small = torch.empty((n, 3, self.N), dtype=torch.int64, device=self.device)
            for s0 in range(0, n, 16384):  # one key per call of the library: chunks of a seeded draw get distinct seeds
                e0 = min(n, s0 + 16384)
                sd = _seed32(seed)
                if sd is not None and s0:
                    sd = _seed32((int.from_bytes(sd, "little") + s0) % (1 << 256))
                check(lib().ppq_sample_encrypt_noise(self.h, sd, _ptr(small[s0:e0]), e0 - s0, _stream()))
        ct = self.empty_cts(n, ell)
        step = max(1, 65535 // (3 * ell))
        for s in range(0, n, step):
            e = min(n, s + step)
            check(lib().ppq_encrypt(self.h, ell, _ptr(b), _ptr(a), _ptr(pt.data[s:e]), _ptr(small[s:e]), _ptr(ct[s:e]),
                                    e - s, _stream()))
        return CiphertextBatch(ct, replace(pt.meta))

    def EncodeEncrypt(self, pk, values, ell=None, small=None, seed=None):
        """MakeCKKSPackedPlaintext + Encrypt in one call (client/src/encryptModelWeights.cpp:109-110 per chunk): the
        encoded coefficients join e0 before the transform, 3*ell instead of 4*ell limb transforms per ciphertext; the
        result is bit-identical to Encrypt(pk, MakeCKKSPackedPlaintext(values), small)."""
        b, a = pk
        ell = self.L if ell is None else ell
        v = values if torch.is_tensor(values) else torch.from_numpy(np.ascontiguousarray(values, dtype=np.float64))
        v = v.to(self.device, dtype=torch.float64).contiguous()
        if v.dim() == 1:
            v = v[None, :]
        n, nv = v.shape
        sf = self.sf_big(0)
        if small is None:
            small = torch.empty((n, 3, self.N), dtype=torch.int64, device=self.device)
            for s0 in range(0, n, 16384):
                e0 = min(n, s0 + 16384)
                sd = _seed32(seed)
                if sd is not None and s0:
                    sd = _seed32((int.from_bytes(sd, "little") + s0) % (1 << 256))
                check(lib().ppq_sample_encrypt_noise(self.h, sd, _ptr(small[s0:e0]), e0 - s0, _stream()))
        ct = self.empty_cts(n, ell)
        check(lib().ppq_encode_encrypt(self.h, ell, _ptr(b), _ptr(a), _ptr(v), nv, sf, _ptr(small), _ptr(ct), n, _stream()))
        return CiphertextBatch(ct, CtMeta(noise_scale_deg=2, level=0, scaling_factor=sf, slots=self.N // 2))

    def KeyGen(self, seed=None, samples=None):
        """cc->KeyGen() (client/src/keyGen.cpp:33).  Returns (sk [L][N], (pk_b, pk_a) [L+K][N] each), CUDA int64, EVAL.
        samples: optional (s [N] int64, e [N] int64, a [L+K][N]) CUDA tensors injected for parity tests; drawn on
        the device otherwise (fresh OS entropy per call; `seed` is for tests only)."""
        if samples is None:
            s = torch.empty(self.N, dtype=torch.int64, device=self.device)
            e = torch.empty_like(s)
            a = torch.empty((self.M, self.N), dtype=torch.int64, device=self.device)
            check(lib().ppq_sample_keygen(self.h, _seed32(seed), _ptr(s), _ptr(e), _ptr(a), _stream()))
        else:
            s, e, a = samples
            if tuple(s.shape) != (self.N,) or tuple(e.shape) != (self.N,) or tuple(a.shape) != (self.M, self.N):
                raise ValueError("KeyGen: samples must be (s [N], e [N], a [L+K][N])")
        sk = torch.empty((self.L, self.N), dtype=torch.int64, device=self.device)
        pkb = torch.empty((self.M, self.N), dtype=torch.int64, device=self.device)
        check(lib().ppq_keygen(self.h, _ptr(s), _ptr(e), _ptr(a), _ptr(sk), _ptr(pkb), _stream()))
        return sk, (pkb, a)

    def ReKeyGen(self, sk_old, pk_new, seed=None, small=None):
        """cc->ReKeyGen(oldPrivateKey, newPublicKey) (client/src/REkeyGen.cpp:52) -> EvalKey resident on the device.
        small: optional CUDA int64 [dnum][3][N] = injected (u, e0, e1) per digit."""
        b, a = pk_new
        if tuple(sk_old.shape) != (self.L, self.N) or tuple(b.shape) != (self.M, self.N) or tuple(a.shape) != tuple(b.shape):
            raise ValueError("ReKeyGen: sk_old must be [L][N] and the public key two [L+K][N] tensors")
        if small is None:
            small = torch.empty((self.dnum, 3, self.N), dtype=torch.int64, device=self.device)
            check(lib().ppq_sample_encrypt_noise(self.h, _seed32(seed), _ptr(small), self.dnum, _stream()))
        elif tuple(small.shape) != (self.dnum, 3, self.N):
            raise ValueError("ReKeyGen: small must be [dnum][3][N]")
        kb = torch.empty((self.dnum, self.M, self.N), dtype=torch.int64, device=self.device)
        ka = torch.empty_like(kb)
        check(lib().ppq_rekeygen(self.h, _ptr(sk_old), _ptr(b), _ptr(a), _ptr(small), _ptr(kb), _ptr(ka), _stream()))
        return EvalKey(self, kb, ka)

    def DecryptCore(self, sk, ct):
        co = torch.empty((ct.n, ct.ell, self.N), dtype=torch.int64, device=self.device)
        step = max(1, 65535 // ct.ell)
        for s in range(0, ct.n, step):
            e = min(ct.n, s + step)
            check(lib().ppq_decrypt_core(self.h, ct.ell, _ptr(sk), _ptr(ct.data[s:e]), _ptr(co[s:e]), e - s, _stream()))
        return co

    def Decrypt(self, sk, ct):
        co = self.DecryptCore(sk, ct)
        vals = torch.empty((ct.n, self.N // 2), dtype=torch.float64, device=self.device)
        step = 8192
        for s in range(0, ct.n, step):
            e = min(ct.n, s + step)
            check(lib().ppq_decode(self.h, ct.ell, _ptr(co[s:e]), float(ct.meta.scaling_factor), _ptr(vals[s:e]), e - s,
                                   _stream()))
        return PlaintextBatch(values=vals, meta=replace(ct.meta))
