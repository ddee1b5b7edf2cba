#!/usr/bin/env python
"""Generate the committed golden fixtures from the REAL artefacts in /root/reference.

Run here (where /root/reference exists); the fixtures travel to the GPU box, the reference does not.

  CC.json                 verbatim copy of server/storage/CC.json (a data file: the serialized OpenFHE
                          CryptoContext that pins N, the 4 Q primes, their 2N-th roots, dnum, techniques)
  openfhe_sk_golden.npz   for each client's real OpenFHE secret key (client_i-private.key, cereal JSON,
                          4 limbs in EVALUATION format): the ternary coefficient vector recovered by the
                          oracle's inverse NTT (identical on all four limbs), the SHA-256 of every serialized
                          limb (little-endian u64 bytes), and the first 8 residues of every limb.
                          Test: NTT(ternary) on limb i must hash to the stored digest  => pins root choice,
                          bit-reversed EVAL ordering and limb consistency against real OpenFHE output.
  weights_golden.npz      sample_weights_c{1,2}.json (float32-exact values, float64 mean/std) and the
                          reference's own decrypted_weights_c{1,2}.json for the same round
                          (end-to-end CKKS tolerance golden: decrypted == (w1+w2)/2 within 1.4e-8).
"""
import hashlib
import json
import os
import shutil
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
import oracle  # noqa: E402

REF = "/root/reference"


def main():
    shutil.copyfile(f"{REF}/server/storage/CC.json", f"{HERE}/CC.json")
    o = oracle.Oracle.from_ccjson(open(f"{HERE}/CC.json").read())

    sk = {}
    for c in (1, 2):
        d = json.load(open(f"{REF}/client/storage/client_{c}/private/client_{c}-private.key"))
        s = d["value0"]["ptr_wrapper"]["data"]["s"]
        assert s["f"] == 0  # EVALUATION
        limbs = [np.array(t["v"]["ptr_wrapper"]["data"]["v"], dtype=np.uint64) for t in s["v"]]
        mods = [int(t["v"]["ptr_wrapper"]["data"]["m"]["v"]) for t in s["v"]]
        assert mods == [int(x) for x in o.moduli[:4]]
        tern = None
        digests, heads = [], []
        for i, l in enumerate(limbs):
            co = o.ntt(i, l, inverse=True)
            q = int(o.moduli[i])
            small = np.where(co > q // 2, co.astype(np.int64) - q, co.astype(np.int64))
            assert set(np.unique(small)) <= {-1, 0, 1}
            if tern is None:
                tern = small
            assert np.array_equal(tern, small)
            digests.append(hashlib.sha256(l.astype("<u8").tobytes()).hexdigest())
            heads.append(l[:8])
        sk[f"c{c}_ternary"] = tern.astype(np.int8)
        sk[f"c{c}_sha256"] = np.array(digests)
        sk[f"c{c}_head"] = np.stack(heads)
    np.savez_compressed(f"{HERE}/openfhe_sk_golden.npz", **sk)

    w = {}
    for c in (1, 2):
        for kind in ("sample", "decrypted"):
            d = json.load(open(f"{REF}/client/storage/client_{c}/private/{kind}_weights_c{c}.json"))
            layers = d["weights_summary"]
            vals = [np.array(t["values"], dtype=np.float64).ravel() for t in layers]
            flat = np.concatenate(vals)
            if kind == "sample":
                assert np.array_equal(flat, flat.astype(np.float32).astype(np.float64))
                flat = flat.astype(np.float32)
            w[f"c{c}_{kind}_values"] = flat
            w[f"c{c}_{kind}_mean"] = np.array([t["mean"] for t in layers], dtype=np.float64)
            w[f"c{c}_{kind}_std"] = np.array([t["std_dev"] for t in layers], dtype=np.float64)
            if c == 1 and kind == "sample":
                w["layers"] = np.array([t["layer"] for t in layers])
                w["shapes"] = np.array([json.dumps(t["shape"]) for t in layers])
                w["sizes"] = np.array([len(v) for v in vals], dtype=np.int64)
    np.savez_compressed(f"{HERE}/weights_golden.npz", **w)
    for f in ("CC.json", "openfhe_sk_golden.npz", "weights_golden.npz"):
        print(f, os.path.getsize(f"{HERE}/{f}"))


if __name__ == "__main__":
    main()
