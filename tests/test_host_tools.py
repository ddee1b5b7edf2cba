"""The four drop-in tools under host/ (C++ over the C ABI) and their file-format code.

CPU part: wire-format self tests, the real OpenFHE private key of the reference parsed by the C++ reader, argv /
exit-code contract, and "no CPU fallback" (without a device the tools must exit 1 at the context load).
GPU part: the reference's round, run through the BINARIES exactly as orchestration/server_fns.sh:62-80 and
client_fns.sh:75-102 call them: encrypt x2 -> changeCipherDomain(1->2) -> aggregate -> changeCipherDomain(2->1) ->
decrypt x2, compared with decrypted_weights_c{1,2}.json within the CKKS tolerance 2^-20.
"""
import json
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "host", "bin")
GOLD = os.path.join(ROOT, "tests", "golden")
REF_SK = "/root/reference/client/storage/client_1/private/client_1-private.key"


@pytest.fixture(scope="module")
def tools():
    import ppqsflhe_b200
    ppqsflhe_b200.build()
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "host")], stdout=subprocess.DEVNULL)
    return BIN


def run(*a, **kw):
    return subprocess.run(list(a), capture_output=True, text=True, **kw)


def fnv(limb):
    h = 1469598103934665603
    for b in np.ascontiguousarray(limb, dtype="<u8").tobytes():
        h = ((h ^ b) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    return h


@pytest.mark.parametrize("mode", ["base64", "blob", "json"])
def test_wire_selftests(tools, mode):
    r = run(os.path.join(tools, "wire_selftest"), mode)
    assert r.returncode == 0, r.stderr


def _sk_eval(o, c):
    tern = np.load(os.path.join(GOLD, "openfhe_sk_golden.npz"))[f"c{c}_ternary"].astype(np.int64)
    return np.stack([o.small_to_eval(r, tern) for r in range(o.M)])


def test_key_files_round_trip_through_the_cpp_reader(tools, tmp_path, oracle_c1):
    """mkkeys writes OpenFHE's cereal-JSON DCRTPoly element layout; privkey reads it back limb for limb."""
    o = oracle_c1
    sk = _sk_eval(o, 1)[:4]
    spec = {"N": o.N, "moduli": [int(x) for x in o.moduli[:4]], "sk": [[int(x) for x in l] for l in sk]}
    (tmp_path / "in.json").write_text(json.dumps(spec))
    assert run(os.path.join(tools, "wire_selftest"), "mkkeys", str(tmp_path / "in.json"), str(tmp_path)).returncode == 0
    r = run(os.path.join(tools, "wire_selftest"), "privkey", str(tmp_path / "private.key"))
    assert r.returncode == 0, r.stderr
    lines = r.stdout.strip().splitlines()
    assert lines[0] == f"limbs 4 N {o.N} format 0"
    for i in range(4):
        assert lines[1 + i] == f"limb {i} q {int(o.moduli[i])} fnv {fnv(sk[i])}"


@pytest.mark.skipif(not os.path.exists(REF_SK), reason="reference tree not present on this machine")
def test_real_openfhe_private_key_parses_to_the_golden_limbs(tools, oracle_c1):
    """The C++ cereal-JSON reader on a REAL OpenFHE file: limbs must equal NTT(ternary golden) on all four primes."""
    o = oracle_c1
    r = run(os.path.join(tools, "wire_selftest"), "privkey", REF_SK)
    assert r.returncode == 0, r.stderr
    sk = _sk_eval(o, 1)
    lines = r.stdout.strip().splitlines()
    assert lines[0] == "limbs 4 N 16384 format 0"
    for i in range(4):
        assert lines[1 + i] == f"limb {i} q {int(o.moduli[i])} fnv {fnv(sk[i])}"


@pytest.mark.parametrize("tool,usage", [
    ("changeCipherDomain", "<cc_path> <rekey_path> <input_encfile> <output_encfile>"),
    ("aggregateEncryptedWeights", "<cc_path> <client2_encfile> <client1to2_encfile> <output_aggfile>"),
    ("encryptModelWeights", "<cc_path> <pubkey_path> <input_weights> <output_encfile>"),
    ("decryptModelWeights", "<cc_path> <privkey_path> <input_encfile> <output_file>"),
    ("keyGen", "<cc_path> <pubkey_out> <privkey_out>"),  # client/src/keyGen.cpp:14-18
    ("REkeyGen", "<cc.json> <client_privkey.json> <peer_pubkey.json> <rekey_out.json>"),  # client/src/REkeyGen.cpp:15-20
])
def test_argv_contract(tools, tool, usage):
    r = run(os.path.join(tools, tool), "only-one-arg")
    assert r.returncode == 1 and "Usage:" in r.stderr and usage in r.stderr


def test_tools_fail_loudly_without_a_device(tools, tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    r = run(os.path.join(tools, "changeCipherDomain"), os.path.join(GOLD, "CC.json"), "nokey", "noin", str(tmp_path / "o.json"))
    assert r.returncode == 1 and "Failed to load CryptoContext" in r.stderr and "no CUDA device" in r.stderr
    assert not (tmp_path / "o.json").exists()
    r = run(os.path.join(tools, "changeCipherDomain"), str(tmp_path / "missing.json"), "a", "b", "c")
    assert r.returncode == 1 and "[recrypt] ERROR" in r.stderr


# ------------------------------------------------------------------------------------------------ GPU: the full round
def _weights_json(path, g, c):
    layers, shapes, sizes = g["layers"], g["shapes"], g["sizes"]
    vals = g[f"c{c}_sample_values"].astype(np.float64)
    out, off = [], 0
    for i in range(len(layers)):
        n = int(sizes[i])
        out.append({"layer": str(layers[i]), "shape": json.loads(str(shapes[i])),
                    "mean": float(g[f"c{c}_sample_mean"][i]), "std_dev": float(g[f"c{c}_sample_std"][i]),
                    "values": [float(x) for x in vals[off:off + n]]})
        off += n
    out.append({"layer": "optimizer/iter", "shape": [1], "mean": 0.0, "std_dev": 0.0, "values": [7.0]})  # must be skipped
    json.dump({"weights_summary": out}, open(path, "w"))
    return out[:-1]


@pytest.mark.gpu
def test_reference_round_through_the_binaries(tools, tmp_path, oracle_c1, ccjson_text):
    import oracle
    o = oracle_c1
    g = np.load(os.path.join(GOLD, "weights_golden.npz"))
    cc_path = tmp_path / "CC.json"
    cc_path.write_text(ccjson_text)
    # scaffolding keys (oracle KeyGen/ReKeyGen with the REAL OpenFHE secrets), written as cereal-JSON key files
    keys = {}
    for c in (1, 2):
        _, _, _, pka = o.keygen(40 + c)
        sk = _sk_eval(o, c)
        e = oracle.sample_gauss(900 + c, o.N)
        pkb = np.stack([(o.small_to_eval(r, e).astype(object) - pka[r].astype(object) * sk[r].astype(object)) % int(o.moduli[r])
                        for r in range(o.M)]).astype(np.uint64)
        keys[c] = (sk, pkb, pka)
    rk = {(1, 2): o.rekeygen(51, keys[1][0], keys[2][1], keys[2][2]), (2, 1): o.rekeygen(52, keys[2][0], keys[1][1], keys[1][2])}
    mod = [int(x) for x in o.moduli]

    def tolist(a):
        return a.tolist()

    for c in (1, 2):
        d = tmp_path / f"c{c}"
        d.mkdir()
        other = 2 if c == 1 else 1
        kb, ka = rk[(c, other)]
        spec = {"N": o.N, "moduli": mod, "sk": tolist(keys[c][0][:o.L]), "pk_b": tolist(keys[c][1]), "pk_a": tolist(keys[c][2]),
                "rk_b": tolist(kb), "rk_a": tolist(ka)}
        (d / "in.json").write_text(json.dumps(spec))
        assert run(os.path.join(tools, "wire_selftest"), "mkkeys", str(d / "in.json"), str(d)).returncode == 0
        plain = _weights_json(d / "weights.json", g, c)
    T = lambda name: os.path.join(tools, name)  # noqa: E731
    env = dict(os.environ, PPQ_SEED="1234")

    def ok(*a):
        r = run(*a, env=env)
        assert r.returncode == 0, r.stdout + r.stderr
        return r.stdout

    for c in (1, 2):
        out = ok(T("encryptModelWeights"), str(cc_path), str(tmp_path / f"c{c}/public.key"), str(tmp_path / f"c{c}/weights.json"),
                 str(tmp_path / f"enc_c{c}.json"))
        assert "Skipping optimizer layer: optimizer/iter" in out and "Batch size from CryptoContext = 8192" in out
    enc = json.load(open(tmp_path / "enc_c1.json"))
    assert list(enc) == ["weights_summary"] and len(enc["weights_summary"]) == 8
    assert sorted(enc["weights_summary"][0]) == ["layer", "mean", "shape", "std_dev", "values"]
    n_ct = sum(2 + len(w["values"]) for w in enc["weights_summary"])
    assert n_ct == 27  # 8 tensors x {mean, std_dev} + 11 value chunks
    # orchestration/server_fns.sh:62-80
    assert "[recrypt] Re-encryption completed successfully" in ok(
        T("changeCipherDomain"), str(cc_path), str(tmp_path / "c1/rekey.key"), str(tmp_path / "enc_c1.json"), str(tmp_path / "c1_to_c2.json"))
    assert "[agg] Aggregation completed successfully" in ok(
        T("aggregateEncryptedWeights"), str(cc_path), str(tmp_path / "enc_c2.json"), str(tmp_path / "c1_to_c2.json"), str(tmp_path / "agg.json"))
    ok(T("changeCipherDomain"), str(cc_path), str(tmp_path / "c2/rekey.key"), str(tmp_path / "agg.json"), str(tmp_path / "agg_to_c1.json"))
    # wire sizes: fresh ct = 2 x 4 limbs, aggregated = 2 x 3 limbs (comm_metrics.csv:10-13 ratio 37.8 MB : 28.4 MB)
    s_enc, s_agg = os.path.getsize(tmp_path / "enc_c1.json"), os.path.getsize(tmp_path / "agg.json")
    assert abs(s_agg / s_enc - 0.75) < 0.01
    # client_fns.sh:90-102
    ok(T("decryptModelWeights"), str(cc_path), str(tmp_path / "c2/private.key"), str(tmp_path / "agg.json"), str(tmp_path / "dec_c2.json"))
    ok(T("decryptModelWeights"), str(cc_path), str(tmp_path / "c1/private.key"), str(tmp_path / "agg_to_c1.json"), str(tmp_path / "dec_c1.json"))
    w1, w2 = g["c1_sample_values"].astype(np.float64), g["c2_sample_values"].astype(np.float64)
    for c in (1, 2):
        dec = json.load(open(tmp_path / f"dec_c{c}.json"))["weights_summary"]
        assert [d["layer"] for d in dec] == [p["layer"] for p in plain]
        assert [d["shape"] for d in dec] == [p["shape"] for p in plain]
        vals = np.concatenate([np.asarray(d["values"]) for d in dec])
        assert vals.shape == w1.shape
        assert np.max(np.abs(vals - (w1 + w2) / 2)) < 2.0 ** -20
        assert np.max(np.abs(vals - g[f"c{c}_decrypted_values"])) < 2.0 ** -20
        means = np.array([d["mean"] for d in dec])
        assert np.max(np.abs(means - (g["c1_sample_mean"] + g["c2_sample_mean"]) / 2)) < 2.0 ** -20
    # a truncated blob must be rejected with exit 1, not decrypted into garbage
    bad = json.load(open(tmp_path / "agg.json"))
    bad["weights_summary"][0]["mean"] = bad["weights_summary"][0]["mean"][:1000]
    json.dump(bad, open(tmp_path / "bad.json", "w"))
    r = run(T("decryptModelWeights"), str(cc_path), str(tmp_path / "c2/private.key"), str(tmp_path / "bad.json"), str(tmp_path / "x.json"))
    assert r.returncode == 1 and "[decrypt] ERROR" in r.stderr


@pytest.mark.gpu
def test_round_with_keys_made_by_the_binaries(tools, tmp_path, ccjson_text):
    """The whole client + server flow with nothing fabricated: keyGen x2 (client_fns.sh key setup), REkeyGen 1->2,
    encryptModelWeights, changeCipherDomain, aggregateEncryptedWeights, decryptModelWeights under client 2's key."""
    g = np.load(os.path.join(GOLD, "weights_golden.npz"))
    cc_path = tmp_path / "CC.json"
    cc_path.write_text(ccjson_text)
    T = lambda name: os.path.join(tools, name)  # noqa: E731

    def ok(seed, *a):
        r = run(*a, env=dict(os.environ, PPQ_SEED=str(seed)))
        assert r.returncode == 0, r.stdout + r.stderr
        return r.stdout

    for c in (1, 2):
        (tmp_path / f"c{c}").mkdir()
        out = ok(7000 + c, T("keyGen"), str(cc_path), str(tmp_path / f"c{c}/public.key"), str(tmp_path / f"c{c}/private.key"))
        assert "[keyGen] Public and Private keys generated" in out and "[keyGen] Keys saved" in out
        _weights_json(tmp_path / f"c{c}/weights.json", g, c)
    # the private key file parses with the reader that parses the reference's real OpenFHE key, and is ternary
    r = run(T("wire_selftest"), "privkey", str(tmp_path / "c1/private.key"))
    assert r.returncode == 0 and r.stdout.splitlines()[0] == "limbs 4 N 16384 format 0"
    out = ok(7100, T("REkeyGen"), str(cc_path), str(tmp_path / "c1/private.key"), str(tmp_path / "c2/public.key"), str(tmp_path / "c1/rekey.key"))
    assert "[ReKeyGen] Re-encryption key generated successfully" in out and "[ReKeyGen] Re-encryption key saved" in out
    for c in (1, 2):
        ok(7200 + c, T("encryptModelWeights"), str(cc_path), str(tmp_path / f"c{c}/public.key"), str(tmp_path / f"c{c}/weights.json"),
           str(tmp_path / f"enc_c{c}.json"))
    ok(0, T("changeCipherDomain"), str(cc_path), str(tmp_path / "c1/rekey.key"), str(tmp_path / "enc_c1.json"), str(tmp_path / "c1_to_c2.json"))
    ok(0, T("aggregateEncryptedWeights"), str(cc_path), str(tmp_path / "enc_c2.json"), str(tmp_path / "c1_to_c2.json"), str(tmp_path / "agg.json"))
    ok(0, T("decryptModelWeights"), str(cc_path), str(tmp_path / "c2/private.key"), str(tmp_path / "agg.json"), str(tmp_path / "dec_c2.json"))
    w1, w2 = g["c1_sample_values"].astype(np.float64), g["c2_sample_values"].astype(np.float64)
    dec = json.load(open(tmp_path / "dec_c2.json"))["weights_summary"]
    vals = np.concatenate([np.asarray(d["values"]) for d in dec])
    assert np.max(np.abs(vals - (w1 + w2) / 2)) < 2.0 ** -20
    # wrong-domain guard: REkeyGen refuses a public key file in place of the private key
    r = run(T("REkeyGen"), str(cc_path), str(tmp_path / "c2/public.key"), str(tmp_path / "c2/public.key"), str(tmp_path / "x.key"))
    assert r.returncode == 1 and "Error loading Client private key" in r.stderr
