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


def test_cereal_ciphertext_codec_selftest(tools):
    """ppq_cereal.hpp: self round trip at every limb count / sharing pattern, damaged blobs rejected, and the two byte
    counts the reference's recorded file sizes pin (orchestration/metrics/comm_metrics.csv:10-13): 1553 bytes around a
    fresh 4-limb ciphertext, 1551 around an aggregated 3-limb one."""
    r = run(os.path.join(tools, "wire_selftest"), "cereal", os.path.join(GOLD, "CC.json"))
    assert r.returncode == 0, r.stdout + r.stderr
    assert "context 676 bytes, 9 pointer ids" in r.stdout


def test_gencc_reproduces_the_reference_cc_json_byte_for_byte(tools, tmp_path):
    """host/genCC on the reference's own config (server/config/config_cc.json: depth 2, scaling 40, batch 8192, INDCPA)
    must write server/storage/CC.json (= tests/golden/CC.json, 15,199 bytes) exactly: ring dimension from the security
    table, the four primes and roots, the 160-bit composite modulus, cereal's field order, ids and number formatting."""
    cfg = tmp_path / "config_cc.json"
    cfg.write_text(json.dumps({"MultiplicativeDepth": 2, "ScalingModSize": 40, "BatchSize": 8192, "PREMode": "INDCPA"}))
    out = tmp_path / "CC.json"
    r = run(os.path.join(tools, "genCC"), str(cfg), str(out))
    assert r.returncode == 0 and "CryptoContext Generated and saved to" in r.stdout, r.stderr
    want = open(os.path.join(GOLD, "CC.json"), "rb").read()
    assert len(want) == 15199 and out.read_bytes() == want
    # the reference's FIXED_NOISE quirk (genCC.cpp:59-60 -> FIXEDMANUAL) is refused, unknown modes exit 1 like the reference
    cfg.write_text(json.dumps({"MultiplicativeDepth": 2, "ScalingModSize": 40, "PREMode": "FIXED_NOISE"}))
    assert run(os.path.join(tools, "genCC"), str(cfg), str(out)).returncode == 1
    cfg.write_text(json.dumps({"PREMode": "nope"}))
    r = run(os.path.join(tools, "genCC"), str(cfg), str(out))
    assert r.returncode == 1 and "Unknown PREMode in config: nope" in r.stderr
    r = run(os.path.join(tools, "genCC"), str(tmp_path / "missing.json"), str(out))
    assert r.returncode == 1 and "Failed to open" in r.stderr
    # a deeper context: loads back through the library's own CC.json reader with the chain the library derives
    cfg.write_text(json.dumps({"MultiplicativeDepth": 6, "ScalingModSize": 50, "BatchSize": 16384, "PREMode": "INDCPA"}))
    assert run(os.path.join(tools, "genCC"), str(cfg), str(out)).returncode == 0
    import oracle
    cc = oracle.parse_ccjson(out.read_text())
    assert cc["logN"] == 15 and len(cc["moduli"]) == 8 and cc["dnum"] == 3


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
    # ... and, with the cereal PortableBinary blobs (default), the files have EXACTLY the sizes the reference recorded for
    # this model (comm_metrics.csv:10-13): 37,806,011 bytes encrypted, 28,368,755 bytes aggregated / domain-changed
    assert s_enc == 37806011 and os.path.getsize(tmp_path / "enc_c2.json") == 37806011
    assert s_agg == 28368755 and os.path.getsize(tmp_path / "agg_to_c1.json") == 28368755
    # the private PPQCT1 blob is still written on request and read back by the same tools
    r = run(T("changeCipherDomain"), str(cc_path), str(tmp_path / "c2/rekey.key"), str(tmp_path / "agg.json"), str(tmp_path / "agg_p.json"),
            env=dict(env, PPQ_BLOB="ppqct1"))
    assert r.returncode == 0 and os.path.getsize(tmp_path / "agg_p.json") < 28368755
    ok(T("decryptModelWeights"), str(cc_path), str(tmp_path / "c1/private.key"), str(tmp_path / "agg_p.json"), str(tmp_path / "dec_p.json"))
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


class _Served:
    """a resident ppq_served process driven over its stdin/stdout line protocol"""

    def __init__(self, tools, cc_path, env):
        self.p = subprocess.Popen([os.path.join(tools, "ppq_served"), str(cc_path)], stdin=subprocess.PIPE, stdout=subprocess.PIPE,
                                  stderr=subprocess.PIPE, text=True, env=env)
        seen = []
        while True:
            line = self.p.stdout.readline()
            assert line, "ppq_served died: " + self.p.stderr.read()
            seen.append(line)
            if line.startswith("[served] ready"):
                break

    def ask(self, line):
        self.p.stdin.write(line + "\n")
        self.p.stdin.flush()
        r = self.p.stdout.readline().strip()
        return r

    def close(self):
        r = self.ask("QUIT")
        self.p.wait(timeout=30)
        return r


def _kv(reply):
    return {k: v for k, v in (t.split("=", 1) for t in reply.split()[1:] if "=" in t)}


@pytest.mark.gpu
def test_resident_server_three_clients_two_rounds(tools, tmp_path, ccjson_text):
    """host/ppq_served: context and re-encryption keys parsed ONCE, then whole rounds per request --
    for 3 clients (the reference hard-codes 2): clients 1, 2 -> domain 3, + client 3's own, x 1/3 with rescale, fan-out
    of the aggregate back to domains 1 and 2.  Round 2 runs without any key re-parse.  The 2-client request is checked
    against the tool-by-tool flow of orchestration/server_fns.sh:62-80 (decrypted outputs identical)."""
    g = np.load(os.path.join(GOLD, "weights_golden.npz"))
    cc_path = tmp_path / "CC.json"
    cc_path.write_text(ccjson_text)
    T = lambda name: os.path.join(tools, name)  # noqa: E731

    def ok(seed, *a):
        r = run(*a, env=dict(os.environ, PPQ_SEED=str(seed), PPQ_QUIET="1"))
        assert r.returncode == 0, r.stdout + r.stderr
        return r.stdout

    w = {}
    for c in (1, 2, 3):
        d = tmp_path / f"c{c}"
        d.mkdir()
        ok(8000 + c, T("keyGen"), str(cc_path), str(d / "public.key"), str(d / "private.key"))
        src = 1 if c != 2 else 2
        plain = _weights_json(d / "weights.json", g, src)
        if c == 3:  # a third model: client 1's weights negated and shifted
            doc = json.load(open(d / "weights.json"))
            for t in doc["weights_summary"]:
                t["values"] = [0.125 - v for v in t["values"]]
                t["mean"] = 0.125 - t["mean"]
            json.dump(doc, open(d / "weights.json", "w"))
        doc = json.load(open(d / "weights.json"))["weights_summary"][:-1]
        w[c] = np.concatenate([np.asarray(t["values"]) for t in doc])
    for a, b in ((1, 3), (2, 3), (3, 1), (3, 2), (1, 2), (2, 1)):
        ok(8100 + 10 * a + b, T("REkeyGen"), str(cc_path), str(tmp_path / f"c{a}/private.key"), str(tmp_path / f"c{b}/public.key"),
           str(tmp_path / f"rk_{a}_{b}.key"))
    for c in (1, 2, 3):
        ok(8200 + c, T("encryptModelWeights"), str(cc_path), str(tmp_path / f"c{c}/public.key"), str(tmp_path / f"c{c}/weights.json"),
           str(tmp_path / f"enc_c{c}.json"))
    sv = _Served(tools, cc_path, dict(os.environ, PPQ_QUIET="1"))
    try:
        for a, b in ((1, 3), (2, 3), (3, 1), (3, 2), (1, 2), (2, 1)):
            r = sv.ask(f"LOAD_REKEY rk{a}{b} {tmp_path}/rk_{a}_{b}.key")
            assert r.startswith("OK rekey"), r
        assert sv.ask("ROUND target=nofile out=x").startswith("ERR"), "a bad request must not kill the server"
        assert sv.ask(f"ROUND target={tmp_path}/enc_c3.json out={tmp_path}/x.json up=nokey:{tmp_path}/enc_c1.json").startswith("ERR unknown rekey")
        rounds = []
        for rnd in (1, 2):
            r = sv.ask(f"ROUND target={tmp_path}/enc_c3.json out={tmp_path}/agg3_r{rnd}.json up=rk13:{tmp_path}/enc_c1.json "
                       f"up=rk23:{tmp_path}/enc_c2.json down=rk31:{tmp_path}/agg1_r{rnd}.json down=rk32:{tmp_path}/agg2_r{rnd}.json")
            assert r.startswith("OK clients=3 cts=27 limbs_out=3"), r
            rounds.append(_kv(r))
        st = _kv(sv.ask("STATS"))
        assert st["rounds"] == "2" and st["rekeys"] == "6" and st["key_parses"] == "6", st  # no key parsed after the LOAD_REKEYs
        # 2-client request == the reference's three tool invocations
        r = sv.ask(f"ROUND target={tmp_path}/enc_c2.json out={tmp_path}/sv_agg.json up=rk12:{tmp_path}/enc_c1.json "
                   f"down=rk21:{tmp_path}/sv_agg_to_c1.json")
        assert r.startswith("OK clients=2 cts=27"), r
    finally:
        assert sv.close() == "OK bye"
    want = (w[1] + w[2] + w[3]) / 3
    for c in (1, 2, 3):
        for rnd in (1, 2):
            ok(0, T("decryptModelWeights"), str(cc_path), str(tmp_path / f"c{c}/private.key"), str(tmp_path / f"agg{c}_r{rnd}.json"),
               str(tmp_path / f"dec{c}_r{rnd}.json"))
            dec = json.load(open(tmp_path / f"dec{c}_r{rnd}.json"))["weights_summary"]
            vals = np.concatenate([np.asarray(d["values"]) for d in dec])
            assert np.max(np.abs(vals - want)) < 2.0 ** -20, (c, rnd)
        assert open(tmp_path / f"agg{c}_r1.json", "rb").read() == open(tmp_path / f"agg{c}_r2.json", "rb").read()  # deterministic
    assert os.path.getsize(tmp_path / "agg3_r1.json") == 28368755
    # tool-by-tool flow for clients 1, 2 and the served result decrypt to the same numbers, bit for bit
    ok(0, T("changeCipherDomain"), str(cc_path), str(tmp_path / "rk_1_2.key"), str(tmp_path / "enc_c1.json"), str(tmp_path / "t_c1_to_c2.json"))
    ok(0, T("aggregateEncryptedWeights"), str(cc_path), str(tmp_path / "enc_c2.json"), str(tmp_path / "t_c1_to_c2.json"), str(tmp_path / "t_agg.json"))
    ok(0, T("changeCipherDomain"), str(cc_path), str(tmp_path / "rk_2_1.key"), str(tmp_path / "t_agg.json"), str(tmp_path / "t_agg_to_c1.json"))
    for who, a, b in ((2, "sv_agg.json", "t_agg.json"), (1, "sv_agg_to_c1.json", "t_agg_to_c1.json")):
        ok(0, T("decryptModelWeights"), str(cc_path), str(tmp_path / f"c{who}/private.key"), str(tmp_path / a), str(tmp_path / "da.json"))
        ok(0, T("decryptModelWeights"), str(cc_path), str(tmp_path / f"c{who}/private.key"), str(tmp_path / b), str(tmp_path / "db.json"))
        assert open(tmp_path / "da.json", "rb").read() == open(tmp_path / "db.json", "rb").read()
    print("ppq_served rounds:", rounds)

    # ---- the same process as the reference's HTTP drop-box (runMserver.cpp:237-285) that calls the GPU path itself:
    # keys and encrypted weights arrive as multipart uploads, the round runs when the last client's file lands, the
    # clients download the results from the reference's URLs
    import http.client
    import socket
    store = tmp_path / "server" / "storage"
    (store / "client_1").mkdir(parents=True)
    (store / "client_2").mkdir(parents=True)
    (store / "CC.json").write_text(ccjson_text)
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    scfg = {"mSConfig": {"SERVER_IP": "127.0.0.1", "SERVER_PORT": port}, "CC": {"path": str(store / "CC.json")},
            "CLIENTS": {"CLIENT_1_PUBLIC": str(store / "client_1/client_1-public.key"),
                        "CLIENT_2_PUBLIC": str(store / "client_2/client_2-public.key"),
                        "CLIENT_1_REKEY": str(store / "client_1/client_1-ReKey.key"),
                        "CLIENT_2_REKEY": str(store / "client_2/client_2-ReKey.key"),
                        "CLIENT_1_ENCRYPTED_WEIGHTS_PATH": str(store / "client_1/encrypted_weights_c1.json"),
                        "CLIENT_2_ENCRYPTED_WEIGHTS_PATH": str(store / "client_2/encrypted_weights_c2.json"),
                        "OUTPUT_DOMAIN_CHANGED_PATH": str(store / "client_2/c1_domainChange_c2.json"),
                        "AGGREGATED_ENCRYPTED_WEIGHTS_PATH": str(store / "client_2/aggregated_weights.json"),
                        "OUTPUT_AGGREGATED_DOMAIN_CHANGED_PATH": str(store / "client_1/c2_domainChange_c1.json")}}
    (tmp_path / "sConfig.json").write_text(json.dumps(scfg))
    hp = subprocess.Popen([T("ppq_served"), "--http", str(tmp_path / "sConfig.json")], stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                          text=True, env=dict(os.environ, PPQ_QUIET="1"))
    try:
        while True:
            line = hp.stdout.readline()
            assert line, "ppq_served --http died"
            if "HTTP server running" in line:
                assert "(2 clients)" in line
                break

        def req(method, url, body=None, ctype=None):
            c = http.client.HTTPConnection("127.0.0.1", port, timeout=120)
            c.request(method, url, body=body, headers={"Content-Type": ctype} if ctype else {})
            r = c.getresponse()
            data = r.read()
            c.close()
            return r.status, data

        def upload(url, path):
            bnd = "----ppqtestboundary7MA4YWxk"
            body = (f"--{bnd}\r\nContent-Disposition: form-data; name=\"client_id\"\r\n\r\nclient\r\n"
                    f"--{bnd}\r\nContent-Disposition: form-data; name=\"file\"; filename=\"f\"\r\n"
                    f"Content-Type: application/octet-stream\r\n\r\n").encode() + open(path, "rb").read() + f"\r\n--{bnd}--\r\n".encode()
            st, data = req("POST", url, body, f"multipart/form-data; boundary={bnd}")
            assert st == 200 and data == b'{"status":"received"}', (url, st, data)

        st, data = req("GET", "/getCC")
        assert st == 200 and data == ccjson_text.encode()
        assert req("GET", "/nothing")[0] == 404 and req("GET", "/download/../../etc/passwd")[0] == 404
        upload("/uploadPubKeyC1", tmp_path / "c1/public.key")
        assert req("GET", "/sendPbKeyC1")[1] == open(tmp_path / "c1/public.key", "rb").read()
        upload("/uploadReKeyC1", tmp_path / "rk_1_2.key")    # 27 MB of JSON each: parsed once, resident from here on
        upload("/uploadReKeyC2", tmp_path / "rk_2_1.key")
        for rnd in (1, 2):
            upload("/uploadEncWeightsC1", tmp_path / "enc_c1.json")
            assert req("GET", "/download/client_2/aggregated_weights.json")[0] == (404 if rnd == 1 else 200)  # nothing runs early
            upload("/uploadEncWeightsC2", tmp_path / "enc_c2.json")                                           # -> the round
            st, agg2 = req("GET", "/download/client_2/aggregated_weights.json")
            st1, agg1 = req("GET", "/download/client_1/c2_domainChange_c1.json")
            assert st == 200 and st1 == 200 and len(agg2) == 28368755 and len(agg1) == 28368755
            # identical to what the stdin server / the three tools produce for the same inputs
            assert agg2 == open(tmp_path / "sv_agg.json", "rb").read()
            assert agg1 == open(tmp_path / "sv_agg_to_c1.json", "rb").read()
        assert req("POST", "/shutdown")[0] == 200
        out = hp.stdout.read()
        assert "round 2: OK clients=2 cts=27" in out, out[-2000:]
        assert out.count("LOAD_REKEY") == 0 and out.count("OK rekey") == 2  # two keys, parsed once each
    finally:
        hp.kill()


@pytest.mark.gpu
def test_cpp_host_aggregates_over_all_gpus_through_the_c_abi(tools):
    """host/multi_gpu_selftest: a plain C++ process (no Python, no torch), one thread per GPU, clients sharded round-robin,
    ppq_reencrypt_aggregate_host_ranks on every rank (local ppq_agg_* + ONE ncclReduce + mod q on the root) -- the root's
    aggregate must equal the single-GPU aggregate of all clients bit for bit.  Uses every GPU the box has (one rank on a
    single-GPU box: the communicator and the reduce entry point still run)."""
    r = run(os.path.join(tools, "multi_gpu_selftest"))
    assert r.returncode == 0, r.stdout + r.stderr
    assert "IDENTICAL to the single-GPU aggregate" in r.stdout
    import torch
    assert f"{torch.cuda.device_count()} GPUs" in r.stdout
