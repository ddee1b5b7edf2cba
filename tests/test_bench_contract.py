"""bench.py contract checks that need no GPU: the reference arm (CPU oracle on the host cores) prints ONE JSON line of
the agreed shape, and the roofline helpers agree with SURVEY.md 8(d)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_json_line():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--config", "c1", "--steps", "1",
                        "--warmup", "1"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "ciphertexts/s" and d["higher_is_better"] is True
    assert d["metric"].startswith("ciphertexts/s ReEncrypt+EvalAdd")
    assert d["value"] > 0 and d["e2e"]["value"] == d["value"]
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and "sample" in cb
    assert d["gpu_launches"] == 0 and "workload" in d["config"]


def test_reference_arm_other_ranks_print_nothing():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--config", "c1", "--gpus", "2"],
                       capture_output=True, text=True, timeout=120, env=env)
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_algorithmic_bytes_match_survey():
    sys.path.insert(0, ROOT)
    import bench
    # C3 (ell=20, K=7, dnum=3, B_k=54) = 291.0 words x N = 152.57 MB per ciphertext
    assert abs(bench.path_alg_bytes(65536, 20, 7, 3, 54) / 1e6 - 152.57) < 0.01
    # C1 (ell=4, K=2, dnum=2, B_k=27): 7.98 MB
    assert abs(bench.path_alg_bytes(16384, 4, 2, 2, 27) / 1e6 - 7.98) < 0.01
    w = bench.stage_alg_words(20, 7, 3, 54, split=True)
    assert len(w) == 11 and w[0] == 40 and w[3] == 2 * 61 and w[10] == 180
    assert bench.stage_alg_words(20, 7, 3, 54, split=False)[4] == 0
    # deferred-ModDown pipeline (ppq_agg_*): the inner product also reads c0 and read-modify-writes the EVAL sum; the
    # conversion runs per group of 8 clients, the last two stages once per slot
    a = bench.stage_alg_words(20, 7, 3, 54, agg=True, group=8, clients=64)
    assert len(a) == 11 and a[:5] == w[:5] and a[6:8] == w[6:8]
    assert abs(a[5] - (61 + 20 + 2 * 3 * 27 / 54 + 20 + 80 + 14)) < 1e-9
    assert a[8] == 14 + 10 and abs(a[9] - 80 / 64) < 1e-9 and abs(a[10] - 120 / 64) < 1e-9
