#!/usr/bin/env python
"""bench.py -- ReEncrypt+EvalAdd aggregation throughput (BASELINE.json metric) on N B200s of one node.

One "step" = one full aggregation round of the workload: every client's ciphertexts are proxy re-encrypted with
that client's re-encryption key (cc->ReEncrypt, server/src/changeCipherDomain.cpp:105) and summed into the
per-slot running aggregate (cc->EvalAdd, server/src/aggregateEncryptedWeights.cpp:106); with N > 1 the clients are
sharded across ranks and the per-rank partial sums are combined by ONE NCCL u64 sum-reduce + a final mod-q.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config c3] [--impl reference]

Prints ONE JSON line (rank 0).  `value` = ciphertexts/s with inputs resident in HBM; `e2e` = the same metric
through the host-buffer C-ABI call (H2D of every ciphertext and D2H of the aggregate inside the timed region).
`--impl reference` times the reference's CPU algorithm (the oracle restatement of OpenFHE: the real OpenFHE is an
un-vendored submodule absent from the reference tree) on the host cores with the same JSON shape.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# benchmark configurations.  slots = ciphertexts per client.
CONFIGS = {
    # name: logN, depth, first, scale, extra, dnum, clients, slots
    "c1": dict(logN=14, depth=2, first=60, scale=40, extra=20, dnum=2, clients=2, slots=27,
               desc="C1: 2-client GRU, CKKS N=2^14, 4 limbs (CC.json parameters)"),
    "c2": dict(logN=15, depth=2, first=60, scale=40, extra=20, dnum=2, clients=8, slots=84,
               desc="C2: 8-client LSTM (~1M params), CKKS N=2^15, 4 limbs"),
    "c3": dict(logN=16, depth=18, first=60, scale=59, extra=20, dnum=3, clients=64, slots=54,
               desc="C3: 64 clients x 54 cts, CKKS N=2^16, 20 limbs, dnum 3"),
    "c4": dict(logN=17, depth=38, first=60, scale=59, extra=20, dnum=3, clients=16, slots=16,
               desc="C4: 16 clients x 16 cts, CKKS N=2^17, 40 limbs, dnum 3"),
}
METRIC = "ciphertexts/s ReEncrypt+EvalAdd at N=2^16"
# Integer-issue floor of the path in us per ciphertext on one B200 ("Integer co-bound": butterflies and
# multiply-adds at 100 % of the fmaheavy pipe with the measured issue intervals), for the deferred-ModDown algorithm.
FLOOR_US = {"c3": 76.0}
UNIT = "ciphertexts/s"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, burst copy)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def bind_near_gpu(local):
    """Run this process on the CPUs NVML reports as local to the GPU (its NUMA node) while the pinned e2e buffers
    are allocated (first touch) and copied from: host->device copies from the far socket are 15-20 % slower.
    Returns (previous affinity, number of CPUs bound) or (None, 0)."""
    try:
        import pynvml
        import torch
        pynvml.nvmlInit()
        pr = torch.cuda.get_device_properties(local)
        bus = "%08x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        h = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = {64 * i + b for i, w in enumerate(words) for b in range(64) if (int(w) >> b) & 1}
        old = os.sched_getaffinity(0)
        cpus &= old
        if cpus and cpus != old:
            os.sched_setaffinity(0, cpus)
            return old, len(cpus)
        return None, len(cpus)
    except Exception:
        return None, 0


def stage_alg_words(ell, K, parts, bk, split=True, agg=False, group=8, clients=1):
    """Algorithmic (compulsory) u64 words per ciphertext, in units of N, for each of the 11 pipeline stages
    (DESIGN.md 'kernels').  bk = ciphertexts sharing one rekey inside a call (key traffic is amortised).
    agg: the deferred-ModDown pipeline (ppq_agg_*): stage 8 runs once per `group` clients, stages 9-10 once per
    slot for all `clients` of a GPU."""
    M = ell + K
    conv = parts * M - ell
    if agg:
        return [
            2 * ell,                                   # k_inv_row(c1):        R ell + W ell
            2 * ell,                                   # k_inv_col(c1):        R ell + W ell
            ell + conv,                                # k_bconv(modup):       R ell (COEFF digits) + W converted limbs
            2 * conv,                                  # k_fwd_col(modup):     R + W converted limbs
            2 * conv,                                  # k_fwd_row(modup):     R + W converted limbs
            conv + ell + 2 * parts * M / bk + ell + 4 * ell + 2 * K,  # k_ip_tma: R digits, own c1, keys/bk, c0; RMW accE; W 2K
            4 * K,                                     # k_inv_row(P):         R 2K + W 2K
            4 * K,                                     # k_inv_col(P):         R 2K + W 2K
            2 * K + 4.0 * ell / group,                 # k_bconv_wide:         R 2K per client; RMW accW per group
            4.0 * ell / clients,                       # k_fwd_col(moddown):   once per slot
            6.0 * ell / clients,                       # k_fwd_row_moddown:    R accW 2l + R accE 2l + W 2l, once per slot
        ]
    return [
        2 * ell,                                   # k_inv_row(c1):        R ell + W ell
        2 * ell,                                   # k_inv_col(c1):        R ell + W ell
        ell + conv,                                # k_bconv(modup):       R ell (COEFF digits) + W converted limbs
        2 * conv,                                  # k_fwd_col(modup):     R + W converted limbs
        (2 * conv) if split else 0,                # k_fwd_row(modup):     R + W converted limbs (split form only)
        conv + ell + 2 * parts * M / bk + 2 * M,   # k_ip_tma: R digits + R own c1 + R keys/bk + W 2(ell+K)
        (4 * K) if split else 0,                   # k_inv_row(P):         R 2K + W 2K (split form only)
        4 * K,                                     # k_inv_col(P):         R 2K + W 2K
        2 * K + 2 * ell,                           # k_bconv(moddown):     R 2K + W 2 ell
        4 * ell,                                   # k_fwd_col(moddown):   R 2 ell + W 2 ell
        9 * ell,                                   # k_fwd_row_moddown:    R conv 2l + R ext 2l + R c0 l + R acc 2l + W 2l
    ]


def path_alg_bytes(N, ell, K, dnum, bk):
    """W_alg = 8 N (13 ell + 4 K + 2 dnum (ell+K) / B_k) bytes per ciphertext (ideal fusion)."""
    return 8.0 * N * (13 * ell + 4 * K + 2.0 * dnum * (ell + K) / bk)


class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.p = None
        self.nv = []       # NVML samples taken by a thread of this process: they start at once, so that a timed
        self.nv_stop = None  # region shorter than nvidia-smi's start-up (8 GPUs: ~0.1 s) still gets its samples
        self.nv_thread = None

    def _nvml_loop(self, h, pynvml):
        bits = {"sw_power_cap": 0x4, "hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40}
        while not self.nv_stop.is_set():
            try:
                sm = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
                mx = pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)
                pw = pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0
                try:
                    r = pynvml.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    r = pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                self.nv.append((float(sm), float(mx), pw, {n for n, b in bits.items() if int(r) & b}))
            except Exception:
                return
            self.nv_stop.wait(0.01)

    def prepare(self):
        """NVML initialisation and the device handle, done BEFORE the warm-up so that polling starts with the timed
        region (at 8 GPUs a 3-step region is 0.13 s; initialising inside it left one sample)."""
        try:
            import pynvml
            import torch
            pynvml.nvmlInit()
            pr = torch.cuda.get_device_properties(self.index)
            bus = "%08x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
            self._h = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
            self._nv = pynvml
        except Exception:
            self._h = None

    def start(self):
        import threading
        if getattr(self, "_h", None) is None and not hasattr(self, "_nv"):
            self.prepare()
        if getattr(self, "_h", None) is not None:
            self.nv_stop = threading.Event()
            self.nv_thread = threading.Thread(target=self._nvml_loop, args=(self._h, self._nv), daemon=True)
            self.nv_thread.start()
        else:
            self.nv_thread = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                       "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.p = None

    def stop(self):
        if self.nv_thread is not None:
            self.nv_stop.set()
            self.nv_thread.join(timeout=2)
        if self.p is None and not self.nv:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        out = ""
        if self.p is not None:
            time.sleep(0.05)
            self.p.terminate()
            try:
                out, _ = self.p.communicate(timeout=5)
            except Exception:
                self.p.kill()
                out = ""
        sm, mx, pw, reasons = [], [], [], set()
        for a, b, c, r in self.nv:
            sm.append(a)
            mx.append(b)
            pw.append(c)
            reasons |= r
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
                pw.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "power_w_max": float(max(pw)),
                "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------ CPU arm
def cpu_run(cfg, steps, warmup, sample_slots=None):
    """The reference's CPU algorithm (oracle restatement of the OpenFHE path) on all host threads, on a bounded
    sample of the workload: ONE client's rekey and `sample_slots` of its ciphertexts per step."""
    import oracle
    o = oracle.Oracle.synthetic(cfg["logN"], cfg["depth"], cfg["first"], cfg["scale"], cfg["extra"], cfg["dnum"])
    T = oracle.max_threads()
    ell = o.L
    ct_bytes = 2 * ell * o.N * 8
    if sample_slots is None:
        sample_slots = max(2, min(4 * T, int((4 << 30) // ct_bytes), 256))
    rng = np.random.default_rng(42)
    q = o.moduli.astype(np.uint64)
    cts = (rng.integers(0, 1 << 62, (1, sample_slots, 2, ell, o.N), dtype=np.uint64) % q[:ell].reshape(1, 1, 1, ell, 1))
    kb = rng.integers(0, 1 << 62, (o.dnum, o.M, o.N), dtype=np.uint64) % q.reshape(1, o.M, 1)
    ka = rng.integers(0, 1 << 62, (o.dnum, o.M, o.N), dtype=np.uint64) % q.reshape(1, o.M, 1)
    times = []
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        o.reencrypt_accumulate_batch(ell, cts, [kb], [ka], nthreads=T)
        dt = time.perf_counter() - t0
        if it >= warmup:
            times.append(dt)
    t = float(np.median(times))
    return {"value": sample_slots / t, "unit": UNIT, "cores": T, "kind": "port",
            "note": "scalar C restatement of the OpenFHE algorithms: Shoup butterflies, 128-bit accumulators reduced by Barrett "
                    "(no 128-bit division in the inner loops since round 2: 1.7-2x the round-1 port), branch-free modular "
                    "add/sub, OpenMP over ciphertext slots; no SIMD -- an AVX-512 NTT such as OpenFHE+HEXL would be faster "
                    "still, so GPU/CPU ratios over this arm remain upper bounds",
            "sample": f"{sample_slots} ciphertexts of 1 client (ell={ell}, N=2^{o.logN}, K={o.K}, dnum={o.dnum}) per step, "
                      f"{steps} steps after {warmup} warm-up, median; OpenFHE-algorithm CPU restatement (oracle/), "
                      f"OpenMP over ciphertexts",
            "ms_per_step": t * 1e3, "_ell": ell, "_K": o.K}


def reference_arm(args, cfg):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps, warmup = max(1, min(args.steps, 3)), max(1, min(args.warmup, 1))
    r = cpu_run(cfg, steps, warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {"workload": cfg["desc"], "impl_note": "CPU, host cores only; each step is a bounded sample"},
        "cpu_baseline": {k: v for k, v in r.items() if not k.startswith("_") and k != "ms_per_step"},
        "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ GPU arm
def _rand_res(torch, dev, shape, q):  # uniform-ish canonical residues
    return torch.randint(0, 1 << 62, shape, dtype=torch.int64, device=dev).remainder_(q)


def int_pipe_of(kernel_name):
    """fmaheavy pipe utilisation (%) of a kernel from the committed ncu --set full capture (profiles/ncu_int_pipe.json)."""
    p = os.path.join(ROOT, "profiles", "ncu_int_pipe.json")
    try:
        return json.load(open(p)).get(kernel_name)
    except Exception:
        return None


def run_aggregation(torch, cc, cfg, clients, steps, warmup, comm=None, world=1, rank=0, profile=True, barrier=None,
                    allmax=None, sampler=None, rounds=1):
    """Device-resident ReEncrypt+EvalAdd of `clients` (this rank's) x cfg['slots'] ciphertexts per step through
    ppq_agg_* (+ ppq_reduce_across_ranks).  Returns timing, per-stage CUDA-event times and the data (for e2e)."""
    from ppqsflhe_b200 import Aggregator, CiphertextBatch, CtMeta, EvalKey
    from ppqsflhe_b200._capi import check, lib
    dev = cc.device
    ell, N, M = cc.L, cc.N, cc.M
    n_slots = cfg["slots"]
    qv = torch.from_numpy(cc.moduli.astype(np.int64)).to(dev)
    # rounds > 1: that many independent aggregations (own ciphertexts and keys) per step, so that a configuration whose
    # inputs would fit the 126 MB L2 still streams every input from HBM
    cts = [CiphertextBatch(_rand_res(torch, dev, (n_slots, 2, ell, N), qv[:ell].view(1, 1, ell, 1)), CtMeta())
           for _ in range(rounds) for _ in clients]
    keys = [EvalKey(cc, _rand_res(torch, dev, (cc.dnum, M, N), qv.view(1, M, 1)), _rand_res(torch, dev, (cc.dnum, M, N), qv.view(1, M, 1)))
            for _ in range(rounds) for _ in clients]
    nc = len(clients)
    acc = CiphertextBatch(torch.zeros((n_slots, 2, ell, N), dtype=torch.int64, device=dev), CtMeta())
    agg = Aggregator(cc, ell, n_slots)
    torch.cuda.synchronize()

    def step():
        for r in range(rounds):
            agg.add_many(cts[r * nc:(r + 1) * nc], keys[r * nc:(r + 1) * nc])
            if nc:
                agg.finish(out=acc.data)
            else:
                acc.data.zero_()
        if comm is not None:
            comm.reduce(acc, root=0)  # one ncclReduce of canonical residues (< 2^60) + mod q on the root

    def timed(fn, k):
        if barrier:
            barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(k):
            fn()
        e1.record()
        torch.cuda.synchronize()
        if barrier:
            barrier()
        ms = e0.elapsed_time(e1)
        return allmax(ms) if allmax else ms

    if sampler is not None:
        sampler.prepare()
    for _ in range(warmup):
        step()
    torch.cuda.synchronize()
    if profile:
        check(lib().ppq_profile_enable(cc.h, 1))
    l0 = int(lib().ppq_launch_count())
    if sampler is not None:
        sampler.start()
    total_ms = timed(step, steps)
    clocks = sampler.stop() if sampler is not None else None
    launches = int(lib().ppq_launch_count()) - l0
    NST = 11
    pms, pl = (C.c_double * NST)(), (C.c_uint64 * NST)()
    if profile:
        check(lib().ppq_profile_read(cc.h, pms, pl, 1))
        check(lib().ppq_profile_enable(cc.h, 0))
    return {"total_ms": total_ms, "stage_ms": list(pms), "stage_launches": list(pl), "launches": launches, "clocks": clocks, "cts": cts,
            "keys": keys, "acc": acc, "agg": agg, "step": step, "timed": timed}


def side_config(torch, name, steps=2, warmup=3):
    """Short device-resident run of another BASELINE.json configuration on this GPU (parity-tested shapes; these are
    context for the headline, not the headline)."""
    from ppqsflhe_b200 import CryptoContext
    cfg = dict(CONFIGS[name])
    cc = CryptoContext.synthetic(cfg["logN"], cfg["depth"], cfg["first"], cfg["scale"], cfg["extra"], cfg["dnum"], torch.cuda.current_device())
    r = None
    try:
        per_round = cfg["clients"] * cfg["slots"] * 2 * cc.L * cc.N * 8
        rounds = max(1, -(-(512 << 20) // per_round))  # inputs per step >= 512 MiB
        r = run_aggregation(torch, cc, cfg, list(range(cfg["clients"])), steps, warmup, profile=False, rounds=rounds)
        n = rounds * cfg["clients"] * cfg["slots"]
        ms = r["total_ms"] / steps
        v = n / (ms * 1e-3)
        peak, _ = peaks()
        w = path_alg_bytes(cc.N, cc.L, cc.K, cc.dnum, cfg["slots"])
        return {"workload": cfg["desc"], "value": v, "unit": UNIT, "ms_per_step": ms, "steps": steps, "warmup": warmup,
                "alg_bytes_per_ct": w, "path_frac": v * w / 1e9 / peak, "limbs": cc.L, "aux_limbs": cc.K,
                "rounds_per_step": rounds,
                "l2": "%d independent aggregations per step: %.0f MiB of inputs per step > 126 MB L2, no flush" %
                      (rounds, n * 2 * cc.L * cc.N * 8 / 2 ** 20)}
    finally:
        del r
        cc.close()
        torch.cuda.empty_cache()


def ntt_sweep_summary(torch, limbs=16):
    """C5 summary (BASELINE.json configs[4]): batched forward/inverse NTT through ppq_ntt_forward / ppq_ntt_inverse,
    logN 12..17 at `limbs` limbs, working set >= 256 MiB, GB/s at the algorithmic 16 N bytes per limb transform."""
    from ppqsflhe_b200 import CryptoContext
    peak, _ = peaks()
    dev = torch.device("cuda", torch.cuda.current_device())
    rows = []
    for logN in (12, 13, 14, 15, 16, 17):
        N = 1 << logN
        cc = CryptoContext.synthetic(logN, limbs - 2, 60, 59, 20, 2, dev.index)
        nl = cc.L
        reps = max(1, (256 << 20) // (nl * N * 8))
        ids = np.tile(np.arange(nl, dtype=np.int32), reps)
        qv = torch.from_numpy(cc.moduli[:nl].astype(np.int64)).to(dev).repeat(reps).view(-1, 1)
        buf = _rand_res(torch, dev, (len(ids), N), qv)
        row = {"logN": logN, "limbs": nl, "polys": int(len(ids)), "MiB": int(buf.numel() * 8 >> 20)}
        for name, inv in (("fwd", False), ("inv", True)):
            for _ in range(3):
                cc.ntt(buf, ids, inverse=inv)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5):
                cc.ntt(buf, ids, inverse=inv)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 5
            row[name + "_gbs"] = 16.0 * N * len(ids) / (ms * 1e-3) / 1e9
            row[name + "_frac"] = row[name + "_gbs"] / peak
        rows.append(row)
        del buf, qv
        cc.close()
        torch.cuda.empty_cache()
    return rows


def client_leg(torch, name, n_ct):
    """Client half of north_star: batched encode+Encrypt (client/src/encryptModelWeights.cpp:109-110) and
    Decrypt+decode (client/src/decryptModelWeights.cpp:108-109) of n_ct full ciphertexts, device-resident.
    Compulsory bytes per ciphertext: N/2 doubles on one side, the 2*ell*N residues of the ciphertext on the other."""
    from ppqsflhe_b200 import CryptoContext
    cfg = CONFIGS[name]
    peak, _ = peaks()
    dev_i = torch.cuda.current_device()
    cc = CryptoContext.synthetic(cfg["logN"], cfg["depth"], cfg["first"], cfg["scale"], cfg["extra"], cfg["dnum"], dev_i)
    try:
        sk, pk = cc.KeyGen()
        vals = torch.empty((n_ct, cc.N // 2), dtype=torch.float64, device=cc.device).normal_(0, 0.05).clamp_(-0.3, 0.3)
        out = {"workload": cfg["desc"], "cts": n_ct, "limbs": cc.L, "logN": cc.logN}
        alg = 16.0 * cc.L * cc.N + 4.0 * cc.N

        def t(fn, reps=5):
            r = fn()
            r = fn()  # second warm-up while the first result is alive: the allocator has both blocks before the timing
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                r = fn()
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1) / reps, r

        ms2, _ = t(lambda: cc.Encrypt(pk, cc.MakeCKKSPackedPlaintext(vals)))
        ms, ct = t(lambda: cc.EncodeEncrypt(pk, vals))
        out["encode_encrypt"] = {"cts_per_s": n_ct / (ms * 1e-3), "ms": ms, "alg_bytes_per_ct": alg,
                                 "hbm_frac": n_ct * alg / (ms * 1e-3) / 1e9 / peak,
                                 "api": "ppq_sample_encrypt_noise + ppq_encode_encrypt (3*ell limb transforms per ct)",
                                 "two_calls_cts_per_s": n_ct / (ms2 * 1e-3),
                                 "two_calls_api": "ppq_encode + ppq_sample_encrypt_noise + ppq_encrypt (4*ell transforms)"}
        ms, pt = t(lambda: cc.Decrypt(sk, ct))
        out["decrypt_decode"] = {"cts_per_s": n_ct / (ms * 1e-3), "ms": ms, "alg_bytes_per_ct": alg,
                                 "hbm_frac": n_ct * alg / (ms * 1e-3) / 1e9 / peak}
        err = float((pt.values - vals).abs().max().item())
        out["max_abs_err"] = err
        return out
    finally:
        cc.close()
        torch.cuda.empty_cache()


def round_tail_leg(torch, name="c3", n_keys=8):
    """What a round does with the aggregate: cc->EvalMult(sum, 1/C) with its rescale
    (server/src/aggregateEncryptedWeights.cpp:83,92,107) and the fan-out of the result to the other key domains
    (orchestration/server_fns.sh:76-80: ppq_reencrypt_fanout, ModUp shared over the keys), device-resident."""
    from ppqsflhe_b200 import CiphertextBatch, CryptoContext, CtMeta, EvalKey
    cfg = CONFIGS[name]
    dev_i = torch.cuda.current_device()
    cc = CryptoContext.synthetic(cfg["logN"], cfg["depth"], cfg["first"], cfg["scale"], cfg["extra"], cfg["dnum"], dev_i)
    try:
        dev, ell, M, N, n = cc.device, cc.L, cc.M, cc.N, cfg["slots"]
        qv = torch.from_numpy(cc.moduli.astype(np.int64)).to(dev)
        agg = CiphertextBatch(_rand_res(torch, dev, (n, 2, ell, N), qv[:ell].view(1, 1, ell, 1)),
                              CtMeta(noise_scale_deg=2, level=0, scaling_factor=cc.sf_big(0)))
        keys = [EvalKey(cc, _rand_res(torch, dev, (cc.dnum, M, N), qv.view(1, M, 1)),
                        _rand_res(torch, dev, (cc.dnum, M, N), qv.view(1, M, 1))) for _ in range(n_keys)]

        def t(fn, reps=3):
            r = fn()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                r = fn()
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1) / reps, r

        ms_mul, scaled = t(lambda: cc.EvalMult(agg, 1.0 / cfg["clients"]))
        ms_fan, outs = t(lambda: cc.ReEncryptFanOut(scaled, keys))
        return {"workload": cfg["desc"], "cts": n, "limbs_in": ell, "limbs_out": scaled.ell, "keys": n_keys,
                "eval_mult_rescale": {"us_per_ct": ms_mul * 1e3 / n, "cts_per_s": n / (ms_mul * 1e-3),
                                      "api": "ppq_eval_mult_const (rescale + constant)"},
                "fanout": {"us_per_ct_key": ms_fan * 1e3 / (n * n_keys), "ct_keys_per_s": n * n_keys / (ms_fan * 1e-3),
                           "api": "ppq_reencrypt_fanout"}}
    finally:
        cc.close()
        torch.cuda.empty_cache()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="c3", choices=sorted(CONFIGS))
    ap.add_argument("--clients", type=int, default=0)
    ap.add_argument("--slots", type=int, default=0)
    ap.add_argument("--wave", type=int, default=0)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the configs / C5 / client blocks (N = 1 only)")
    args = ap.parse_args()
    cfg = dict(CONFIGS[args.config])
    if args.clients:
        cfg["clients"] = args.clients
    if args.slots:
        cfg["slots"] = args.slots
    if args.impl == "reference":
        return reference_arm(args, cfg)
    if args.warmup < 3:
        args.warmup = 3

    import torch
    import torch.distributed as dist
    from ppqsflhe_b200 import Communicator, CryptoContext, EvalKey
    from ppqsflhe_b200._capi import check, lib, vp
    from ppqsflhe_b200.aggregate import clients_of_rank, plan_shards

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    assert world == args.gpus or world == 1, "launch with torchrun --nproc-per-node N for --gpus N"

    cc = CryptoContext.synthetic(cfg["logN"], cfg["depth"], cfg["first"], cfg["scale"], cfg["extra"], cfg["dnum"], local)
    if args.wave:
        cc.set_wave(args.wave)
    ell, K, N, M = cc.L, cc.K, cc.N, cc.M
    parts = cc.num_parts(ell)
    n_slots = cfg["slots"]
    my_clients = clients_of_rank(cfg["clients"], world, rank)
    torch.manual_seed(42 + rank)

    comm = None
    barrier = allmax = None
    if world > 1:
        # the library's own communicator (ppq_comm_*): rank 0 makes the NCCL id, torch.distributed only carries it
        idt = torch.zeros(128, dtype=torch.uint8, device=dev)
        if rank == 0:
            idt.copy_(torch.frombuffer(bytearray(Communicator.unique_id()), dtype=torch.uint8))
        dist.broadcast(idt, 0)
        comm = Communicator(cc, world, rank, bytes(idt.cpu().numpy().tobytes()))
        barrier = dist.barrier

        def allmax(ms):
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())

    R = run_aggregation(torch, cc, cfg, my_clients, args.steps, args.warmup, comm, world, rank, True, barrier, allmax,
                        ClockSampler(local))
    clocks = R["clocks"]
    total_ms, stage_ms, stage_launches = R["total_ms"], R["stage_ms"], R["stage_launches"]
    cts, keys, timed = R["cts"], R["keys"], R["timed"]
    n_ct_step = cfg["clients"] * n_slots
    ms_per_step = total_ms / args.steps
    value = n_ct_step / (ms_per_step * 1e-3)

    # ---- roofline.  `frac` is SURVEY 8(d)'s figure: whole-path W_alg (ideal fusion) x ct/s per GPU against the HBM
    # peak.  The dominant kernel (largest CUDA-event share inside the timed region) is reported beside it against the
    # bytes of ITS pass (`pass_frac`), with its integer-pipe utilisation from the committed ncu capture.
    peak, peak_src = peaks()
    names = [lib().ppq_profile_stage_name(i).decode() for i in range(11)]
    dom = int(np.argmax(stage_ms))
    words = stage_alg_words(ell, K, parts, n_slots, agg=True, clients=max(1, len(my_clients)))
    local_ct = len(my_clients) * n_slots * args.steps
    dom_bytes = words[dom] * 8.0 * N * local_ct
    dom_gbs = dom_bytes / (stage_ms[dom] * 1e-3) / 1e9 if stage_ms[dom] > 0 else 0.0
    traffic = None
    traffic_path = None
    tpath = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    cts_per_launch = local_ct / max(1, stage_launches[dom])
    if os.path.exists(tpath):
        try:
            tj = json.load(open(tpath)).get(args.config, {})
            per_ct = tj.get("dram_bytes_per_ct", {}).get(names[dom])
            traffic = per_ct * cts_per_launch if per_ct is not None else None
            traffic_path = tj.get("path_dram_bytes_per_ct")
        except Exception:
            traffic = None
    w_alg = path_alg_bytes(N, ell, K, cc.dnum, n_slots)
    per_gpu = value / world
    us_per_ct = 1e6 / per_gpu if per_gpu > 0 else None
    roofline = {
        "bound": "hbm", "achieved": per_gpu * w_alg / 1e9, "peak": peak, "unit": "GB/s",
        "frac": per_gpu * w_alg / 1e9 / peak, "peak_source": peak_src,
        "alg_bytes_per_ct": w_alg, "us_per_ct": us_per_ct,
        "note": "whole ReEncrypt+EvalAdd path per GPU: ct/s x SURVEY 8(d) W_alg = 8N(13 ell + 4K + 2 dnum (ell+K)/B_k) "
                "(ideal fusion) against the measured HBM peak",
        "traffic": traffic_path, "traffic_note": "dram__bytes read+write per ciphertext summed over the path's kernels "
                                                 "(ncu --set full, profiles/ncu_traffic.json)",
        "kernel": names[dom], "pass_achieved": dom_gbs, "pass_frac": dom_gbs / peak,
        "pass_alg_bytes_per_ct": words[dom] * 8.0 * N, "pass_traffic_per_launch": traffic,
        "launch_avg_ms": stage_ms[dom] / max(1, stage_launches[dom]), "cts_per_launch": cts_per_launch,
        "int_pipe": int_pipe_of(names[dom]),
        "int_pipe_note": "sm__inst_executed_pipe_fmaheavy % of the dominant kernel (committed ncu capture); the path is "
                         "integer-issue bound (DESIGN.md 4), so this, not the HBM fraction, is the binding roofline",
        "floor_us_per_ct": FLOOR_US.get(args.config),
        "stage_share": {n: (m / max(1e-9, sum(stage_ms))) for n, m in zip(names, stage_ms)},
        "stage_us_per_ct": {n: (m * 1e3 / max(1, local_ct)) for n, m in zip(names, stage_ms)},
    }
    gpu_launches = R["launches"]

    # ---- standalone batched NTT / iNTT (second half of BASELINE.json's metric: "NTT HBM GB/s"), rank 0 only.
    ntt = None
    if rank == 0:
        reps = 20
        qv = torch.from_numpy(cc.moduli.astype(np.int64)).to(dev)
        ids = np.tile(np.arange(M, dtype=np.int32), reps)
        buf = _rand_res(torch, dev, (len(ids), N), qv.repeat(reps).view(-1, 1))
        res = {}
        for name, inv in (("fwd", False), ("inv", True)):
            cc.ntt(buf, ids, inverse=inv)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5):
                cc.ntt(buf, ids, inverse=inv)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 5
            res[name] = {"ms": ms, "gbs": 16.0 * N * len(ids) / (ms * 1e-3) / 1e9, "limb_ntt_per_s": len(ids) / (ms * 1e-3)}
        ntt = {"logN": cfg["logN"], "polys": int(len(ids)), "bytes": int(buf.numel() * 8), "alg_bytes_per_transform": 16 * N,
               "fwd_gbs": res["fwd"]["gbs"], "inv_gbs": res["inv"]["gbs"], "fwd_frac": res["fwd"]["gbs"] / peak,
               "inv_frac": res["inv"]["gbs"] / peak, "fwd_limb_ntt_per_s": res["fwd"]["limb_ntt_per_s"],
               "inv_limb_ntt_per_s": res["inv"]["limb_ntt_per_s"], "api": "ppq_ntt_forward / ppq_ntt_inverse (C ABI)"}
        del buf
        torch.cuda.empty_cache()

    # ---- e2e: host buffers in, aggregate out, every copy inside the timed region, through ONE C-ABI call per step
    # (ppq_reencrypt_aggregate_host; N > 1: ppq_reencrypt_aggregate_host_ranks, which ends in the NCCL reduce).
    # The copy-only leg runs the same uploads/download on the same streams without kernels: the box's ceiling.
    e2e = None
    if not args.no_e2e:
        blk = n_slots * 2 * ell * N
        pblk = n_slots * int(lib().ppq_packed_words(cc.h, ell))
        old_aff, near_cpus = bind_near_gpu(local)
        pool_n = max(1, min(len(my_clients), 4))
        nk = len(keys)
        qv_all = torch.from_numpy(cc.moduli.astype(np.int64)).to(dev)

        def e2e_leg(packed):
            """one format of the host buffers: [n_slots][2][ell][N] u64, or packed residues (bitlen(q_i) bits each)"""
            words = pblk if packed else blk
            pool = [torch.empty(words, dtype=torch.int64).pin_memory() for _ in range(pool_n)]
            for i, p in enumerate(pool):
                if i < len(cts):
                    if packed:  # the deserialiser's output format; made here by the device packer
                        d_pk = torch.empty(words, dtype=torch.int64, device=dev)
                        check(lib().ppq_pack(cc.h, ell, vp(cts[i].data.data_ptr()), vp(d_pk.data_ptr()), n_slots, None))
                        torch.cuda.synchronize()
                        p.copy_(d_pk)
                        del d_pk
                    else:
                        p.copy_(cts[i].data.view(-1))
            h_acc = torch.empty(words, dtype=torch.int64).pin_memory()
            d2h = words * 8 if rank == 0 else 0
            sfx = "_packed" if packed else ""
            f_one = getattr(lib(), "ppq_reencrypt_aggregate_host" + sfx)
            f_ranks = getattr(lib(), "ppq_reencrypt_aggregate_host" + sfx + "_ranks")
            f_copy = getattr(lib(), "ppq_copy_ceiling_host" + sfx)
            # Shards for the host-buffer path.  The ranks of one box do not all copy at the same rate (8-GPU B200 box, all
            # ranks copying: 23 GB/s on GPUs 0-3, 35 GB/s on GPUs 4-7, tools/copy_rates.py), and the step ends with the
            # slowest rank: every rank times one copy-only pass of an equal shard while all ranks copy, the per-client
            # costs are gathered and ppq_plan_shards gives each rank the clients it can move in the same time.
            nk_e, shards, costs = nk, None, None
            if world > 1:
                ptrs0 = (vp * max(1, nk))(*[pool[i % pool_n].data_ptr() for i in range(nk)])
                mine_ms = []
                for _ in range(2):  # the second pass is the measurement
                    dist.barrier()
                    torch.cuda.synchronize()
                    t0 = time.perf_counter()
                    check(f_copy(cc.h, ell, ptrs0, nk, n_slots, vp(h_acc.data_ptr())))  # synchronous
                    mine_ms.append((time.perf_counter() - t0) * 1e3)
                ct = torch.tensor([mine_ms[-1] / max(1, nk) if nk else 1e9], device=dev, dtype=torch.float64)
                allc = [torch.zeros_like(ct) for _ in range(world)]
                dist.all_gather(allc, ct)
                costs = [float(c.item()) for c in allc]
                shards = plan_shards(cfg["clients"], costs)
                nk_e = shards[rank]
            while len(keys) < nk_e:  # a rank that takes more clients holds their re-encryption keys too
                keys.append(EvalKey(cc, _rand_res(torch, dev, (cc.dnum, M, N), qv_all.view(1, M, 1)),
                                    _rand_res(torch, dev, (cc.dnum, M, N), qv_all.view(1, M, 1))))
            rks_e = (vp * max(1, nk_e))(*[k.h for k in keys[:nk_e]])
            ptrs = (vp * max(1, nk_e))(*[pool[i % pool_n].data_ptr() for i in range(nk_e)])
            h2d = nk_e * words * 8
            if world == 1:
                def e2e_step():
                    check(f_one(cc.h, rks_e, ell, ptrs, nk_e, n_slots, vp(h_acc.data_ptr())))
            else:
                def e2e_step():
                    check(f_ranks(cc.h, comm.h, 0, rks_e, ell, ptrs, nk_e, n_slots, vp(h_acc.data_ptr())))

            def copy_step():
                check(f_copy(cc.h, ell, ptrs, nk_e, n_slots, vp(h_acc.data_ptr())))
                if world > 1:
                    # the real step ends in a collective, so no rank starts its next uploads while another still copies;
                    # without the same rendezvous here the early ranks take bandwidth from the late ones and the
                    # "ceiling" comes out BELOW the real leg (171 vs 198 GB/s on the 8-GPU box)
                    dist.barrier()

            for _ in range(args.warmup):
                e2e_step()
            copy_step()
            c_ms = timed(copy_step, args.steps)       # the copy-only leg runs before AND after the e2e leg; the faster
            e_ms = timed(e2e_step, args.steps)        # of the two is the ceiling (host-side contention varies run to run)
            c_ms = min(c_ms, timed(copy_step, args.steps))
            tot = torch.tensor([float(h2d), float(d2h)], device=dev, dtype=torch.float64)
            if world > 1:
                dist.all_reduce(tot)
            bytes_step = float(tot[0].item() + tot[1].item())
            e_val = n_ct_step / (e_ms / args.steps * 1e-3)
            c_val = n_ct_step / (c_ms / args.steps * 1e-3)
            name = "ppq_reencrypt_aggregate_host" + sfx
            return {"value": e_val, "unit": UNIT, "h2d_bytes_per_step": int(tot[0].item()),
                    "d2h_bytes_per_step": int(tot[1].item()), "ms_per_step": e_ms / args.steps,
                    "api": f"{name} (C ABI, pinned host buffers)" if world == 1 else
                           f"{name}_ranks (C ABI, pinned host buffers, ends in ppq_reduce_across_ranks)",
                    "host_format": ("packed residues: bitlen(q_i) bits per residue of limb i, unpacked on the device "
                                    "(include/ppq_b200.h)") if packed else "[n_slots][2][ell][N] u64",
                    "note": f"every client's {words * 8 >> 20} MiB upload is copied each step from a pool of {pool_n} "
                            "pinned buffers",
                    "host_cpus_near_gpu": near_cpus,
                    "shards": shards, "shard_cost_ms_per_client": costs,
                    "shard_note": None if shards is None else
                    "clients per rank from ppq_plan_shards over each rank's measured copy-only time per client (all ranks "
                    "copying at once); the device-resident `value` keeps equal shards",
                    "copy_gbs": bytes_step / (e_ms / args.steps * 1e-3) / 1e9,
                    "h2d_ceiling_gbs": bytes_step / (c_ms / args.steps * 1e-3) / 1e9,
                    "ceiling_value": c_val, "frac_of_copy_ceiling": e_val / c_val,
                    "ceiling_note": "ppq_copy_ceiling_host" + sfx + ": the same pinned pools, streams and copies with no "
                                    "kernel launched"}

        # headline: packed host buffers (the path is bound by the host->device copy, and a residue of limb i carries
        # bitlen(q_i) bits); the u64 form of the same call is reported beside it
        e2e_u64 = e2e_leg(False)
        e2e = e2e_leg(True)
        e2e["u64_host_buffers"] = e2e_u64
        if old_aff is not None:
            os.sched_setaffinity(0, old_aff)  # the CPU baseline below uses every host core again

    # ---- the other BASELINE.json configurations, C5 and the client half: N = 1 only, after the headline is measured
    del R, cts, keys
    torch.cuda.empty_cache()
    extras = {}
    if world == 1 and not args.no_extras:
        for name in ("c1", "c2", "c4"):
            if name == args.config:
                continue
            try:
                extras[name] = side_config(torch, name)
            except Exception as e:  # an extra must never take the headline down
                extras[name] = {"error": repr(e)}
        try:
            extras["c5_ntt_sweep"] = ntt_sweep_summary(torch)
        except Exception as e:
            extras["c5_ntt_sweep"] = {"error": repr(e)}
        client = {}
        for name, n_ct in (("c1", 216), ("c3", 54)):
            try:
                client[name] = client_leg(torch, name, n_ct)
            except Exception as e:
                client[name] = {"error": repr(e)}
        extras["client"] = client
        try:
            extras["round_tail"] = round_tail_leg(torch)
        except Exception as e:
            extras["round_tail"] = {"error": repr(e)}

    # ---- CPU baseline on rank 0, N = 1 only
    cpu = None
    if world == 1 and not args.no_cpu:
        r = cpu_run(cfg, 3, 1)  # >= 3 repetitions, median (about 1.3 s each at C3 on 16 threads)
        cpu = {k: v for k, v in r.items() if not k.startswith("_") and k != "ms_per_step"}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "u64", "data": "synthetic",
            "config": {"workload": cfg["desc"], "clients": cfg["clients"], "cts_per_client": n_slots, "logN": cfg["logN"],
                       "limbs": ell, "aux_limbs": K, "dnum": cc.dnum, "parallelism": f"clients sharded over {world} GPU(s)",
                       "l2": "inputs per step (%.1f GiB per GPU) >> 126 MB L2; no flush needed" %
                             (len(my_clients) * n_slots * 2 * ell * N * 8 / 2 ** 30),
                       "residues": "uniform random canonical residues (timing is data independent)",
                       "algorithm": "ppq_agg_*: per client ModUp + key inner product + inverse NTT of the P limbs; the "
                                    "linear tail of ModDown (P->Q conversion, forward NTT, subtraction) runs on sums over "
                                    "clients; output bit-identical to per-client ReEncrypt then EvalAdd"},
            "roofline": roofline, "ntt": ntt, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": gpu_launches, "clocks": clocks,
            "configs": {k: v for k, v in extras.items() if k in ("c1", "c2", "c3", "c4")},
            "c5_ntt_sweep": extras.get("c5_ntt_sweep"), "client": extras.get("client"),
            "round_tail": extras.get("round_tail"),
        }
        print(json.dumps(line), flush=True)
    if comm is not None:
        comm.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
