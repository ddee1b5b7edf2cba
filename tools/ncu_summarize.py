#!/usr/bin/env python
"""Summarise ncu outputs into small committed text files under profiles/.

    python tools/ncu_summarize.py launches gpurun_out/X_launches.csv  > profiles/rNN_launches.md
    python tools/ncu_summarize.py full     gpurun_out/X.ncu-rep       > profiles/rNN_ncu_full.md
"""
import collections
import csv
import io
import re
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "time"),
    ("launch__grid_size", "grid"),
    ("launch__registers_per_thread", "regs"),
    ("dram__bytes_read.sum", "dram_rd"),
    ("dram__bytes_write.sum", "dram_wr"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
    ("sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed", "fmaheavy%"),
    ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu%"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps%"),
    ("smsp__inst_executed.sum", "warp_inst"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem_conf"),
    ("lts__t_sector_hit_rate.pct", "l2hit%"),
]
STALLS = ["math_pipe_throttle", "not_selected", "wait", "dispatch_stall", "long_scoreboard", "short_scoreboard", "barrier",
          "mio_throttle", "lg_throttle", "selected"]


def short(name):
    name = re.sub(r"\(.*", "", name).replace("void ", "").replace("ppq::", "")
    return name.strip()


def launches(path):
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    hdr = rows[0]
    ki, vi, gi = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Grid Size")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        k = short(r[ki])
        a = agg.setdefault(k, [0, 0.0, r[gi]])
        a[0] += 1
        a[1] += float(r[vi].replace(",", "")) / 1e3
    tot = sum(a[1] for a in agg.values())
    print("| kernel | launches | total us | avg us | share | grid of first |")
    print("|---|---:|---:|---:|---:|---|")
    for k, (n, us, g) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{k}` | {n} | {us:.1f} | {us / n:.1f} | {us / tot:.3f} | {g} |")
    print(f"\n{sum(a[0] for a in agg.values())} launches, {tot / 1e3:.2f} ms (ncu-serialised, cold cache: compare SHARES, not absolutes)")


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    print("| kernel | " + " | ".join(s for _, s in KEYS) + " | top stalls (pc samples) |")
    print("|---|" + "---:|" * len(KEYS) + "---|")
    for r in data:
        cells = []
        for k, _ in KEYS:
            if k in idx:
                v = r[idx[k]].replace(",", "")
                try:
                    f = float(v)
                    v = f"{f:.4g}" if abs(f) < 1e6 else f"{f:.4e}"
                except ValueError:
                    pass
                cells.append(f"{v} {units[idx[k]]}".strip())
            else:
                cells.append("-")
        st = []
        for s in STALLS:
            k = f"smsp__pcsamp_warps_issue_stalled_{s}"
            if k in idx:
                st.append((float(r[idx[k]].replace(",", "") or 0), s))
        tot = sum(v for v, _ in st) or 1
        st = ", ".join(f"{s} {v / tot:.0%}" for v, s in sorted(st, reverse=True)[:4])
        print(f"| `{short(r[idx['Kernel Name']])}` | " + " | ".join(cells) + f" | {st} |")


def traffic(path, n_cts=54, tag=""):
    """One wave of ReEncrypt (15 launches in pipeline order, as captured by `--launch-skip 15 --launch-count 15` on
    tools/stage_times.py --clients 1 --reps 1) -> profiles/ncu_traffic.json layout: DRAM bytes per ct and stage."""
    import json
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    mult = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    seq = []
    for r in data:
        b = 0.0
        for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            b += float(r[idx[k]].replace(",", "")) * mult[units[idx[k]]]
        seq.append((short(r[idx["Kernel Name"]]), b))
    # rotate so that the wave starts at its first kernel (the inverse ROW pass that follows a ModDown pass)
    names = [n for n, _ in seq]
    start = next(i for i in range(len(seq)) if names[i].startswith("k_inv_row") and names[i - 1].startswith("k_fwd_row_moddown"))
    seq = seq[start:] + seq[:start]
    end = next(i for i in range(len(seq)) if seq[i][0].startswith("k_fwd_row_moddown"))
    seq = seq[: end + 1]  # exactly one wave
    stages = ["k_inv_row(c1)", "k_inv_col(c1,scale)", "k_bconv(modup)", "k_fwd_col(modup)", "k_fwd_row(modup)",
              "k_ip_tma|k_ip_agg", "k_inv_row(P)", "k_inv_col(P,scale)", "k_bconv(moddown)", "k_fwd_col(moddown)",
              "k_fwd_row_moddown"]
    per = dict.fromkeys(stages, 0.0)
    n_cts = int(n_cts)
    si, prev = -1, None
    for n, b in seq:  # consecutive launches of the same kernel family (the conversions) belong to one stage
        fam = n.split("<")[0]
        if fam != prev or not fam.startswith("k_bconv"):
            si += 1
        prev = fam
        per[stages[si]] += b / n_cts
    print(json.dumps({"c3": {"wave_cts": n_cts, "source": tag or path, "dram_bytes_per_ct": per}}, indent=1))


def traffic_agg(path, n_slots=54, cap_clients=8, job_clients=64, group=8, nb=4, tag=""):
    """Deferred-ModDown pipeline (ppq_agg_*): the capture holds the LAST client's add of a step of `cap_clients`
    clients, the group conversion (k_bconv_wide x2) and ppq_agg_finish (k_fwd_col2, k_fwd_row_moddown2), in launch
    order.  Prints the profiles/ncu_traffic.json + ncu_int_pipe.json content: DRAM bytes per ciphertext and stage, the
    path total for a GPU that aggregates `job_clients` clients, and fmaheavy % per stage."""
    import json
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    mult = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    n_slots, cap_clients, job_clients, group, nb = int(n_slots), int(cap_clients), int(job_clients), int(group), int(nb)
    stages = ["k_inv_row(c1)", "k_inv_col(c1,scale)", "k_bconv(modup)", "k_fwd_col(modup)", "k_fwd_row(modup)",
              "k_ip_tma|k_ip_agg", "k_inv_row(P)", "k_inv_col(P,scale)", "k_bconv(moddown)", "k_fwd_col(moddown)",
              "k_fwd_row_moddown"]
    per_launch = dict.fromkeys(stages, 0.0)
    fma = {}
    tms = dict.fromkeys(stages, 0.0)
    si, prev = -1, None
    for r in data:
        n = short(r[idx["Kernel Name"]])
        fam = n.split("<")[0]
        if fam != prev or not (fam.startswith("k_bconv") or fam.startswith("k_inv_row")):
            si += 1  # consecutive launches of the conversions / of the per-client inverse ROW pass are one stage
        prev = fam
        b = sum(float(r[idx[k]].replace(",", "")) * mult[units[idx[k]]] for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
        per_launch[stages[si]] += b
        f = float(r[idx["sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed"]].replace(",", ""))
        t = float(r[idx["gpu__time_duration.sum"]].replace(",", ""))
        fma[stages[si]] = max(fma.get(stages[si], 0.0), f)
        tms[stages[si]] += t
    assert si == len(stages) - 1, f"expected one add + group conversion + finish, got {si + 1} stages"
    div = {s: n_slots * nb for s in stages}  # the per-pass kernels cover nb clients x n_slots ciphertexts
    div["k_bconv(moddown)"] = n_slots * group
    per_ct_cap = {}
    path = 0.0
    for s in stages:
        d = div[s]
        if s in ("k_fwd_col(moddown)", "k_fwd_row_moddown"):
            per_ct_cap[s] = per_launch[s] / (n_slots * cap_clients)
            path += per_launch[s] / (n_slots * job_clients)
        else:
            per_ct_cap[s] = per_launch[s] / d
            path += per_launch[s] / d
    print(json.dumps({"traffic": {"c3": {"wave_cts": n_slots, "source": tag or path, "capture_clients": cap_clients,
                                         "job_clients": job_clients, "group": group, "clients_per_pass": nb, "dram_bytes_per_ct": per_ct_cap,
                                         "path_dram_bytes_per_ct": path}},
                      "int_pipe": fma}, indent=1))


if __name__ == "__main__":
    {"launches": launches, "full": full, "traffic": traffic, "traffic_agg": traffic_agg}[sys.argv[1]](*sys.argv[2:])
