#!/usr/bin/env python
"""Writes / updates profiles/ncu_summary.json from an `ncu --set full` capture: the measured per-kernel figures that bench.py
quotes in its `roofline` block (DRAM bytes per launch, FP64 pipe, issue slots, stall shares, spills, bank conflicts).

    python tools/ncu_summary.py <report.ncu-rep | raw.csv> --key "gstep_kernel<256,2,2>" --workload obstacle20 --envs 592 \
           --profile profiles/r02c_gstep256_raw.csv [--stalls "barrier=0.40,short_sb=0.14"] [--commit <sha>]

A .ncu-rep is read with `ncu -i ... --page raw --csv` (and `--page source --csv` for the stall shares when --stalls is not
given); a raw CSV is one that command wrote earlier.
"""
import argparse
import collections
import csv
import io
import json
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def raw_rows(path):
    if path.endswith(".ncu-rep"):
        text = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    else:
        text = open(path).read()
    rows = list(csv.reader(io.StringIO(text)))
    return rows[0], rows[1], rows[2]


def stall_shares(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
    hdr, tot = None, collections.Counter()
    for r in csv.reader(io.StringIO(out)):
        if not r:
            continue
        if r[0] in ("Address", "Line No", "#"):
            hdr = r
            continue
        if hdr is None or len(r) != len(hdr):
            continue
        for k, v in zip(hdr, r):
            if k.startswith("stall_") and "Not Issued" not in k and v.isdigit():
                tot[k[6:]] += int(v)
    n = sum(tot.values()) or 1
    return {k: round(v / n, 3) for k, v in tot.most_common(6)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("report")
    ap.add_argument("--key", required=True)
    ap.add_argument("--workload", required=True)
    ap.add_argument("--envs", type=int, required=True)
    ap.add_argument("--profile", required=True, help="the file under profiles/ that holds this capture's raw page")
    ap.add_argument("--stalls", default="")
    ap.add_argument("--commit", default="")
    a = ap.parse_args()
    hdr, units, vals = raw_rows(a.report)

    def get(name, scale_units=True):
        for h, u, v in zip(hdr, units, vals):
            if h == name or h.endswith("." + name):
                x = float(v.replace(",", ""))
                if scale_units:
                    x *= {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "us": 1e-3, "ms": 1.0, "ns": 1e-6, "s": 1e3}.get(u, 1.0)
                return x
        return None

    ld, ldc = get("l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum"), get("l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum")
    entry = {
        "workload": a.workload, "envs": a.envs, "profile": a.profile,
        "commit": a.commit or subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip(),
        "duration_ms": get("gpu__time_duration.sum"),
        "dram_bytes_per_launch": int((get("dram__bytes_read.sum") or 0) + (get("dram__bytes_write.sum") or 0)),
        "fp64_pipe_frac": round((get("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active") or 0) / 100, 4),
        "issue_frac": round((get("smsp__issue_active.avg.pct_of_peak_sustained_active") or 0) / 100, 4),
        "warps_per_sm": round((get("sm__warps_active.avg.pct_of_peak_sustained_active") or 0) / 100 * 64, 1),
        "registers_per_thread": get("launch__registers_per_thread", False),
        "local_ld_st": [int(get("smsp__sass_inst_executed_op_local_ld.sum", False) or 0), int(get("smsp__sass_inst_executed_op_local_st.sum", False) or 0)],
        "shared_ld_conflict_share": round(ldc / ld, 3) if ld and ldc is not None else None,
        "note": "ncu replays the launch cold-cache and serialised: durations are for shares, not for throughput",
    }
    if a.stalls:
        entry["stall_shares"] = {k: float(v) for k, v in (kv.split("=") for kv in a.stalls.split(","))}
    elif a.report.endswith(".ncu-rep"):
        entry["stall_shares"] = stall_shares(a.report)
    path = os.path.join(ROOT, "profiles", "ncu_summary.json")
    doc = {"kernels": {}}
    if os.path.exists(path):
        doc = json.load(open(path))
    doc["kernels"][a.key] = entry
    json.dump(doc, open(path, "w"), indent=1)
    print(json.dumps(entry, indent=1))


if __name__ == "__main__":
    main()
