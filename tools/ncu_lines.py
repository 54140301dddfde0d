#!/usr/bin/env python
"""Per-source-line summary of an ncu report's source page (instructions executed, stall samples).

    python tools/ncu_lines.py gpurun_out/prof.ncu-rep [top_n] [kernel-substring]
"""
import csv
import subprocess
import sys


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    cur_file, hdr, lines, first_kernel = None, None, [], None
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            cur_file = r[1].split("/")[-1]
            continue
        if r[0] == "Function Name":
            if first_kernel is None:
                first_kernel = r[1]
            cur_kernel = r[1]
            continue
        if r[0] == "Line No":
            hdr = r
            continue
        if hdr is None or r[0] == "" or cur_kernel != first_kernel:
            continue
        d = dict(zip(hdr, r))
        try:
            inst = int(d["Instructions Executed"]); samp = int(d["# Samples"])
        except (KeyError, ValueError):
            continue
        stalls = {k[6:]: int(v) for k, v in d.items() if k.startswith("stall_") and "Not Issued" not in k and v.isdigit() and int(v)}
        lines.append((cur_file, int(r[0]), r[1].strip()[:90], inst, samp, stalls))
    tot_i = sum(l[3] for l in lines) or 1
    tot_s = sum(l[4] for l in lines) or 1
    print(f"kernel {first_kernel}: total warp-instructions {tot_i}, samples {tot_s}")
    print("--- by instructions executed")
    for f, n, src, inst, samp, st in sorted(lines, key=lambda l: -l[3])[:top]:
        print(f"{100 * inst / tot_i:5.1f}% inst {100 * samp / tot_s:5.1f}% samp  {f}:{n:<4d} {src}")
    print("--- by stall samples")
    for f, n, src, inst, samp, st in sorted(lines, key=lambda l: -l[4])[:top]:
        top_st = ", ".join(f"{k}={v}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:3])
        print(f"{100 * samp / tot_s:5.1f}% samp {100 * inst / tot_i:5.1f}% inst  {f}:{n:<4d} {src[:60]}  [{top_st}]")


if __name__ == "__main__":
    main()
