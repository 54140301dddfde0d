#!/usr/bin/env python
"""Instruction-cache footprint of a kernel from an ncu report: distinct SASS instructions (16 B each) executed at
least `frac` x (launches of the hottest... no: x the per-unit count) times, grouped by source function.

    python tools/ncu_hotcode.py report.ncu-rep units file.cuh[,file2.cuh]
`units` = how many times the loop body ran in the launch (e.g. envs x substeps)."""
import csv, subprocess, sys, collections, re, os

rep, units, files = sys.argv[1], float(sys.argv[2]), sys.argv[3].split(",")
marks = {}
for f in files:
    path = f if os.path.exists(f) else os.path.join("modularevogym_b200/csrc", f)
    cur = "top"
    for n, line in enumerate(open(path), 1):
        m2 = re.match(r"^(?:template.*>\s*)?__(?:device|global)__.*?(\w+)\(", line)
        if m2: cur = m2.group(1)
        marks[(os.path.basename(f), n)] = cur
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
cur = None; line = None; kern = first = None
seen = {}
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if r[0] == "Function Name": kern = r[1]; first = first or kern; continue
    if r[0] == "Line No": continue
    if kern != first: continue
    if r[0] != "": line = int(r[0]); continue
    if len(r) > 7 and r[2].startswith("0x"):
        try: ex = int(r[7])
        except ValueError: continue
        seen[r[2]] = (marks.get((cur, line), cur), ex)
print(first, "distinct SASS:", len(seen), "=", len(seen) * 16 / 1024, "KB")
for frac in (0.5, 0.1, 0.01, 0.0):
    by = collections.Counter()
    for a, (g, ex) in seen.items():
        if ex > frac * units: by[g] += 1
    tot = sum(by.values())
    print(f"--- executed > {frac} x per unit: {tot} instr = {tot*16/1024:.1f} KB")
    print("   " + ", ".join(f"{g} {c*16/1024:.1f}" for g, c in by.most_common(12)))
