#!/usr/bin/env python
"""Aggregate an ncu source page by marked code regions ("// =====" comments and function starts).
    python tools/ncu_regions.py report.ncu-rep file.cuh[,file2.cuh]"""
import csv, subprocess, sys, collections, re, os

rep = sys.argv[1]
files = sys.argv[2].split(",")
marks = {}
for f in files:
    path = f if os.path.exists(f) else os.path.join("modularevogym_b200/csrc", f)
    cur = "top"
    for n, line in enumerate(open(path), 1):
        m = re.search(r"// =+ (.*)", line)
        if m: cur = m.group(1)[:40]
        m2 = re.match(r"^(?:template.*>\s*)?__(?:device|global)__.*?(\w+)\(", line)
        if m2: cur = "fn " + m2.group(1)
        marks[(os.path.basename(f), n)] = cur
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = cur = None
kern = first = None
inst = collections.Counter(); samp = collections.Counter(); stalls = collections.defaultdict(collections.Counter)
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if r[0] == "Function Name": kern = r[1]; first = first or kern; continue
    if r[0] == "Line No": hdr = r; continue
    if hdr is None or r[0] == "" or kern != first: continue
    d = dict(zip(hdr, r))
    try: i = int(d["Instructions Executed"]); s = int(d["# Samples"])
    except Exception: continue
    g = marks.get((cur, int(r[0])), cur)
    inst[g] += i; samp[g] += s
    for k, v in d.items():
        if k.startswith("stall_") and "Not Issued" not in k and v.isdigit(): stalls[g][k[6:]] += int(v)
ti = sum(inst.values()) or 1; ts = sum(samp.values()) or 1
print(first, "total inst", ti, "samples", ts)
for g, v in inst.most_common():
    top = ", ".join(f"{k}={100*c/ts:.1f}" for k, c in stalls[g].most_common(3))
    print(f"{g:42s} inst {100*v/ti:5.1f}%  samples {100*samp[g]/ts:5.1f}%   [{top}]")
tot = collections.Counter()
for g in stalls: tot.update(stalls[g])
print({k: round(100*v/sum(tot.values()), 1) for k, v in tot.most_common(9)})
