#!/usr/bin/env python
"""Instruction shares per (file, line range) phase. usage: ncu_phases.py rep file:lo-hi=name ..."""
import csv, subprocess, sys
rep = sys.argv[1]
specs = []
for a in sys.argv[2:]:
    rng, name = a.split("="); f, r = rng.split(":"); lo, hi = r.split("-"); specs.append((f, int(lo), int(hi), name))
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
fname = ""; h = None; acc = {}; tot = 0; byfile = {}
for r in rows:
    if not r: continue
    if r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if r[0] == "Line No": h = r; ci = h.index("Instructions Executed"); sa = h.index("# Samples"); continue
    if h is None or len(r) <= ci or r[2] != "-": continue
    try: n = int(r[ci]); s = int(r[sa] or 0); l = int(r[0])
    except ValueError: continue
    tot += n
    key = "other:" + fname
    for f, lo, hi, name in specs:
        if f == fname and lo <= l <= hi: key = name; break
    a = acc.setdefault(key, [0, 0]); a[0] += n; a[1] += s
ts = sum(a[1] for a in acc.values())
for k, (n, s) in sorted(acc.items(), key=lambda kv: -kv[1][0]):
    print(f"{n/tot*100:5.1f}% inst {s/max(ts,1)*100:5.1f}% samp  {k}")
