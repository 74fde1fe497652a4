#!/usr/bin/env python
"""Per-source-line instruction / stall-sample shares from `ncu -i X.ncu-rep --page source --print-source cuda,sass --csv`.
usage: python tools/ncu_lines.py report.ncu-rep [top]"""
import csv, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 50
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
fname = ""; h = None; out = []
for r in rows:
    if not r: continue
    if r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if r[0] == "Line No": h = r; ci = h.index("Instructions Executed"); sa = h.index("# Samples"); continue
    if h is None or len(r) <= ci or r[2] != "-": continue   # cuda line rows have Address "-"
    try: out.append((int(r[ci]), int(r[sa] or 0), fname, r[0], r[1].strip()[:100]))
    except ValueError: pass
tot = sum(o[0] for o in out); ts = sum(o[1] for o in out)
print("total warp-instructions", tot, "samples", ts)
for n, s, f, l, src in sorted(out, reverse=True)[:top]:
    print(f"{n/tot*100:5.1f}% inst {s/max(ts,1)*100:5.1f}% samp  {f}:{l}: {src}")
