#!/usr/bin/env python
"""Per-source-line shared-memory wavefronts / global L1 tag requests from an ncu --set full report."""
import csv, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
fname = ""; h = None; out = []
for r in rows:
    if not r: continue
    if r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if r[0] == "Line No":
        h = r; ci = h.index("Instructions Executed"); ws = h.index("L1 Wavefronts Shared"); wi = h.index("L1 Wavefronts Shared Ideal")
        tg = h.index("L1 Tag Requests Global"); continue
    if h is None or len(r) <= ws or r[2] != "-": continue
    try: out.append((int(r[ws] or 0), int(r[wi] or 0), int(r[tg] or 0), int(r[ci] or 0), fname, r[0], r[1].strip()[:90]))
    except ValueError: pass
tw = sum(o[0] for o in out); tt = sum(o[2] for o in out); ti = sum(o[3] for o in out)
print("shared wavefronts", tw, "ideal", sum(o[1] for o in out), "global tag requests", tt, "instructions", ti)
for w, i, t, n, f, l, src in sorted(out, reverse=True)[:top]:
    print(f"smem {w/tw*100:5.1f}% (ideal {i/max(w,1)*100:3.0f}%) gtag {t/max(tt,1)*100:5.1f}% inst {n/ti*100:5.1f}%  {f}:{l}: {src}")
print("-- by global tag requests")
for w, i, t, n, f, l, src in sorted(out, key=lambda o: -o[2])[:8]:
    print(f"gtag {t/max(tt,1)*100:5.1f}%  {f}:{l}: {src}")
