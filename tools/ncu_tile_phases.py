#!/usr/bin/env python
"""Phase shares (warp-instructions, stall samples) of the tile forward kernel from an ncu --set full capture with
--import-source on.  Phases are found from marker comments in search_fwd_tile.cuh, so the table follows the source.
usage: python tools/ncu_tile_phases.py report.ncu-rep units   (units = (query, frame slot) groups one warp pass handles)"""
import csv, re, subprocess, sys, os

def lines(rep):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    fname = ""; h = None; out = []
    for r in rows:
        if not r: continue
        if r[0] == "File Path": fname = r[1].split("/")[-1]; continue
        if r[0] == "Line No": h = r; ci = h.index("Instructions Executed"); sa = h.index("# Samples"); continue
        if h is None or len(r) <= ci or r[2] != "-": continue
        try: out.append((fname, int(r[0]), int(r[ci]), int(r[sa] or 0)))
        except ValueError: pass
    return out

src = open(os.path.join(os.path.dirname(__file__), "..", "stnls_b200", "csrc", "search_fwd_tile.cuh")).read().splitlines()
def find(pat, start=0):
    for i in range(start, len(src)):
        if pat in src[i]: return i + 1
    raise SystemExit("marker not found: " + pat)
loop = find("for (int st = 0; st < W_t; st++) {", find("// ---- frame slots: TMA box"))
marks = [("core (interpolate + accumulate)", find("__device__ __forceinline__ float acc1"), find("// ---- pool compaction")),
         ("tile_merge (pool compaction)", find("// ---- pool compaction"), find("__global__ void __launch_bounds__(TILE_THREADS")),
         ("prologue (geometry, q tile, TMA issue)", find("// ---- which tile ----"), loop),
         ("frame head (mbarrier wait, mirror fill)", loop, find("for (int pass = 0;", loop)),
         ("q patch load", find("// query patch, this lane's channel"), find("// one integer-spaced (sub-)window")),
         ("flush (transpose-sum, key, append)", find("auto flush = [&]"), find("tile_core<WH, WW, PS, DIST, FP, FL>(bp")),
         ("output (final compaction call, dists/inds)", find("// ---- outputs: the sorted list"), find("// ---- queries this tile could not stage"))]
rep = sys.argv[1]; units = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
L = lines(rep)
tot = sum(x[2] for x in L); ts = sum(x[3] for x in L)
ph = {}
for f, l, n, s in L:
    k = "other files (intrinsics, geometry helpers)"
    if f == "search_fwd_tile.cuh":
        k = "tile: pass loop glue"
        for name, a, b in marks:
            if a <= l < b: k = name; break
    elif f == "search_fwd_direct.cuh": k = "direct_query fallback"
    ph.setdefault(k, [0, 0]); ph[k][0] += n; ph[k][1] += s
print(f"{os.path.basename(rep)}: {tot} warp-instructions, {tot/units:.1f} per unit, {ts} samples")
for k, (n, s) in sorted(ph.items(), key=lambda kv: -kv[1][0]):
    print(f"  {k:42s} {n/tot*100:5.1f}% inst ({n/units:8.1f}/unit) {s/max(ts,1)*100:5.1f}% samples")
