import sys,json
for l in sys.stdin:
    if l.startswith("{"):
        d=json.loads(l); print(round(d["ms_per_step"],3), "e2e", round(d["e2e"]["ms_per_step"],3), {k:round(v["ms_per_launch"],4) for k,v in d["kernels"].items()})
    else: print(l.rstrip()[-300:])
