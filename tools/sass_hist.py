#!/usr/bin/env python
"""SASS opcode histogram of the hot kernels in libstnls_b200.so (cuobjdump -sass): which Blackwell instructions the
library really contains -- UTMALDG (TMA), SYNCS (mbarrier), RED (vector reductions), FFMA/FADD/FMUL, LDS/STS, SHFL ...
usage: python tools/sass_hist.py [kernel-name regex ...] > profiles/sass_hot_kernels.txt"""
import collections, re, subprocess, sys, os
so = os.path.join(os.path.dirname(__file__), "..", "stnls_b200", "libstnls_b200.so")
pats = sys.argv[1:] or ["search_fwd_tile_kernel", "tile_merge", "search_bwd_kernel", "wpsum_bwd_kernel", "wpsum_fwd_tile_kernel",
                        "seg_copy_kernel", "search_fwd_tab_kernel"]
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
funcs = re.split(r"\n\s*Function : ", txt)[1:]
demangle = lambda n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()
KEYS = ["UTMALDG", "UTMASTG", "UBLKCP", "SYNCS", "UTCHMMA", "LDTM", "REDG", "ATOMG", "ATOMS", "LDG", "STG", "LDS", "STS", "LDSM",
        "FFMA", "FADD", "FMUL", "HFMA2", "IMAD", "IADD3", "LOP3", "SHFL", "REDUX", "VOTE", "BAR", "WARPSYNC", "MOV", "BRA",
        "LDL", "STL", "DADD", "DMUL", "DFMA", "F2I", "I2F"]
print("# cuobjdump -sass stnls_b200/libstnls_b200.so: static instruction counts per kernel (sm_100a)")
print("# UTMALDG = cp.async.bulk.tensor (TMA load), SYNCS = mbarrier ops, REDG = red.global.add(.v4.f32)")
tot_all = collections.Counter()
for f in funcs:
    name = f.split("\n")[0].strip()
    if not any(re.search(p, name) for p in pats):
        continue
    ins = re.findall(r"/\*[0-9a-f]{4,6}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", f)
    c = collections.Counter(ins)
    tot_all.update(c)
    print(f"\n{demangle(name)[:150]}\n  total {len(ins)}  " + "  ".join(f"{k}={c[k]}" for k in KEYS if c[k]))
print("\n# all listed kernels together: " + "  ".join(f"{k}={tot_all[k]}" for k in KEYS if tot_all[k]))
