#!/usr/bin/env python
"""One small forward+backward of every hot kernel for compute-sanitizer:
    compute-sanitizer --tool memcheck  python tools/sanitize_case.py
    compute-sanitizer --tool racecheck python tools/sanitize_case.py
Covers the tile forward kernel (TMA box, mirror fill at the border, shared-memory pool, per-warp transpose scratch),
the table kernel (int path), search backward, weighted patch sum forward (tile kernel: 9 conflict-free phases) and
backward, the softmax kernels and the layout kernels -- at sizes the tools finish in a minute."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch as th
import stnls_b200

dev = th.device("cuda:0")
g = th.Generator().manual_seed(0)


def run(F, HD, H, W, T, ws, wt, ps, K, s1, dist, act, itype):
    q, k, v = [th.randn(1, T, F, H, W, generator=g).to(dev).requires_grad_(True) for _ in range(3)]
    shape = (1, 1, T, 2 * wt, 2, H, W)
    fl = th.rand(shape, generator=g) / 2 + th.randint(-3, 3, shape, generator=g) + 0.2
    fl = (fl.round() if itype == "int" else fl).to(dev).requires_grad_(itype == "float")
    search = stnls_b200.search.NonLocalSearch(ws, wt, ps, K, nheads=HD, stride0=1, stride1=s1, dist_type=dist,
                                              self_action=act, itype=itype)
    agg = stnls_b200.agg.WeightedPatchSum(ps=ps, stride0=1, itype=itype)
    d, i = search(q, k, fl)
    w = stnls_b200.normz.softmax.softmax_scaled(d, -10.0 if dist == "l2" else 10.0)
    out = agg(v, w, i)
    out.sum().backward()
    th.cuda.synchronize()
    assert bool(th.isfinite(out).all()) and bool(th.isfinite(q.grad).all())
    print("ok", F, HD, H, W, ws, ps, s1, dist, itype, flush=True)


run(16, 2, 35, 37, 5, 5, 2, 3, 10, 0.5, "l2", "anchor", "float")   # C2 shape: tile kernel, Fp = 8, stride1 = 0.5
run(32, 2, 36, 34, 3, 9, 1, 3, 20, 1.0, "prod", None, "float")     # C4 shape: tile kernel, Fp = 16
run(32, 2, 36, 34, 3, 9, 1, 1, 9, 1.0, "prod", None, "float")      # C3 shape
run(16, 2, 24, 24, 5, 5, 2, 3, 10, 1.0, "l2", "anchor", "int")     # C1 shape: table kernel, int
print("sanitize_case done")
