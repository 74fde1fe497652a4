"""Generate golden vectors by running the REFERENCE's own kernels (oracle/_ref, built from
/root/reference/lib/csrc by oracle/build_ref.py) on a GPU.  Run on the B200 box:

    python tests/golden/make_golden.py gpurun_out/golden      # then copy *.npz into tests/golden/

Each .npz holds the seeded inputs and the reference outputs (the full candidate tensors the reference
kernel writes, the anchored top-K results, gradients).  tests/test_golden.py replays them through the CPU
oracle (-m "not gpu") and tests/test_ref_parity_gpu.py through the CUDA path.
"""
import os
import sys

import numpy as np
import torch as th

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import ref_cuda  # noqa: E402
from util import make_flows, make_vids  # noqa: E402

NLS = [
    dict(name="nls_int_prod", itype="int", dist_type="prod", stride1=1, flow="int", self_action="anchor", k=6),
    dict(name="nls_int_l2", itype="int", dist_type="l2", stride1=1, flow="int", self_action=None, k=6),
    dict(name="nls_float_l2_half", itype="float", dist_type="l2", stride1=0.5, flow="float", self_action="anchor", k=6),
    dict(name="nls_float_prod_1p1", itype="float", dist_type="prod", stride1=1.1, flow="float", self_action=None, k=6),
]
B, T, F, H, W, HD, ws, wt, ps = 1, 4, 8, 14, 12, 2, 3, 1, 3


def np_(t):
    return None if t is None else t.detach().cpu().numpy()


def main(outdir):
    os.makedirs(outdir, exist_ok=True)
    dev = th.device("cuda:0")
    for c in NLS:
        q, k, v = make_vids(B, T, F, H, W, seed=101)
        flows = make_flows(B, 1, T, 2 * wt + 1, H, W, c["flow"], seed=102)
        kw = dict(nheads=HD, stride0=1, stride1=c["stride1"], dist_type=c["dist_type"],
                  self_action=c["self_action"], itype=c["itype"])
        d, i, (df, if_) = ref_cuda.nls_forward(q.to(dev), k.to(dev), flows.to(dev), ws, wt, ps, c["k"],
                                               return_full=True, **kw)
        gen = th.Generator().manual_seed(103)
        gd = th.randn(d.shape, generator=gen)
        gi = th.randn(i.shape, generator=gen)
        gd = th.where(th.isfinite(d.cpu()), gd, th.zeros_like(gd))
        flt = c["itype"] == "float"
        g0, g1, gf = ref_cuda.search_backward(q.to(dev), k.to(dev), flows.to(dev), gd.to(dev), gi.to(dev), i, wt, ps,
                                              nheads=HD, stride0=1, dist_type=c["dist_type"], itype=c["itype"],
                                              flows_grad=flt, dists=d)
        out = dict(q=np_(q), k=np_(k), flows=np_(flows), dists=np_(d), inds=np_(i), dists_full=np_(df),
                   inds_full=np_(if_), gd=np_(gd), gi=np_(gi), g0=np_(g0), g1=np_(g1))
        if gf is not None:
            out["gflows"] = np_(gf)
        # refinement of the search result (wr=3) and its gradients
        rkw = dict(nheads=HD, stride0=1, stride1=c["stride1"], dist_type=c["dist_type"],
                   self_action=c["self_action"], itype=c["itype"])
        dr, ir, ks, rf = ref_cuda.refine_forward(q.to(dev), k.to(dev), i[..., :3, :].contiguous(), ws, wt, 3, 8, -1,
                                                 ps, **rkw)
        out.update(r_dists=np_(dr), r_inds=np_(ir))
        if ks is not None:
            out.update(r_kselect=np_(ks), r_reflect=np_(rf))
            gdr = th.randn(dr.shape, generator=gen)
            gir = th.randn(ir.shape, generator=gen)
            gdr = th.where(th.isfinite(dr.cpu()), gdr, th.zeros_like(gdr))
            r0, r1, rfl = ref_cuda.search_backward(q.to(dev), k.to(dev), i[..., :3, :].contiguous(), gdr.to(dev),
                                                   gir.to(dev), ir, wt, ps, nheads=HD, stride0=1,
                                                   dist_type=c["dist_type"], itype="float", flows_grad=True,
                                                   kselect=ks, reflect=rf, refine=True)
            out.update(r_gd=np_(gdr), r_gi=np_(gir), r_g0=np_(r0), r_g1=np_(r1), r_gflows=np_(rfl))
        # weighted patch sum on the search result
        w = th.softmax((-1 if c["dist_type"] == "l2" else 1) * d, -1)
        o, counts = ref_cuda.gather_add_forward(v.to(dev), w, i, ps, 1, itype=c["itype"])
        go = th.randn(o.shape, generator=gen)
        gv, gw, gfa = ref_cuda.gather_add_backward(go.to(dev), v.to(dev), w, i, counts, ps, 1, itype=c["itype"])
        out.update(v=np_(v), w=np_(w), agg_out=np_(o), counts=np_(counts), go=np_(go), gv=np_(gv), gw=np_(gw))
        if gfa is not None:
            out["gf_agg"] = np_(gfa)
        np.savez_compressed(os.path.join(outdir, c["name"] + ".npz"), **out)
        print("wrote", c["name"], {k_: v_.shape for k_, v_ in out.items() if v_ is not None and v_.ndim > 5})


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/golden")
