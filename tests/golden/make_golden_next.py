"""Golden vectors for the next-row components (SURVEY.md 8f N1 search_flow, N3 NonLocalGather), produced by
the REFERENCE's own kernels (oracle/_ref) on a GPU.  Run on the B200 box:

    python tests/golden/make_golden_next.py gpurun_out/golden      # then copy next_*.npz into tests/golden/

tests/test_golden_next.py replays them through the CPU oracle (-m "not gpu")."""
import os
import sys

import numpy as np
import torch as th

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import ref_cuda  # noqa: E402


def np_(t):
    return None if t is None else t.detach().cpu().numpy()


def flow_case(outdir, dev, name, B, T, H, W, wt, stride0, scale, seed):
    g = th.Generator().manual_seed(seed)
    ff = th.randn(B, T, 2, H, W, generator=g) * scale
    bf = th.randn(B, T, 2, H, W, generator=g) * scale
    fl = ref_cuda.search_flow_forward(ff.to(dev), bf.to(dev), wt, stride0)
    gfl = th.randn(fl.shape, generator=g)
    gf, gb = ref_cuda.search_flow_backward(gfl.to(dev), ff.to(dev), bf.to(dev), fl, wt, stride0)
    np.savez_compressed(os.path.join(outdir, name + ".npz"), fflow=np_(ff), bflow=np_(bf), flows=np_(fl),
                        gflows=np_(gfl), gfflow=np_(gf), gbflow=np_(gb), wt=wt, stride0=stride0)
    print("wrote", name, tuple(fl.shape))


def gather_case(outdir, dev, name, itype, ps, stride0, seed):
    B, T, F, H, W, HD, K = 1, 3, 4, 12, 10, 2, 3
    g = th.Generator().manual_seed(seed)
    v = th.randn(B, T, F * HD, H, W, generator=g)
    nH, nW = (H - 1) // stride0 + 1, (W - 1) // stride0 + 1
    w = th.rand(B, HD, T, nH, nW, K, generator=g)
    inds = th.zeros(B, HD, T, nH, nW, K, 3)
    inds[..., 0] = th.randint(-1, 2, inds[..., 0].shape, generator=g).float()
    inds[..., 1:] = th.randint(-4, 5, inds[..., 1:].shape, generator=g).float()
    if itype == "float":
        inds[..., 1:] += th.rand(inds[..., 1:].shape, generator=g) - 0.5
    st, counts = ref_cuda.gather_forward(v.to(dev), w.to(dev), inds.to(dev), ps, stride0, itype=itype)
    gs = th.randn(st.shape, generator=g)
    gv, gw, gi = ref_cuda.gather_backward(gs.to(dev), v.to(dev), w.to(dev), inds.to(dev), counts, ps, stride0,
                                          itype=itype)
    out = dict(v=np_(v), w=np_(w), inds=np_(inds), stack=np_(st), counts=np_(counts[0, 0]), gs=np_(gs), gv=np_(gv),
               gw=np_(gw), ps=ps, stride0=stride0)
    if gi is not None:
        out["gi"] = np_(gi)
    np.savez_compressed(os.path.join(outdir, name + ".npz"), **out)
    print("wrote", name, tuple(st.shape))


def main(outdir):
    os.makedirs(outdir, exist_ok=True)
    dev = th.device("cuda:0")
    flow_case(outdir, dev, "next_search_flow_wt2", 1, 5, 12, 10, 2, 1, 1.5, 201)
    flow_case(outdir, dev, "next_search_flow_wt1_s2", 2, 4, 11, 13, 1, 2, 3.0, 202)
    gather_case(outdir, dev, "next_gather_int", "int", 3, 1, 203)
    gather_case(outdir, dev, "next_gather_float", "float", 3, 1, 204)


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/golden")
