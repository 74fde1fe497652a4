"""CPU: replay the committed golden vectors -- outputs of the REFERENCE's own kernels on a B200
(tests/golden/make_golden.py) -- through the CPU oracle.  This pins the oracle without a GPU.
int path: every candidate distance bit-exact, every index equal.  float path: 1e-5.  Gradients 1e-3
(the reference accumulates them with atomics)."""
import glob
import os

import numpy as np
import pytest
import torch as th

from util import assert_close, assert_topk_equal

HERE = os.path.dirname(os.path.abspath(__file__))
FILES = sorted(glob.glob(os.path.join(HERE, "golden", "nls_*.npz")))
CFG = {
    "nls_int_prod": dict(itype="int", dist_type="prod", stride1=1, self_action="anchor", k=6),
    "nls_int_l2": dict(itype="int", dist_type="l2", stride1=1, self_action=None, k=6),
    "nls_float_l2_half": dict(itype="float", dist_type="l2", stride1=0.5, self_action="anchor", k=6),
    "nls_float_prod_1p1": dict(itype="float", dist_type="prod", stride1=1.1, self_action=None, k=6),
}
HD, ws, wt, ps = 2, 3, 1, 3


def load(path):
    z = np.load(path)
    return {k: th.from_numpy(z[k]) for k in z.files}


def test_golden_files_present():
    assert len(FILES) == 4


@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(f)[:-4] for f in FILES])
def test_oracle_reproduces_reference(path, oracle):
    c = CFG[os.path.basename(path)[:-4]]
    z = load(path)
    exact = c["itype"] == "int"
    flt = not exact
    kw = dict(nheads=HD, stride0=1, stride1=c["stride1"], dist_type=c["dist_type"], self_action=c["self_action"],
              itype=c["itype"])
    d, i, (df, if_) = oracle.nls_forward(z["q"], z["k"], z["flows"], ws, wt, ps, c["k"], return_full=True, **kw)
    if exact:
        assert th.equal(df, z["dists_full"]) and th.equal(if_, z["inds_full"])
    else:
        fin = th.isfinite(df)
        assert th.equal(fin, th.isfinite(z["dists_full"]))
        assert_close(df[fin], z["dists_full"][fin], 1e-5, 1e-5, "dists_full")
        assert_close(if_, z["inds_full"], 1e-5, 0, "inds_full")
    assert_topk_equal(d, i, z["dists"], z["inds"], exact=exact, what="top-K")
    # search backward at the reference's indices
    g0, g1, gf = oracle.search_backward(z["q"], z["k"], z["flows"], z["gd"], z["gi"], z["inds"], wt, ps, nheads=HD,
                                        stride0=1, dist_type=c["dist_type"], itype=c["itype"], flows_grad=flt)
    assert_close(g0, z["g0"], 1e-3, 1e-3, "grad_q")
    assert_close(g1, z["g1"], 1e-3, 1e-3, "grad_k")
    if flt:
        assert_close(gf, z["gflows"], 2e-3, 2e-3, "grad_flows")
    # refinement (wr=3) of the first three neighbours
    fl3 = z["inds"][..., :3, :].contiguous()
    dr, ir, ks, rf = oracle.refine_forward(z["q"], z["k"], fl3, ws, wt, 3, 8, -1, ps, **kw)
    assert_topk_equal(dr, ir, z["r_dists"], z["r_inds"], exact=exact, what="refine top-K")
    if flt:
        assert th.equal(rf.bool(), z["r_reflect"].bool())
        r0, r1, rfl = oracle.search_backward(z["q"], z["k"], fl3, z["r_gd"], z["r_gi"], z["r_inds"], wt, ps, nheads=HD,
                                             stride0=1, dist_type=c["dist_type"], itype="float", flows_grad=True,
                                             kselect=z["r_kselect"], reflect=z["r_reflect"], refine=True)
        assert_close(r0, z["r_g0"], 1e-3, 1e-3, "refine grad_q")
        assert_close(r1, z["r_g1"], 1e-3, 1e-3, "refine grad_k")
        assert_close(rfl, z["r_gflows"], 2e-3, 2e-3, "refine grad_flows")
    # weighted patch sum
    out, counts = oracle.gather_add_forward(z["v"], z["w"], z["inds"], ps, 1, itype=c["itype"])
    assert th.equal(counts, z["counts"])
    assert_close(out, z["agg_out"], 1e-5, 1e-5, "wpsum out")
    gv, gw, gfa = oracle.gather_add_backward(z["go"], z["v"], z["w"], z["inds"], counts, ps, 1, itype=c["itype"])
    assert_close(gv, z["gv"], 1e-3, 1e-3, "wpsum grad_v")
    assert_close(gw, z["gw"], 1e-3, 1e-3, "wpsum grad_w")
    if flt:
        assert_close(gfa, z["gf_agg"], 2e-3, 2e-3, "wpsum grad_flows")
