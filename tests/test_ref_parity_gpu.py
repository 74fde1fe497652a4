"""GPU: pin the CPU oracle AND the CUDA path against the REFERENCE's own kernels (oracle/_ref, compiled
from /root/reference/lib/csrc by oracle/build_ref.py; SURVEY.md 8c "O1").  Skipped when the prebuilt
reference library did not travel to the box."""
import pytest
import torch as th

import stnls_b200
from util import assert_close, assert_topk_equal, make_flows, make_vids

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def refc(cuda):
    from oracle import ref_cuda
    if not ref_cuda.available():
        pytest.skip("oracle/_ref/stnls_cuda_ref*.so not present")
    ref_cuda.mod()
    return ref_cuda


CASES = [
    dict(name="int_prod", itype="int", dist_type="prod", stride1=1, flow="int", self_action="anchor", ws=5, wt=2, ps=3),
    dict(name="int_l2", itype="int", dist_type="l2", stride1=1, flow="int", self_action=None, ws=5, wt=2, ps=3),
    dict(name="int_l2_s2", itype="int", dist_type="l2", stride1=2, flow="int", self_action="anchor", ws=3, wt=1, ps=1,
         stride0=2),
    dict(name="float_l2_half", itype="float", dist_type="l2", stride1=0.5, flow="float", self_action="anchor", ws=5,
         wt=2, ps=3),
    dict(name="float_prod_1p1", itype="float", dist_type="prod", stride1=1.1, flow="float", self_action=None, ws=3,
         wt=1, ps=3, stride0=2),
    dict(name="float_prod_raw_ps3", itype="float", dist_type="prod", stride1=1.0, flow="float", self_action="anchor", ws=5,
         wt=1, ps=3, F=32),
    dict(name="float_prod_raw_ps1", itype="float", dist_type="prod", stride1=1.0, flow="float", self_action=None, ws=5,
         wt=1, ps=1, F=32),
    dict(name="float_bigflow", itype="float", dist_type="l2", stride1=1.0, flow="float", self_action="anchor", ws=3,
         wt=1, ps=3, flow_scale=6),
]
B, T, F, H, W, HD, K = 1, 5, 8, 24, 20, 2, 8


def inputs(c):
    q, k, v = make_vids(B, T, c.get("F", F), H, W, seed=61)
    s0 = c.get("stride0", 1)
    nH, nW = (H - 1) // s0 + 1, (W - 1) // s0 + 1
    flows = make_flows(B, 1, T, 2 * c["wt"] + 1, nH, nW, c["flow"], seed=62) * c.get("flow_scale", 1)
    kw = dict(nheads=HD, stride0=s0, stride1=c["stride1"], dist_type=c["dist_type"],
              self_action=c["self_action"], itype=c["itype"])
    return q, k, v, flows, kw


@pytest.mark.parametrize("c", CASES, ids=[c["name"] for c in CASES])
def test_reference_vs_oracle_vs_cuda_forward(c, cuda, oracle, refc):
    q, k, v, flows, kw = inputs(c)
    exact = c["itype"] == "int"
    dr, ir, (dfr, ifr) = refc.nls_forward(q.to(cuda), k.to(cuda), flows.to(cuda), c["ws"], c["wt"], c["ps"], K,
                                          return_full=True, **kw)
    do, io, (dfo, ifo) = oracle.nls_forward(q, k, flows, c["ws"], c["wt"], c["ps"], K, return_full=True, **kw)
    # (1) the oracle restates the reference kernel: every candidate, before anchor/top-K
    if exact:
        assert th.equal(dfr.cpu(), dfo), "oracle dists_full != reference (int path must be bit-exact)"
        assert th.equal(ifr.cpu(), ifo), "oracle inds_full != reference"
    else:
        fin = th.isfinite(dfo)
        assert th.equal(th.isfinite(dfr.cpu()), fin)
        assert_close(dfr.cpu()[fin], dfo[fin], 1e-5, 1e-5, "oracle dists_full vs reference")
        assert_close(ifr, ifo, 1e-5, 0, "oracle inds_full vs reference")
    assert_topk_equal(do, io, dr, ir, exact=exact, what="oracle vs reference top-K")
    # (2) the CUDA path against the reference
    search = stnls_b200.search.NonLocalSearch(c["ws"], c["wt"], c["ps"], K, **kw)
    d, i = search(q.to(cuda), k.to(cuda), flows.to(cuda))
    assert_topk_equal(d, i, dr, ir, exact=exact, what="cuda vs reference top-K")


MENU = [("anchor_each", "all"), ("anchor_each", "each"), ("remove", "all"), ("remove_ref_frame", "each"),
        ("anchor_and_remove_ref_frame", "all"), ("anchor", "each"), (None, "none")]


@pytest.mark.parametrize("itype", ["int", "float"])
@pytest.mark.parametrize("self_action,topk_mode", MENU, ids=[f"{a}-{m}" for a, m in MENU])
def test_reference_menu(self_action, topk_mode, itype, cuda, oracle, refc):
    """self_action / topk_mode menus: the reference's own anchor_self_time / anchor_self kernels + selections
    (oracle/ref_cuda.py) against the oracle and against the CUDA path."""
    c = dict(CASES[0] if itype == "int" else CASES[3], ws=3, wt=1, flow_scale=2)
    q, k, v, flows, kw = inputs(c)
    kw["self_action"] = self_action
    kk = -1 if topk_mode == "none" else 4
    exact = itype == "int"
    dr, ir = refc.nls_forward(q.to(cuda), k.to(cuda), flows.to(cuda), c["ws"], c["wt"], c["ps"], kk,
                              topk_mode=topk_mode, **kw)
    do, io = oracle.nls_forward(q, k, flows, c["ws"], c["wt"], c["ps"], kk, topk_mode=topk_mode, **kw)
    assert_topk_equal(do, io, dr, ir, exact=exact, what="oracle vs reference")
    search = stnls_b200.search.NonLocalSearch(c["ws"], c["wt"], c["ps"], kk, topk_mode=topk_mode, **kw)
    d, i = search(q.to(cuda), k.to(cuda), flows.to(cuda))
    assert_topk_equal(d, i, dr, ir, exact=exact, what="cuda vs reference")


@pytest.mark.parametrize("c", CASES, ids=[c["name"] for c in CASES])
def test_reference_backward(c, cuda, oracle, refc):
    q, k, v, flows, kw = inputs(c)
    flt = c["itype"] == "float"
    search = stnls_b200.search.NonLocalSearch(c["ws"], c["wt"], c["ps"], K, **kw)
    qg, kg = q.to(cuda).requires_grad_(True), k.to(cuda).requires_grad_(True)
    fg = flows.to(cuda).requires_grad_(flt)
    d, i = search(qg, kg, fg)
    gen = th.Generator().manual_seed(63)
    gd = th.randn(d.shape, generator=gen)
    gi = th.randn(i.shape, generator=gen)
    gd = th.where(th.isfinite(d.detach().cpu()), gd, th.zeros_like(gd))
    loss = (d * gd.to(cuda)).sum()
    if flt:
        loss = loss + (i * gi.to(cuda)).sum()
    loss.backward()
    bkw = dict(nheads=HD, stride0=kw["stride0"], dist_type=c["dist_type"], itype=c["itype"], flows_grad=flt)
    r0, r1, rf = refc.search_backward(q.to(cuda), k.to(cuda), flows.to(cuda), gd.to(cuda), gi.to(cuda), i.detach(),
                                      c["wt"], c["ps"], dists=d.detach(), **bkw)
    o0, o1, of = oracle.search_backward(q, k, flows, gd, gi, i.detach().cpu(), c["wt"], c["ps"], **bkw)
    to5 = lambda g: g.permute(0, 2, 1, 3, 4, 5).reshape(B, T, c.get("F", F), H, W)
    for name, ours, ref_, orc in (("grad_q", qg.grad, r0, o0), ("grad_k", kg.grad, r1, o1)):
        assert_close(orc, ref_, 1e-3, 1e-3, "oracle vs reference " + name)
        assert_close(ours, to5(ref_), 1e-3, 1e-3, "cuda vs reference " + name)
    if flt:
        assert_close(of, rf, 2e-3, 2e-3, "oracle vs reference grad_flows")
        assert_close(fg.grad, rf, 2e-3, 2e-3, "cuda vs reference grad_flows")


@pytest.mark.parametrize("itype,anchor", [("int", None), ("int", "anchor"), ("float", None), ("float", "anchor")])
def test_reference_refine(itype, anchor, cuda, oracle, refc):
    q, k, v = make_vids(B, T, F, H, W, seed=71)
    gen = th.Generator().manual_seed(72)
    Ks, wr, Kr = 5, 3, 9
    fl = th.zeros(B, HD, T, H, W, Ks, 3)
    fl[..., 0] = th.randint(-2, 3, fl[..., 0].shape, generator=gen).float()
    fl[..., 1:] = th.randint(-6, 7, fl[..., 1:].shape, generator=gen).float()
    flt = itype == "float"
    if flt:
        fl[..., 1:] += th.rand(fl[..., 1:].shape, generator=gen) / 2 + 0.2
    kw = dict(nheads=HD, stride0=1, stride1=0.5 if flt else 1, dist_type="l2", self_action=anchor, itype=itype)
    dr, ir, ksr, rfr = refc.refine_forward(q.to(cuda), k.to(cuda), fl.to(cuda), 3, 1, wr, Kr, -1, 3, **kw)
    do, io, kso, rfo = oracle.refine_forward(q, k, fl, 3, 1, wr, Kr, -1, 3, **kw)
    assert_topk_equal(do, io, dr, ir, exact=not flt, what="oracle vs reference refine")
    search = stnls_b200.search.RefineSearch(3, 1, wr, Kr, -1, 3, **kw)
    qg, kg = q.to(cuda).requires_grad_(True), k.to(cuda).requires_grad_(True)
    fg = fl.to(cuda).requires_grad_(flt)
    d, i = search(qg, kg, fg)
    assert_topk_equal(d, i, dr, ir, exact=not flt, what="cuda vs reference refine")
    if flt:
        assert th.equal(rfr.cpu().bool(), rfo.bool())
        same = (i.detach() - ir).abs().amax(-1) < 1e-4
        if bool(same.all()):
            gd = th.randn(d.shape, generator=gen)
            gi = th.randn(i.shape, generator=gen)
            gd = th.where(th.isfinite(d.detach().cpu()), gd, th.zeros_like(gd))
            ((d * gd.to(cuda)).sum() + (i * gi.to(cuda)).sum()).backward()
            r0, r1, rf = refc.search_backward(q.to(cuda), k.to(cuda), fl.to(cuda), gd.to(cuda), gi.to(cuda), ir, 1, 3,
                                              nheads=HD, stride0=1, dist_type="l2", itype="float", flows_grad=True,
                                              kselect=ksr, reflect=rfr, refine=True)
            to5 = lambda g: g.permute(0, 2, 1, 3, 4, 5).reshape(B, T, F, H, W)
            assert_close(qg.grad, to5(r0), 1e-3, 1e-3, "cuda vs reference refine grad_q")
            assert_close(kg.grad, to5(r1), 1e-3, 1e-3, "cuda vs reference refine grad_k")
            assert_close(fg.grad, rf, 2e-3, 2e-3, "cuda vs reference refine grad_flows")


@pytest.mark.parametrize("itype", ["int", "float"])
@pytest.mark.parametrize("ps,stride0", [(3, 1), (5, 2), (1, 1)])
def test_reference_gather_add(itype, ps, stride0, cuda, oracle, refc):
    gen = th.Generator().manual_seed(81)
    Kk = 4
    v = th.randn(B, T, F, H, W, generator=gen)
    nH, nW = (H - 1) // stride0 + 1, (W - 1) // stride0 + 1
    w = th.rand(B, HD, T, nH, nW, Kk, generator=gen)
    fl = th.zeros(B, HD, T, nH, nW, Kk, 3)
    fl[..., 0] = th.randint(-2, 3, fl[..., 0].shape, generator=gen).float()
    fl[..., 1:] = th.randint(-5, 6, fl[..., 1:].shape, generator=gen).float()
    flt = itype == "float"
    if flt:
        fl[..., 1:] += th.rand(fl[..., 1:].shape, generator=gen) - 0.5
    outr, cr = refc.gather_add_forward(v.to(cuda), w.to(cuda), fl.to(cuda), ps, stride0, itype=itype)
    outo, co = oracle.gather_add_forward(v, w, fl, ps, stride0, itype=itype)
    assert th.equal(cr.cpu(), co)
    assert_close(outo, outr, 1e-4, 1e-4, "oracle vs reference wpsum")
    agg = stnls_b200.agg.WeightedPatchSum(ps, stride0, itype=itype)
    vg, wg = v.to(cuda).requires_grad_(True), w.to(cuda).requires_grad_(True)
    fg = fl.to(cuda).requires_grad_(flt)
    out = agg(vg, wg, fg)
    assert_close(out, outr, 1e-4, 1e-4, "cuda vs reference wpsum")
    go = th.randn(out.shape, generator=gen)
    (out * go.to(cuda)).sum().backward()
    gv, gw, gf = refc.gather_add_backward(go.to(cuda), v.to(cuda), w.to(cuda), fl.to(cuda), cr, ps, stride0, itype=itype)
    assert_close(vg.grad, gv.permute(0, 2, 1, 3, 4, 5).reshape(B, T, F, H, W), 1e-3, 1e-3, "cuda vs reference grad_v")
    assert_close(wg.grad, gw, 1e-3, 1e-3, "cuda vs reference grad_w")
    if flt:
        assert_close(fg.grad, gf, 2e-3, 2e-3, "cuda vs reference grad_flows")
