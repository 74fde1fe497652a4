"""GPU parity tests: the CUDA path (through the Python front end -> C ABI -> sm_100a kernels) against
the CPU oracle (oracle/ref.py) on the same seeded inputs.

Tolerances: itype="int" forward -- dists bit-exact, indices exactly equal.  itype="float" forward --
dists rtol/atol 1e-4, indices equal wherever distances are clearly separated.  Gradients (atomics,
not bit-reproducible) rtol 1e-3 / atol 1e-3 relative to the oracle's fp32 accumulation.
"""
import pytest
import torch as th

import stnls_b200
from util import assert_close, assert_topk_equal, make_flows, make_vids

pytestmark = pytest.mark.gpu

NLS_CASES = [
    # README C1 / C2 shapes, shrunk spatially
    dict(name="readme_int_l2", B=1, T=5, F=16, H=32, W=32, HD=2, ws=5, wt=2, ps=3, k=10, stride0=1, stride1=1,
         dist_type="l2", self_action="anchor", itype="int", flow="int"),
    dict(name="readme_int_prod", B=1, T=5, F=16, H=32, W=32, HD=2, ws=5, wt=2, ps=3, k=10, stride0=1, stride1=1,
         dist_type="prod", self_action="anchor", itype="int", flow="int"),
    dict(name="readme_float_l2", B=1, T=5, F=16, H=32, W=32, HD=2, ws=5, wt=2, ps=3, k=10, stride0=1,
         stride1=0.5, dist_type="l2", self_action="anchor", itype="float", flow="float"),
    dict(name="readme_float_prod", B=1, T=5, F=16, H=32, W=32, HD=2, ws=5, wt=2, ps=3, k=10, stride0=1,
         stride1=0.5, dist_type="prod", self_action="anchor", itype="float", flow="float"),
    # ragged: odd sizes, channel padding (F_head=3), stride0>1, no reflection, window not re-centred
    dict(name="ragged_int", B=2, T=3, F=3, H=21, W=17, HD=1, ws=3, wt=1, ps=1, k=5, stride0=2, stride1=1,
         dist_type="prod", self_action=None, itype="int", flow="int", reflect_bounds=False, full_ws=False),
    dict(name="ragged_float", B=2, T=4, F=6, H=19, W=23, HD=2, ws=3, wt=1, ps=3, k=-1, stride0=2, stride1=1.1,
         dist_type="l2", self_action=None, itype="float", flow="float"),
    dict(name="keepall_anchor_int", B=1, T=3, F=4, H=16, W=16, HD=1, ws=3, wt=1, ps=3, k=-1, stride0=1, stride1=1,
         dist_type="l2", self_action="anchor", itype="int", flow="int"),
    dict(name="bigK_fallback", B=1, T=3, F=4, H=16, W=16, HD=1, ws=7, wt=0, ps=3, k=40, stride0=1, stride1=1,
         dist_type="prod", self_action="anchor", itype="int", flow="zero"),
    dict(name="dilation_adj", B=1, T=3, F=8, H=20, W=20, HD=2, ws=5, wt=1, ps=3, k=6, stride0=1, stride1=2,
         dist_type="prod", self_action=None, itype="int", flow="int", dilation=2, use_adj=True),
    dict(name="flows_wt_minus_1", B=1, T=5, F=4, H=16, W=16, HD=2, ws=3, wt=2, ps=3, k=4, stride0=1, stride1=1,
         dist_type="l2", self_action="anchor", itype="float", flow="float", nF=4),
    dict(name="rvrt_like", B=1, T=4, F=16, H=24, W=24, HD=4, ws=9, wt=1, ps=1, k=9, stride0=1, stride1=1,
         dist_type="prod", self_action=None, itype="float", flow="float"),
    # raw mode of the forward (float, prod, stride1 == 1): interior windows skip the interpolated lattice
    dict(name="raw_ps1_f16", B=1, T=4, F=64, H=40, W=36, HD=4, ws=9, wt=1, ps=1, k=9, stride0=1, stride1=1,
         dist_type="prod", self_action=None, itype="float", flow="float"),
    dict(name="raw_ps1_f8_anchor", B=1, T=3, F=16, H=30, W=30, HD=2, ws=5, wt=1, ps=1, k=7, stride0=2, stride1=1,
         dist_type="prod", self_action="anchor", itype="float", flow="float"),
    dict(name="raw_ps3_f16", B=1, T=4, F=32, H=40, W=44, HD=2, ws=9, wt=1, ps=3, k=20, stride0=1, stride1=1,
         dist_type="prod", self_action="anchor", itype="float", flow="float"),
    dict(name="raw_ps3_f16_keepall", B=1, T=3, F=16, H=24, W=24, HD=1, ws=3, wt=1, ps=3, k=-1, stride0=1, stride1=1,
         dist_type="prod", self_action=None, itype="float", flow="float"),
    dict(name="big_flows", B=1, T=3, F=4, H=12, W=12, HD=1, ws=5, wt=1, ps=3, k=8, stride0=1, stride1=1,
         dist_type="l2", self_action="anchor", itype="int", flow="int", flow_scale=5),
]


def build_case(c):
    q, k, v = make_vids(c["B"], c["T"], c["F"], c["H"], c["W"], seed=7)
    nH, nW = (c["H"] - 1) // c["stride0"] + 1, (c["W"] - 1) // c["stride0"] + 1
    W_t = 2 * c["wt"] + 1
    flows = make_flows(c["B"], 1, c["T"], c.get("nF", W_t), nH, nW, c["flow"]) * c.get("flow_scale", 1)
    kw = dict(nheads=c["HD"], stride0=c["stride0"], stride1=c["stride1"], dist_type=c["dist_type"],
              dilation=c.get("dilation", 1), self_action=c["self_action"],
              reflect_bounds=c.get("reflect_bounds", True), full_ws=c.get("full_ws", True),
              use_adj=c.get("use_adj", False), itype=c["itype"])
    return q, k, v, flows, kw


@pytest.mark.parametrize("c", NLS_CASES, ids=[c["name"] for c in NLS_CASES])
def test_nls_forward(c, cuda, oracle):
    q, k, v, flows, kw = build_case(c)
    d_ref, i_ref = oracle.nls_forward(q, k, flows, c["ws"], c["wt"], c["ps"], c["k"], **kw)
    search = stnls_b200.search.NonLocalSearch(c["ws"], c["wt"], c["ps"], c["k"], **kw)
    d, i = search(q.to(cuda), k.to(cuda), flows.to(cuda))
    assert_topk_equal(d, i, d_ref, i_ref, exact=(c["itype"] == "int"), what=c["name"])


# the rest of the reference's self_action / topk_mode menus (impl/non_local_search.py:88-139): keep-all kernel +
# per-frame anchoring kernel (csrc/anchor.cu) + stable selections
MENU = [("anchor_each", "all"), ("anchor_each", "each"), ("remove", "all"), ("remove_ref_frame", "all"),
        ("remove_ref_frame", "each"), ("anchor_and_remove_ref_frame", "all"), ("anchor_and_remove_ref_frame", "each"),
        (None, "each"), ("anchor", "each"), (None, "none"), ("anchor_each", "none")]


def menu_case(itype, nF=None):
    c = dict(name="menu", B=1, T=4, F=8, H=18, W=14, HD=2, ws=3, wt=1, ps=3, k=4, stride0=1,
             stride1=1 if itype == "int" else 0.5, dist_type="l2", self_action=None, itype=itype, flow=itype,
             flow_scale=2)
    if nF is not None:
        c["nF"] = nF
    return c


@pytest.mark.parametrize("nF", [3, 2])
@pytest.mark.parametrize("itype", ["int", "float"])
@pytest.mark.parametrize("self_action,topk_mode", MENU, ids=[f"{a}-{m}" for a, m in MENU])
def test_nls_menu_forward(self_action, topk_mode, itype, nF, cuda, oracle):
    c = menu_case(itype, nF)
    q, k, v, flows, kw = build_case(c)
    kw["self_action"] = self_action
    kk = -1 if topk_mode == "none" else c["k"]
    d_ref, i_ref = oracle.nls_forward(q, k, flows, c["ws"], c["wt"], c["ps"], kk, topk_mode=topk_mode, **kw)
    search = stnls_b200.search.NonLocalSearch(c["ws"], c["wt"], c["ps"], kk, topk_mode=topk_mode, **kw)
    d, i = search(q.to(cuda), k.to(cuda), flows.to(cuda))
    assert_topk_equal(d, i, d_ref, i_ref, exact=(itype == "int"), what=f"{self_action}/{topk_mode}/{itype}")


@pytest.mark.parametrize("itype", ["int", "float"])
@pytest.mark.parametrize("self_action,topk_mode", [("anchor_each", "each"), ("remove_ref_frame", "all"), ("remove", "all")])
def test_nls_menu_backward(self_action, topk_mode, itype, cuda, oracle):
    c = menu_case(itype)
    q, k, v, flows, kw = build_case(c)
    kw["self_action"] = self_action
    search = stnls_b200.search.NonLocalSearch(c["ws"], c["wt"], c["ps"], c["k"], topk_mode=topk_mode, **kw)
    qg, kg = q.to(cuda).requires_grad_(True), k.to(cuda).requires_grad_(True)
    d, i = search(qg, kg, flows.to(cuda))
    gd = th.randn(d.shape, generator=th.Generator().manual_seed(12))
    gd = th.where(th.isfinite(d.detach().cpu()), gd, th.zeros_like(gd))
    (d * gd.to(cuda)).sum().backward()
    g0, g1, _ = oracle.search_backward(q, k, flows, gd, th.zeros(i.shape), i.detach().cpu(), c["wt"], c["ps"],
                                       nheads=c["HD"], stride0=c["stride0"], dist_type=c["dist_type"], itype=itype)
    B, T, F, H, W = q.shape
    to5 = lambda g: g.permute(0, 2, 1, 3, 4, 5).reshape(B, T, F, H, W)
    assert_close(qg.grad, to5(g0), 1e-3, 1e-3, "menu grad_q")
    assert_close(kg.grad, to5(g1), 1e-3, 1e-3, "menu grad_k")


BWD_CASES = [c for c in NLS_CASES if c["name"] in
             ("readme_int_l2", "readme_int_prod", "readme_float_l2", "readme_float_prod", "ragged_int",
              "ragged_float", "dilation_adj", "flows_wt_minus_1", "rvrt_like", "raw_ps1_f16", "raw_ps3_f16")]


@pytest.mark.parametrize("flows_grad", [False, True])
@pytest.mark.parametrize("c", BWD_CASES, ids=[c["name"] for c in BWD_CASES])
def test_nls_backward(c, flows_grad, cuda, oracle):
    if flows_grad and c["itype"] == "int":
        pytest.skip("int flows are not differentiable")
    q, k, v, flows, kw = build_case(c)
    search = stnls_b200.search.NonLocalSearch(c["ws"], c["wt"], c["ps"], c["k"], **kw)
    qg, kg = q.to(cuda).requires_grad_(True), k.to(cuda).requires_grad_(True)
    fg = flows.to(cuda).requires_grad_(flows_grad)
    d, i = search(qg, kg, fg)
    gen = th.Generator().manual_seed(11)
    gd = th.randn(d.shape, generator=gen)
    gi = th.randn(i.shape, generator=gen)
    gd = th.where(th.isfinite(d.detach().cpu()), gd, th.zeros_like(gd))
    if c["itype"] == "float":
        loss = (d * gd.to(cuda)).sum() + (i * gi.to(cuda)).sum()
    else:
        loss = (d * gd.to(cuda)).sum()
    loss.backward()
    # oracle backward at OUR indices (isolates the backward kernels from forward tie-breaks)
    g0, g1, gf = oracle.search_backward(q, k, flows, gd, gi, i.detach().cpu(), c["wt"], c["ps"], nheads=c["HD"],
                                        stride0=c["stride0"], dist_type=c["dist_type"],
                                        dilation=c.get("dilation", 1),
                                        reflect_bounds=c.get("reflect_bounds", True),
                                        use_adj=c.get("use_adj", False), itype=c["itype"], flows_grad=flows_grad)
    B, T, F, H, W = q.shape
    to5 = lambda g: g.permute(0, 2, 1, 3, 4, 5).reshape(B, T, F, H, W)
    assert_close(qg.grad, to5(g0), 1e-3, 1e-3, c["name"] + " grad_q")
    assert_close(kg.grad, to5(g1), 1e-3, 1e-3, c["name"] + " grad_k")
    if flows_grad:
        assert_close(fg.grad, gf, 2e-3, 2e-3, c["name"] + " grad_flows")


REFINE_CASES = [
    dict(name="refine_int", itype="int", wr=3, k=9, Ks=5, self_action=None, dist_type="l2", stride0=1, stride1=1),
    dict(name="refine_int_anchor", itype="int", wr=3, k=7, Ks=4, self_action="anchor", dist_type="prod", stride0=2,
         stride1=1),
    dict(name="refine_float", itype="float", wr=3, k=9, Ks=5, self_action=None, dist_type="l2", stride0=1,
         stride1=0.5),
    dict(name="refine_float_anchor", itype="float", wr=3, k=10, Ks=6, self_action="anchor", dist_type="prod",
         stride0=2, stride1=1.0),
    dict(name="refine_keepall", itype="float", wr=2, k=-1, Ks=3, self_action="anchor", dist_type="l2", stride0=1,
         stride1=1.0),
    dict(name="refine_manyK", itype="int", wr=1, k=-1, Ks=40, self_action=None, dist_type="l2", stride0=2, stride1=1),
    # topk_mode="each": top-k inside every group of wr*wr candidates (impl/refinement.py:99-106)
    dict(name="refine_each_int", itype="int", wr=3, k=3, Ks=4, self_action=None, dist_type="l2", stride0=1, stride1=1,
         topk_mode="each"),
    dict(name="refine_each_float_anchor", itype="float", wr=3, k=2, Ks=5, self_action="anchor_each", dist_type="prod",
         stride0=2, stride1=0.5, topk_mode="each"),
]


def refine_inputs(c, oracle):
    B, T, F, H, W, HD = 1, 4, 8, 20, 18, 2
    q, k, v = make_vids(B, T, F, H, W, seed=21)
    nH, nW = (H - 1) // c["stride0"] + 1, (W - 1) // c["stride0"] + 1
    gen = th.Generator().manual_seed(22)
    Ks = c["Ks"]
    fl = th.zeros(B, HD, T, nH, nW, Ks, 3)
    fl[..., 0] = th.randint(-2, 3, fl[..., 0].shape, generator=gen).float()
    fl[..., 1:] = th.randint(-6, 7, fl[..., 1:].shape, generator=gen).float()
    if c["itype"] == "float":
        fl[..., 1:] += th.rand(fl[..., 1:].shape, generator=gen) / 2 + 0.2
    return q, k, v, fl, HD


@pytest.mark.parametrize("c", REFINE_CASES, ids=[c["name"] for c in REFINE_CASES])
def test_refine_forward(c, cuda, oracle):
    q, k, v, fl, HD = refine_inputs(c, oracle)
    kw = dict(nheads=HD, stride0=c["stride0"], stride1=c["stride1"], dist_type=c["dist_type"],
              self_action=c["self_action"], itype=c["itype"], topk_mode=c.get("topk_mode", "all"))
    d_ref, i_ref, ks_ref, rf_ref = oracle.refine_forward(q, k, fl, 3, 1, c["wr"], c["k"], -1, 3, **kw)
    search = stnls_b200.search.RefineSearch(3, 1, c["wr"], c["k"], -1, 3, **kw)
    d, i = search(q.to(cuda), k.to(cuda), fl.to(cuda))
    assert_topk_equal(d, i, d_ref, i_ref, exact=(c["itype"] == "int"), what=c["name"])


@pytest.mark.parametrize("c", [c for c in REFINE_CASES if c["name"] != "refine_manyK"],
                         ids=[c["name"] for c in REFINE_CASES if c["name"] != "refine_manyK"])
def test_refine_backward(c, cuda, oracle):
    q, k, v, fl, HD = refine_inputs(c, oracle)
    kw = dict(nheads=HD, stride0=c["stride0"], stride1=c["stride1"], dist_type=c["dist_type"],
              self_action=c["self_action"], itype=c["itype"], topk_mode=c.get("topk_mode", "all"))
    flt = c["itype"] == "float"
    # oracle forward provides kselect / reflect for the oracle backward
    d_ref, i_ref, ks_ref, rf_ref = oracle.refine_forward(q, k, fl, 3, 1, c["wr"], c["k"], -1, 3, **kw)
    search = stnls_b200.search.RefineSearch(3, 1, c["wr"], c["k"], -1, 3, **kw)
    qg, kg = q.to(cuda).requires_grad_(True), k.to(cuda).requires_grad_(True)
    fg = fl.to(cuda).requires_grad_(flt)
    d, i = search(qg, kg, fg)
    gen = th.Generator().manual_seed(23)
    gd = th.randn(d.shape, generator=gen)
    gi = th.randn(i.shape, generator=gen)
    gd = th.where(th.isfinite(d.detach().cpu()), gd, th.zeros_like(gd))
    loss = (d * gd.to(cuda)).sum()
    if flt:
        loss = loss + (i * gi.to(cuda)).sum()
    loss.backward()
    if flt:
        # only compare when both forwards picked the same candidates (no near-tie swaps)
        same = (i.detach().cpu() - i_ref).abs().amax(-1) < 1e-4
        assert float(same.float().mean()) > 0.99
        if not bool(same.all()):
            pytest.skip("near-tie swap between CUDA and oracle forward; backward compared in NLS tests")
    g0, g1, gf = oracle.search_backward(q, k, fl, gd, gi, i_ref, 1, 3, nheads=HD, stride0=c["stride0"],
                                        dist_type=c["dist_type"], itype=c["itype"], flows_grad=flt,
                                        kselect=ks_ref, reflect=rf_ref, refine=True)
    B, T, F, H, W = q.shape
    to5 = lambda g: g.permute(0, 2, 1, 3, 4, 5).reshape(B, T, F, H, W)
    assert_close(qg.grad, to5(g0), 1e-3, 1e-3, c["name"] + " grad_q")
    assert_close(kg.grad, to5(g1), 1e-3, 1e-3, c["name"] + " grad_k")
    if flt:
        assert_close(fg.grad, gf, 2e-3, 2e-3, c["name"] + " grad_flows")


def test_refine_is_fixed_point_of_search(cuda):
    """tests/search/test_refinement.py:93-106 on the CUDA path."""
    q, k, _ = make_vids(1, 4, 8, 24, 24, seed=31)
    flows = make_flows(1, 1, 4, 3, 24, 24, "float")
    for itype in ("int", "float"):
        s = stnls_b200.search.NonLocalSearch(5, 1, 3, 7, nheads=2, stride1=1, itype=itype)
        r = stnls_b200.search.RefineSearch(5, 1, 1, -1, -1, 3, nheads=2, stride0=1, stride1=1, itype=itype)
        d, i = s(q.to(cuda), k.to(cuda), flows.to(cuda))
        dr, ir = r(q.to(cuda), k.to(cuda), i)
        assert th.allclose(d, dr, atol=1e-4, rtol=1e-4)
        assert th.allclose(i.float(), ir.float(), atol=1e-5)


AGG_CASES = [
    dict(name="agg_int_ps3", itype="int", ps=3, stride0=1, K=5, F=6, HD=2),
    dict(name="agg_float_ps3", itype="float", ps=3, stride0=1, K=5, F=6, HD=2),
    dict(name="agg_int_ps1", itype="int", ps=1, stride0=1, K=10, F=3, HD=1),
    dict(name="agg_float_ps5_s2", itype="float", ps=5, stride0=2, K=4, F=8, HD=2),
    dict(name="agg_float_ps5_s3", itype="float", ps=5, stride0=3, K=5, F=3, HD=1),
    dict(name="agg_int_noreflect", itype="int", ps=3, stride0=2, K=3, F=4, HD=1, reflect_bounds=False),
    dict(name="agg_float_dil2", itype="float", ps=3, stride0=1, K=3, F=4, HD=1, dilation=2),
]


def agg_inputs(c):
    B, T, H, W = 1, 4, 16, 16
    gen = th.Generator().manual_seed(41)
    v = th.randn(B, T, c["F"], H, W, generator=gen)
    nH, nW = (H - 1) // c["stride0"] + 1, (W - 1) // c["stride0"] + 1
    w = th.rand(B, c["HD"], T, nH, nW, c["K"], generator=gen)
    fl = th.zeros(B, c["HD"], T, nH, nW, c["K"], 3)
    fl[..., 0] = th.randint(-2, 3, fl[..., 0].shape, generator=gen).float()
    fl[..., 1:] = th.randint(-5, 6, fl[..., 1:].shape, generator=gen).float()
    if c["itype"] == "float":
        fl[..., 1:] += th.rand(fl[..., 1:].shape, generator=gen) - 0.5
    return v, w, fl


@pytest.mark.parametrize("c", AGG_CASES, ids=[c["name"] for c in AGG_CASES])
def test_wpsum_forward_backward(c, cuda, oracle):
    v, w, fl = agg_inputs(c)
    kw = dict(dilation=c.get("dilation", 1), reflect_bounds=c.get("reflect_bounds", True), itype=c["itype"])
    out_ref, counts = oracle.gather_add_forward(v, w, fl, c["ps"], c["stride0"], **kw)
    agg = stnls_b200.agg.WeightedPatchSum(c["ps"], c["stride0"], **kw)
    flt = c["itype"] == "float"
    vg, wg = v.to(cuda).requires_grad_(True), w.to(cuda).requires_grad_(True)
    fg = fl.to(cuda).requires_grad_(flt)
    out = agg(vg, wg, fg)
    assert_close(out, out_ref, 1e-4, 1e-4, c["name"] + " out")
    go = th.randn(out_ref.shape, generator=th.Generator().manual_seed(42))
    (out * go.to(cuda)).sum().backward()
    gv, gw, gf = oracle.gather_add_backward(go, v, w, fl, counts, c["ps"], c["stride0"], **kw)
    B, T, F, H, W = v.shape
    assert_close(vg.grad, gv.permute(0, 2, 1, 3, 4, 5).reshape(B, T, F, H, W), 1e-3, 1e-3, c["name"] + " grad_v")
    assert_close(wg.grad, gw, 1e-3, 1e-3, c["name"] + " grad_w")
    if flt:
        assert_close(fg.grad[..., 1:], gf[..., 1:], 2e-3, 2e-3, c["name"] + " grad_flows")
        assert float(fg.grad[..., 0].abs().max()) == 0.0


def test_readme_pipeline_backward(cuda, oracle):
    """README.md:25-57 end to end (search -> softmax -> WeightedPatchSum) forward + backward, against the
    oracle chained by hand (softmax through torch autograd on the CPU)."""
    B, T, F, H, W, HD = 1, 5, 16, 24, 24, 2
    ws, wt, ps, K, s0, s1 = 5, 2, 3, 10, 1, 0.5
    q, k, v = make_vids(B, T, F, H, W, seed=51)
    flows = make_flows(B, 1, T, 2 * wt, H, W, "float", seed=52)
    search = stnls_b200.search.NonLocalSearch(ws, wt, ps, K, nheads=HD, stride0=s0, stride1=s1,
                                              self_action="anchor", itype="float")
    agg = stnls_b200.agg.WeightedPatchSum(ps=ps, stride0=s0, itype="float")
    qg, kg, vg = [x.to(cuda).requires_grad_(True) for x in (q, k, v)]
    fg = flows.to(cuda).requires_grad_(True)
    dists, inds = search(qg, kg, fg)
    weights = th.softmax(-dists, -1)
    out = agg(vg, weights, inds)
    go = th.randn(out.shape, generator=th.Generator().manual_seed(53))
    (out * go.to(cuda)).sum().backward()

    d_ref, i_ref = oracle.nls_forward(q, k, flows, ws, wt, ps, K, nheads=HD, stride0=s0, stride1=s1,
                                      self_action="anchor", itype="float")
    assert_topk_equal(dists, inds, d_ref, i_ref, exact=False, what="readme fwd")
    i_use = inds.detach().cpu()
    d_leaf = dists.detach().cpu().requires_grad_(True)
    w_ref = th.softmax(-d_leaf, -1)
    out_ref, counts = oracle.gather_add_forward(v, w_ref.detach(), i_use, ps, s0, itype="float")
    assert_close(out, out_ref, 2e-4, 2e-4, "readme out")
    gv, gw, gf = oracle.gather_add_backward(go, v, w_ref.detach(), i_use, counts, ps, s0, itype="float")
    w_ref.backward(gw)
    g0, g1, gfl = oracle.search_backward(q, k, flows, d_leaf.grad, gf, i_use, wt, ps, nheads=HD, stride0=s0,
                                         itype="float", flows_grad=True)
    to5 = lambda g: g.permute(0, 2, 1, 3, 4, 5).reshape(B, T, F, H, W)
    assert_close(vg.grad, to5(gv), 1e-3, 1e-3, "readme grad_v")
    assert_close(qg.grad, to5(g0), 2e-3, 2e-3, "readme grad_q")
    assert_close(kg.grad, to5(g1), 2e-3, 2e-3, "readme grad_k")
    assert_close(fg.grad, gfl, 5e-3, 5e-3, "readme grad_flows")


def test_layout_roundtrip_and_errors(cuda):
    from stnls_b200 import ops
    x = th.randn(2, 3, 12, 9, 11, device=cuda)
    cl = ops.pack_cl(x, 4)
    assert cl.shape == (8, 3, 9, 11, 4)
    y = ops.unpack_cl(cl, x.shape, 4)
    assert th.equal(x, y)
    x6 = th.randn(2, 2, 3, 5, 7, 6, device=cuda)
    assert th.equal(ops.unpack_cl(ops.pack_cl(x6, 2), x6.shape, 2), x6)
    with pytest.raises(RuntimeError):
        ops.pack_cl(th.randn(1, 2, 4, 8, 8), 1)  # CPU tensor: no fallback
    with pytest.raises(NotImplementedError):
        stnls_b200.search.NonLocalSearch(3, 0, 1, 1, pt=2)(x, x)
    with pytest.raises(AssertionError):
        stnls_b200.search.NonLocalSearch(3, 2, 1, 1, nheads=4)(x, x)  # T=3 < W_t=5


@pytest.mark.parametrize("dist_type", ["l2", "prod"])
@pytest.mark.parametrize("stride1", [1.0, 0.5])
def test_fwd_anchor_properties(stride1, dist_type, cuda):
    """The reference's own anchoring test (tests/search/test_non_local_search_bilin2d.py:519-667, test_fwd_anchor):
    with topk_mode="each", slot 0 of every frame group of "anchor_each" is the candidate the flow points at --
    the window centre of an un-recentred (full_ws=False) keep-all search -- whatever k and full_ws are; group 0
    agrees with "anchor"; "anchor_and_remove_ref_frame" gives the same anchors for the non-reference frames."""
    ws, wt, ps, stride0, HD = 3, 1, 3, 1, 1
    B, T, F, H, W = 1, 5, 8, 32, 32
    W_t = 2 * wt + 1
    gen = th.Generator().manual_seed(123)
    vid0 = (th.rand((B, T, F, H, W), generator=gen) - 0.5).to(cuda)
    vid1 = (th.rand((B, T, F, H, W), generator=gen) - 0.5).to(cuda)
    nH, nW = (H - 1) // stride0 + 1, (W - 1) // stride0 + 1
    flows = (2 * th.rand((B, T, W_t - 1, 2, nH, nW), generator=gen)).to(cuda)

    def run(k, full_ws, self_action, topk_mode="each"):
        s = stnls_b200.search.NonLocalSearch(ws, wt, ps, k, HD, stride0=stride0, stride1=stride1, full_ws=full_ws,
                                             self_action=self_action, dist_type=dist_type, topk_mode=topk_mode,
                                             itype="float")
        return s(vid0, vid1, flows)

    vshape = (B, HD, T, nH, nW, W_t)
    d0, i0 = run(-1, False, None)
    d0 = d0.view(vshape + (ws * ws,))[..., ws // 2 * ws + ws // 2]
    i0 = i0.view(vshape + (ws * ws, 3))[..., ws // 2 * ws + ws // 2, :]
    firsts = [(d0, i0)]
    for k, full_ws in ((3, False), (5, True)):
        d, i = run(k, full_ws, "anchor_each")
        firsts.append((d.view(vshape + (k,))[..., 0], i.view(vshape + (k, 3))[..., 0, :]))
    for a in range(3):
        for b in range(a + 1, 3):
            assert th.allclose(firsts[a][0], firsts[b][0], 1e-3, 1e-3, equal_nan=True)
            assert th.allclose(firsts[a][1], firsts[b][1], 1e-3, 1e-3, equal_nan=True)
    d3, i3 = run(8, True, "anchor")
    d3, i3 = d3.view(vshape + (8,))[..., 0], i3.view(vshape + (8, 3))[..., 0, :]
    for d, i in firsts:
        assert th.allclose(d[..., 0], d3[..., 0], 1e-3, 1e-3, equal_nan=True)
        assert th.allclose(i[..., 0, :], i3[..., 0, :], 1e-3, 1e-3, equal_nan=True)
    d4, i4 = run(8, True, "anchor_and_remove_ref_frame")
    d4 = d4.view((B, HD, T, nH, nW, W_t - 1, 8))[..., 0]
    i4 = i4.view((B, HD, T, nH, nW, W_t - 1, 8, 3))[..., 0, :]
    # spatial part of the anchors of the non-reference frames: reflect(ref + flow) - ref
    assert th.allclose(i4[..., 1:], firsts[2][1][..., 1:, 1:], 1e-3, 1e-3)
    assert th.allclose(d4, firsts[2][0][..., 1:], 1e-3, 1e-3)
    # anchors equal the reflected flow target
    grid_h = th.arange(nH, device=cuda).view(1, 1, nH, 1) * stride0
    grid_w = th.arange(nW, device=cuda).view(1, 1, 1, nW) * stride0
    for si in range(1, W_t):
        fh, fw = flows[:, :, si - 1, 1] + grid_h, flows[:, :, si - 1, 0] + grid_w
        fh = th.where(fh > H - 1, 2 * (H - 1) - fh, fh).abs()
        fw = th.where(fw > W - 1, 2 * (W - 1) - fw, fw).abs()
        got = firsts[2][1][:, 0, :, :, :, si]
        assert th.allclose(got[..., 1], fh - grid_h, 1e-3, 1e-3)
        assert th.allclose(got[..., 2], fw - grid_w, 1e-3, 1e-3)


@pytest.mark.parametrize("itype", ["int", "float"])
@pytest.mark.parametrize("dist_type", ["l2", "prod"])
def test_refine_anchor_each_reproduces_search(itype, dist_type, cuda):
    """tests/search/test_refinement.py:225-305 (test_anchor_fwd) on the CUDA path: refining the per-frame top-k of
    a search with k=1, topk_mode="each", self_action="anchor_each" gives the search's own results back, with and
    without window re-centring."""
    ws, wt, wr, ps, stride0, HD = 3, 1, 1, 3, 1, 2
    stride1 = 1 if itype == "int" else 0.5
    B, T, F, H, W = 1, 5, 8, 32, 32
    W_t = 2 * wt + 1
    gen = th.Generator().manual_seed(321)
    vid0 = (th.rand((B, T, F, H, W), generator=gen) - 0.5).to(cuda)
    vid1 = (th.rand((B, T, F, H, W), generator=gen) - 0.5).to(cuda)
    flows = (2 * th.rand((B, T, W_t - 1, 2, H, W), generator=gen)).to(cuda)
    k0 = 5
    search0 = stnls_b200.search.NonLocalSearch(ws, wt, ps, k0, HD, stride0=stride0, stride1=stride1, full_ws=True,
                                               self_action=None, dist_type=dist_type, topk_mode="each", itype=itype)
    d0, i0 = search0(vid0, vid1, flows)
    assert d0.shape[-1] == W_t * k0
    for full_ws in (False, True):
        refine = stnls_b200.search.RefineSearch(ws, wt, wr, 1, -1, ps, HD, stride0=stride0, stride1=stride1,
                                                full_ws=full_ws, self_action="anchor_each", dist_type=dist_type,
                                                topk_mode="each", itype=itype)
        dr, ir = refine(vid0, vid1, i0)
        assert dr.shape == d0.shape
        assert th.allclose(d0, dr, 1e-2, 1e-3, equal_nan=True)
        assert th.allclose(i0.float(), ir.float(), 1e-3, 1e-3, equal_nan=True)


@pytest.mark.parametrize("stride0,stride1", [(1, 1), (2, 1), (2, 2)])
def test_normalize_bwd(stride0, stride1, cuda, oracle):
    """normalize_bwd=True divides the video gradients by the patch-overlap counts (lib/stnls/search/shared.py:59-97)."""
    c = dict(name="normz", B=1, T=3, F=4, H=17, W=20, HD=2, ws=3, wt=1, ps=3, k=4, stride0=stride0, stride1=stride1,
             dist_type="l2", self_action=None, itype="int", flow="int")
    q, k, v, flows, kw = build_case(c)
    grads = []
    for normalize in (False, True):
        search = stnls_b200.search.NonLocalSearch(c["ws"], c["wt"], c["ps"], c["k"], normalize_bwd=normalize, **kw)
        qg, kg = q.to(cuda).requires_grad_(True), k.to(cuda).requires_grad_(True)
        d, i = search(qg, kg, flows.to(cuda))
        gd = th.randn(d.shape, generator=th.Generator().manual_seed(5))
        gd = th.where(th.isfinite(d.detach().cpu()), gd, th.zeros_like(gd))
        (d * gd.to(cuda)).sum().backward()
        grads.append((qg.grad.cpu(), kg.grad.cpu()))
    e0, e1 = oracle.normz_bwd(grads[0][0].clone(), grads[0][1].clone(), c["ps"], stride0, stride1)
    # the two backward passes are separate launches accumulating with float atomics: their sums differ by rounding
    # in the order of arrival (a few 1e-6 at gradients of magnitude ~40), hence atol 1e-5
    assert_close(grads[1][0], e0, 1e-5, 1e-5, "normalize_bwd grad_q")
    assert_close(grads[1][1], e1, 1e-5, 1e-5, "normalize_bwd grad_k")
    assert not th.allclose(grads[1][0], grads[0][0])


@pytest.mark.parametrize("dist_type", ["l2", "prod"])
def test_fwd_float_vs_int(dist_type, cuda):
    """tests/search/test_non_local_search_bilin2d.py:58-150 (test_fwd_vs_int) on the CUDA path: with integer stride1
    and integer optical flow the bilinear search equals the integer search, called the way the reference test calls
    it -- forward(q, k, fflow, bflow), flows accumulated by nn.search_flow inside the module."""
    ws, wt, ps, k, HD, stride0, stride1 = 3, 1, 3, -1, 2, 1, 1
    B, T, F, H, W = 1, 3, 1, 10, 10
    gen = th.Generator().manual_seed(0)
    vid = th.rand((B, T, HD * F, H, W), generator=gen).to(cuda)
    fflow = th.clamp(10 * th.randn((B, T, 2, H, W), generator=gen), -10, 10).round().to(cuda)
    bflow = th.clamp(10 * th.randn((B, T, 2, H, W), generator=gen), -10, 10).round().to(cuda)
    out = {}
    for itype in ("float", "int"):
        search = stnls_b200.search.NonLocalSearch(ws, wt, ps, k, HD, dist_type=dist_type, stride0=stride0,
                                                  stride1=stride1, reflect_bounds=True, full_ws=True,
                                                  self_action=None, itype=itype)
        out[itype] = search(vid, vid, fflow, bflow)
    d_te, i_te = out["float"]
    d_gt, i_gt = out["int"]
    ok = d_gt.abs() < 1e30
    assert bool((ok == (d_te.abs() < 1e30)).all())
    diff = ((d_te - d_gt).abs() / (d_gt.abs() + 1e-5))[ok]
    assert float(diff.mean()) < 5e-3 and float(diff.max()) < 1e-4 + 5e-3
    assert bool(((i_te - i_gt.float()).abs()[ok] < 1e-3).all())
