"""GPU parity of search_flow: CUDA path vs the CPU oracle and vs the reference's own
kernels (oracle/_ref), plus the caller contract NonLocalSearch(q, k, fflow, bflow) ==
NonLocalSearch(q, k, search_flow(fflow, bflow)) (lib/stnls/search/non_local_search.py:173-175)."""
import pytest
import torch as th

import stnls_b200
from util import assert_close, make_vids

pytestmark = pytest.mark.gpu

CASES = [dict(B=1, T=5, H=16, W=12, wt=2, stride0=1, scale=1.5),
         dict(B=2, T=4, H=11, W=13, wt=1, stride0=2, scale=3.0),
         dict(B=1, T=3, H=9, W=9, wt=2, stride0=1, scale=0.7),      # W_t = min(2wt+1, T) = 3
         dict(B=1, T=7, H=20, W=24, wt=3, stride0=4, scale=6.0)]


def flows_in(c, seed=7):
    g = th.Generator().manual_seed(seed)
    ff = th.randn(c["B"], c["T"], 2, c["H"], c["W"], generator=g) * c["scale"]
    bf = th.randn(c["B"], c["T"], 2, c["H"], c["W"], generator=g) * c["scale"]
    return ff, bf


@pytest.mark.parametrize("c", CASES, ids=[f"T{c['T']}wt{c['wt']}s{c['stride0']}" for c in CASES])
def test_search_flow_vs_oracle(c, cuda, oracle):
    ff, bf = flows_in(c)
    ffg, bfg = ff.to(cuda).requires_grad_(True), bf.to(cuda).requires_grad_(True)
    fl = stnls_b200.nn.search_flow(ffg, bfg, c["wt"], c["stride0"])
    fl_ref = oracle.search_flow_forward(ff, bf, c["wt"], c["stride0"])
    assert_close(fl, fl_ref, 1e-5, 1e-5, "accumulated flows")
    gfl = th.randn(fl_ref.shape, generator=th.Generator().manual_seed(8))
    (fl * gfl.to(cuda)).sum().backward()
    gf, gb = oracle.search_flow_backward(gfl, ff, bf, fl_ref, c["wt"], c["stride0"])
    assert_close(ffg.grad, gf, 1e-3, 1e-3, "grad_fflow")
    assert_close(bfg.grad, gb, 1e-3, 1e-3, "grad_bflow")


@pytest.mark.parametrize("c", CASES[:2], ids=["wt2", "wt1_s2"])
def test_search_flow_vs_reference_kernels(c, cuda, oracle):
    from oracle import ref_cuda
    if not ref_cuda.available() or not hasattr(ref_cuda.mod(), "search_flow_forward"):
        pytest.skip("oracle/_ref without the search_flow kernels")
    ff, bf = flows_in(c, seed=9)
    fl_r = ref_cuda.search_flow_forward(ff.to(cuda), bf.to(cuda), c["wt"], c["stride0"])
    fl_o = oracle.search_flow_forward(ff, bf, c["wt"], c["stride0"])
    assert_close(fl_o, fl_r, 1e-5, 1e-5, "oracle vs reference flows")
    ffg, bfg = ff.to(cuda).requires_grad_(True), bf.to(cuda).requires_grad_(True)
    fl = stnls_b200.nn.search_flow(ffg, bfg, c["wt"], c["stride0"])
    assert_close(fl, fl_r, 1e-5, 1e-5, "cuda vs reference flows")
    gfl = th.randn(fl_o.shape, generator=th.Generator().manual_seed(10))
    gf_r, gb_r = ref_cuda.search_flow_backward(gfl.to(cuda), ff.to(cuda), bf.to(cuda), fl_r, c["wt"], c["stride0"])
    gf_o, gb_o = oracle.search_flow_backward(gfl, ff, bf, fl_o, c["wt"], c["stride0"])
    assert_close(gf_o, gf_r, 1e-3, 1e-3, "oracle vs reference grad_fflow")
    assert_close(gb_o, gb_r, 1e-3, 1e-3, "oracle vs reference grad_bflow")
    (fl * gfl.to(cuda)).sum().backward()
    assert_close(ffg.grad, gf_r, 1e-3, 1e-3, "cuda vs reference grad_fflow")
    assert_close(bfg.grad, gb_r, 1e-3, 1e-3, "cuda vs reference grad_bflow")


def test_wt0_gives_empty_flows(cuda):
    ff = th.zeros(1, 3, 2, 8, 8, device=cuda)
    fl = stnls_b200.nn.search_flow(ff, ff, 0, 2)
    assert tuple(fl.shape) == (1, 3, 0, 2, 4, 4)


def test_search_accepts_fflow_bflow(cuda):
    B, T, F, H, W, HD = 1, 5, 8, 16, 16, 2
    q, k = make_vids(B, T, F, H, W, seed=11, n=2)
    c = dict(B=B, T=T, H=H, W=W, scale=1.0)
    ff, bf = flows_in(c, seed=12)
    search = stnls_b200.search.NonLocalSearch(3, 1, 3, 5, nheads=HD, stride0=1, stride1=1.0, self_action="anchor",
                                              itype="float")
    ffg = ff.to(cuda).requires_grad_(True)
    d4, i4 = search(q.to(cuda), k.to(cuda), ffg, bf.to(cuda))
    fl = stnls_b200.nn.search_flow(ff.to(cuda), bf.to(cuda), 1, 1)
    d3, i3 = search(q.to(cuda), k.to(cuda), fl)
    assert th.equal(d4, d3) and th.equal(i4, i3)
    (d4[th.isfinite(d4)].sum() + i4.sum()).backward()       # gradient reaches the optical flow through both outputs
    assert ffg.grad is not None and bool(th.isfinite(ffg.grad).all()) and float(ffg.grad.abs().sum()) > 0
