"""GPU: the BASELINE.json configurations at their FULL sizes.

C1 / C2 (README shapes) are small enough for the CPU oracle: direct parity (C1 bit-exact).  C3 / C4 are checked
on a frame slice of the full resolution through size-independent properties of the domain, as the reference's
own tests do: distances come back sorted (tests/search/test_non_local_search_bilin2d.py:512-517), every kept
offset points inside the temporal window, and RefineSearch(wr=1) is a fixed point of the search result
(tests/search/test_refinement.py:93-106) -- which re-derives every kept distance with a second kernel path.
"""
import pytest
import torch as th

import stnls_b200
from util import assert_close, assert_topk_equal

pytestmark = pytest.mark.gpu


def readme_inputs(H, W, itype, seed=0):
    g = th.Generator().manual_seed(seed)
    B, T, F = 1, 5, 16
    q, k, v = [th.randn(B, T, F, H, W, generator=g) for _ in range(3)]
    shape = (B, 1, T, 4, 2, H, W)
    if itype == "int":
        flows = th.clamp(10 * th.randn(shape, generator=g), -3, 3).round()
    else:
        flows = th.rand(shape, generator=g) / 2 + th.randint(-3, 3, shape, generator=g) + 0.2
    return q, k, v, flows


def test_c1_full_size_bit_exact(cuda, oracle):
    """configs[0]: B=1,T=5,F=16,64x64, ws=5,wt=2,ps=3,K=10,HD=2, stride1=1, itype=int, search+agg forward."""
    q, k, v, flows = readme_inputs(64, 64, "int")
    kw = dict(nheads=2, stride0=1, stride1=1, self_action="anchor", itype="int")
    d_ref, i_ref = oracle.nls_forward(q, k, flows, 5, 2, 3, 10, **kw)
    d, i = stnls_b200.search.NonLocalSearch(5, 2, 3, 10, **kw)(q.to(cuda), k.to(cuda), flows.to(cuda))
    assert_topk_equal(d, i, d_ref, i_ref, exact=True, what="C1")
    w = th.softmax(-10 * d, -1)
    out = stnls_b200.agg.WeightedPatchSum(3, 1, itype="int")(v.to(cuda), w, i)
    out_ref, _ = oracle.gather_add_forward(v, w.cpu(), i_ref, 3, 1, itype="int")
    assert_close(out, out_ref, 1e-5, 1e-5, "C1 aggregated video")


def test_c2_full_size_forward(cuda, oracle):  # noqa
    """configs[1]: 128x128, stride1=0.5, itype=float (the bench workload): search + agg forward vs the oracle."""
    q, k, v, flows = readme_inputs(128, 128, "float")
    kw = dict(nheads=2, stride0=1, stride1=0.5, self_action="anchor", itype="float")
    d_ref, i_ref = oracle.nls_forward(q, k, flows, 5, 2, 3, 10, **kw)
    d, i = stnls_b200.search.NonLocalSearch(5, 2, 3, 10, **kw)(q.to(cuda), k.to(cuda), flows.to(cuda))
    # 163 840 queries: a handful of near-ties across the K-th slot are expected (the kept candidate's distance still
    # matches); every other slot must agree wherever distances are separated
    assert_topk_equal(d, i, d_ref, i_ref, exact=False, what="C2", last_slot_clear=False)
    nsw = int(((i.cpu() - i_ref).abs().amax(-1) > 1e-3).sum())
    print(f"[C2 vs oracle] {nsw} of {i_ref[..., 0].numel()} kept slots differ (near-tie swaps)")
    assert nsw <= 32, f"C2: {nsw} index swaps against the oracle"
    w = th.softmax(-10 * d_ref, -1)
    out = stnls_b200.agg.WeightedPatchSum(3, 1, itype="float")(v.to(cuda), w.to(cuda), i_ref.to(cuda))
    out_ref, _ = oracle.gather_add_forward(v, w, i_ref, 3, 1, itype="float")
    assert_close(out, out_ref, 1e-4, 1e-4, "C2 aggregated video")


def slice_properties(cuda, T, F, H, W, HD, ws, wt, ps, K, name):
    g = th.Generator(device="cuda").manual_seed(5)
    q = th.randn(1, T, F, H, W, device=cuda, generator=g)
    k = th.randn(1, T, F, H, W, device=cuda, generator=g)
    shape = (1, 1, T, 2 * wt, 2, H, W)
    flows = th.rand(shape, device=cuda, generator=g) / 2 + th.randint(-3, 3, shape, device=cuda, generator=g) + 0.2
    search = stnls_b200.search.NonLocalSearch(ws, wt, ps, K, nheads=HD, stride0=1, stride1=1.0, dist_type="prod",
                                              self_action=None, itype="float")
    d, i = search(q, k, flows)
    assert tuple(d.shape) == (1, HD, T, H, W, K) and tuple(i.shape) == (1, HD, T, H, W, K, 3)
    assert bool(th.isfinite(d).all()), name + ": every query has K valid candidates at this size"
    assert bool((d[..., 1:] <= d[..., :-1]).all()), name + ": dists must be sorted (descending for prod)"
    dt = i[..., 0]
    tq = th.arange(T, device=cuda).view(1, 1, T, 1, 1, 1)
    assert bool(((tq + dt) >= 0).all() and ((tq + dt) <= T - 1).all()) and float(dt.abs().max()) <= 2 * wt
    refine = stnls_b200.search.RefineSearch(ws, wt, 1, -1, -1, ps, nheads=HD, stride0=1, stride1=1.0,
                                            dist_type="prod", itype="float")
    dr, ir = refine(q, k, i)
    # two kernel paths with different fp32 summation orders over ps*ps*F_head = up to 144 products of N(0,1)
    # values: 5e-3 absolute covers the tail of ~1e9 samples (observed max 2.7e-3 at |d| ~ 80)
    assert_close(dr, d, 5e-3, 1e-4, name + ": refine(wr=1) must reproduce the kept distances")
    assert bool(((ir - i).abs().amax(-1) < 1e-4).float().mean() > 0.999)


def test_c3_full_resolution_slice(cuda):
    """configs[2] (RVRT-style): F=64, 256x256, ws=9, wt=1, ps=1, K=9, HD=4 -- a 4-frame slice."""
    slice_properties(cuda, 4, 64, 256, 256, 4, 9, 1, 1, 9, "C3")


def test_c4_full_resolution_slice(cuda):
    """configs[3] (1080p): F=64, 1080x1920, ws=9, wt=3, ps=3, K=20, HD=4 -- the 7 frames one rank holds."""
    free, _ = th.cuda.mem_get_info()
    if free < 60e9:
        pytest.skip("needs ~50 GB of device memory")
    slice_properties(cuda, 7, 64, 1080, 1920, 4, 9, 3, 3, 20, "C4")
    th.cuda.empty_cache()
