"""CPU: pin the oracle on the literal known-answer vectors the reference holds for this path
(SURVEY.md 8c): dev/trange.py:33-41 (temporal window order) and dev/ws_offset.py:37-41 (search
window offsets), plus the reference's property tests restated on the oracle itself."""
import ctypes

import pytest
import torch as th

from util import make_flows, make_vids

TRANGE = [((0, 2, 5), [0, 1, 2, 3, 4]), ((1, 2, 5), [1, 2, 3, 4, 0]), ((2, 2, 5), [2, 3, 4, 1, 0]),
          ((2, 1, 3), [2, 1, 0]), ((1, 1, 3), [1, 2, 0]), ((3, 2, 7), [3, 4, 5, 2, 1]),
          ((3, 3, 7), [3, 4, 5, 6, 2, 1, 0])]


@pytest.mark.parametrize("args,gt", TRANGE)
def test_time_window_kat(oracle, args, gt):
    ti, wt, T = args
    out = (ctypes.c_int * (2 * wt + 1))()
    oracle.lib().oracle_time_window(ti, wt, T, out)
    assert list(out) == gt


WSOFF = [((1.1, 0.8), (1, 0)), ((0.8, 0.8), (0, 0)), ((0.2, 0.2), (0, 0))]


@pytest.mark.parametrize("args,gt", WSOFF)
def test_search_offset_kat(oracle, args, gt):
    out = (ctypes.c_float * 2)()
    oracle.lib().oracle_search_offset_f(ctypes.c_float(args[0]), ctypes.c_float(args[1]), ctypes.c_float(1.0),
                                        3, 10, 10, out)
    assert (out[0], out[1]) == gt


def test_float_equals_int_at_integer_flows(oracle):
    """tests/search/test_non_local_search_bilin2d.py:58-142 restated on the oracle."""
    q, k = make_vids(1, 3, 4, 16, 16, n=2)
    flows = make_flows(1, 1, 3, 3, 16, 16, "int")
    for dt in ("prod", "l2"):
        di, ii = oracle.nls_forward(q, k, flows, 3, 1, 3, -1, nheads=2, dist_type=dt, itype="int")
        df, if_ = oracle.nls_forward(q, k, flows, 3, 1, 3, -1, nheads=2, dist_type=dt, itype="float")
        assert th.equal(ii.float(), if_)
        assert th.allclose(di, df, atol=1e-5, rtol=1e-5)


def test_refine_fixed_point(oracle):
    """tests/search/test_refinement.py:93-106: RefineSearch(wr=1,k=-1) on the inds of NonLocalSearch
    reproduces its dists."""
    q, k = make_vids(1, 3, 4, 16, 16, n=2)
    flows = make_flows(1, 1, 3, 3, 16, 16, "float")
    for itype in ("int", "float"):
        d, i = oracle.nls_forward(q, k, flows, 3, 1, 3, 5, nheads=2, stride1=1, itype=itype)
        dr, ir, _, _ = oracle.refine_forward(q, k, i, 3, 1, 1, -1, -1, 3, nheads=2, stride0=1, stride1=1, itype=itype)
        assert th.allclose(d, dr, atol=1e-4, rtol=1e-4)
        assert th.allclose(i.float(), ir.float(), atol=1e-5)


def test_topk_monotone_and_anchor(oracle):
    """test_non_local_search_bilin2d.py:464-517 (monotone dists) and :646-672 (anchored slot 0)."""
    q, k = make_vids(1, 3, 4, 16, 16, n=2)
    flows = make_flows(1, 1, 3, 3, 16, 16, "float")
    d, i = oracle.nls_forward(q, k, flows, 5, 1, 3, 7, nheads=1, stride1=0.5, dist_type="l2",
                              self_action="anchor", itype="float")
    assert bool((d[..., 2:] >= d[..., 1:-1]).all())
    # slot 0 holds the candidate nearest to the query itself; its offset is reported as exactly zero
    # unless that candidate already sat in slot 0 (anchor_self_kernel.cu:101-113 swaps with itself)
    l1 = i.abs().sum(-1)
    assert bool((l1[..., :1] <= l1 + 1e-6).all())
    assert float((l1[..., 0] == 0).float().mean()) > 0.9
    d, i = oracle.nls_forward(q, k, flows, 5, 1, 3, 7, nheads=1, dist_type="prod", itype="int")
    assert bool((d[..., 1:] <= d[..., :-1]).all())


def test_gather_add_matches_python_loops(oracle):
    """tests/agg/test_gather_add.py:106-128 in spirit: the weighted patch sum equals a literal
    python restatement of the definition on a tiny case (int offsets, interior + border pixels)."""
    th.manual_seed(0)
    B, HD, T, F, H, W, K, ps = 1, 1, 2, 2, 6, 6, 2, 3
    v = th.randn(B, HD, T, F, H, W)
    w = th.rand(B, HD, T, H, W, K)
    fl = th.randint(-2, 3, (B, HD, T, H, W, K, 3)).int()
    fl[..., 0] = th.randint(-1, 2, fl[..., 0].shape)
    out, counts = oracle.gather_add_forward(v, w, fl, ps, 1, itype="int")
    refl = lambda x, n: -x if x < 0 else (2 * (n - 1) - x if x > n - 1 else x)
    exp = th.zeros_like(v)
    cnt = th.zeros(H, W)
    for t in range(T):
        for h in range(H):
            for ww in range(W):
                for pi in range(ps):
                    for pj in range(ps):
                        rh, rw = h + pi - 1, ww + pj - 1
                        if not (0 <= rh < H and 0 <= rw < W):
                            continue
                        if t == 0:
                            cnt[rh, rw] += 1
                        for kk in range(K):
                            n0 = refl(t + int(fl[0, 0, t, h, ww, kk, 0]), T)
                            n1 = refl(refl(h + int(fl[0, 0, t, h, ww, kk, 1]), H) + pi - 1, H)
                            n2 = refl(refl(ww + int(fl[0, 0, t, h, ww, kk, 2]), W) + pj - 1, W)
                            if not (0 <= n0 < T and 0 <= n1 < H and 0 <= n2 < W):
                                continue
                            exp[0, 0, t, :, rh, rw] += w[0, 0, t, h, ww, kk] * v[0, 0, n0, :, n1, n2]
    assert th.equal(cnt.int(), counts)
    assert th.allclose(out, exp / cnt, atol=1e-5, rtol=1e-5)


def test_oracle_grads_match_finite_differences(oracle):
    """gradcheck in the spirit of test_non_local_search_bilin2d.py:309-379 (eps 1e-2, atol 1e-2) for the
    prod distance wrt q and k at fixed indices (float path)."""
    q, k = make_vids(1, 3, 2, 8, 8, n=2, seed=3)
    flows = make_flows(1, 1, 3, 3, 8, 8, "float", seed=4)
    kw = dict(nheads=1, stride1=1.0, dist_type="prod", itype="float")
    d, i = oracle.nls_forward(q, k, flows, 3, 1, 3, 4, **kw)
    gd = th.randn(d.shape, generator=th.Generator().manual_seed(5))
    g0, g1, _ = oracle.search_backward(q, k, flows, gd, None, i, 1, 3, nheads=1, dist_type="prod", itype="float")

    def loss(qq, kk):
        dd, _, _, _ = oracle.refine_forward(qq, kk, i, 3, 1, 1, -1, -1, 3, nheads=1, stride0=1, dist_type="prod",
                                            itype="float")
        return float((dd.double() * gd.double()).sum())
    eps = 1e-2
    gen = th.Generator().manual_seed(6)
    for which, g in ((0, g0), (1, g1)):
        for _ in range(6):
            idx = tuple(int(th.randint(0, s, (1,), generator=gen)) for s in q.shape)
            args = [q.clone(), k.clone()]
            args[which][idx] += eps
            up = loss(*args)
            args[which][idx] -= 2 * eps
            dn = loss(*args)
            num = (up - dn) / (2 * eps)
            ana = float(g.reshape(q.shape[0], 1, *q.shape[1:])[idx[0], 0, idx[1], idx[2], idx[3], idx[4]])
            assert abs(num - ana) < 1e-2 + 1e-2 * abs(ana), (which, idx, num, ana)
