"""GPU: the tile forward kernel (csrc/search_fwd_tile.cuh) against the CPU oracle on shapes chosen to hit its special
cases: frames that are not a multiple of the 8x8 tile, border tiles (mirror fill of the TMA box), window shifts at the
frame border (full_ws), stride1 = 0.5 sub-windows, both channel widths (Fp 8 / 16), stride0 = 2, anchoring, flows too
wild for the staged box (queries deferred to the generic path inside the same launch) and non-finite flows.
Every case also runs with STNLS_FWD_TILE=0 (table kernel): both must match the oracle, float tolerance 1e-4."""
import os

import pytest
import torch as th

import stnls_b200
from util import assert_topk_equal

pytestmark = pytest.mark.gpu

CASES = [
    # name, F, HD, H, W, T, ws, wt, ps, K, stride0, stride1, dist, self_action, flow_scale
    ("c2like_fp8_half_l2_anchor", 16, 2, 43, 37, 5, 5, 2, 3, 10, 1, 0.5, "l2", "anchor", 1.0),
    ("fp8_half_prod", 16, 2, 40, 40, 3, 5, 1, 3, 7, 1, 0.5, "prod", None, 1.0),
    ("fp8_ws5_s1_l2", 8, 1, 35, 46, 3, 5, 1, 3, 6, 1, 1.0, "l2", "anchor", 1.0),
    ("fp8_ws3_prod", 16, 2, 33, 34, 3, 3, 1, 3, 5, 1, 1.0, "prod", None, 1.0),
    ("c4like_fp16_ws9_ps3_prod", 32, 2, 41, 38, 3, 9, 1, 3, 20, 1, 1.0, "prod", None, 1.0),
    ("c3like_fp16_ws9_ps1_prod", 32, 2, 36, 45, 3, 9, 1, 1, 9, 1, 1.0, "prod", None, 1.0),
    ("fp16_ws9_ps1_l2_anchor", 16, 1, 36, 36, 3, 9, 1, 1, 9, 1, 1.0, "l2", "anchor", 1.0),
    ("fp16_ws5_ps3_l2", 32, 2, 37, 33, 5, 5, 2, 3, 32, 1, 1.0, "l2", None, 1.0),
    ("fp16_ws5_ps1_prod", 16, 1, 34, 34, 3, 5, 1, 1, 4, 1, 1.0, "prod", "anchor", 1.0),
    ("fp16_half_l2", 32, 2, 38, 41, 3, 5, 1, 3, 8, 1, 0.5, "l2", "anchor", 1.0),
    ("fp16_ws3_l2", 16, 1, 34, 39, 3, 3, 1, 3, 9, 1, 1.0, "l2", None, 1.0),
    ("stride0_2", 16, 2, 58, 62, 3, 5, 1, 3, 6, 2, 1.0, "l2", "anchor", 1.0),
    ("wild_flows_deferred", 16, 2, 48, 48, 3, 5, 1, 3, 6, 1, 0.5, "l2", "anchor", 4.0),
]


def inputs(F, H, W, T, wt, stride0, flow_scale, seed):
    g = th.Generator().manual_seed(seed)
    q, k = [th.randn(1, T, F, H, W, generator=g) for _ in range(2)]
    nH, nW = (H - 1) // stride0 + 1, (W - 1) // stride0 + 1
    shape = (1, 1, T, 2 * wt, 2, nH, nW)
    flows = (th.rand(shape, generator=g) / 2 + th.randint(-3, 3, shape, generator=g) + 0.2) * flow_scale
    return q, k, flows


def run_both(search, q, k, flows, cuda):
    res = []
    for tile in ("1", "0"):
        os.environ["STNLS_FWD_TILE"] = tile
        try:
            res.append(search(q.to(cuda), k.to(cuda), flows.to(cuda)))
        finally:
            os.environ.pop("STNLS_FWD_TILE", None)
    return res


@pytest.mark.parametrize("c", CASES, ids=[c[0] for c in CASES])
def test_tile_kernel_vs_oracle(c, cuda, oracle):
    name, F, HD, H, W, T, ws, wt, ps, K, s0, s1, dist, act, fs = c
    q, k, flows = inputs(F, H, W, T, wt, s0, fs, seed=hash(name) % 1000)
    kw = dict(nheads=HD, stride0=s0, stride1=s1, dist_type=dist, self_action=act, itype="float")
    d_ref, i_ref = oracle.nls_forward(q, k, flows, ws, wt, ps, K, **kw)
    search = stnls_b200.search.NonLocalSearch(ws, wt, ps, K, **kw)
    for (d, i), which in zip(run_both(search, q, k, flows, cuda), ("tile", "table")):
        assert_topk_equal(d, i, d_ref, i_ref, exact=False, what=f"{name} [{which}]", last_slot_clear=False)
        nsw = int(((i.cpu() - i_ref).abs().amax(-1) > 1e-3).sum())
        assert nsw <= 4 + 1e-4 * i_ref[..., 0].numel(), f"{name} [{which}]: {nsw} index swaps against the oracle"


def test_tile_kernel_no_full_ws_and_zero_flows(cuda, oracle):
    """full_ws=False leaves part of the window outside the frame (invalid candidates, +-inf); zero flows make every
    query of a tile share one window start (exact ties between frames are broken by candidate index)"""
    q, k, flows = inputs(16, 40, 36, 3, 1, 1, 0.0, seed=5)
    for full_ws, dist in ((False, "l2"), (True, "prod")):
        kw = dict(nheads=2, stride0=1, stride1=1.0, dist_type=dist, self_action="anchor", full_ws=full_ws, itype="float")
        d_ref, i_ref = oracle.nls_forward(q, k, flows, 5, 1, 3, 12, **kw)
        search = stnls_b200.search.NonLocalSearch(5, 1, 3, 12, **kw)
        for (d, i), which in zip(run_both(search, q, k, flows, cuda), ("tile", "table")):
            assert_topk_equal(d, i, d_ref, i_ref, exact=False, what=f"full_ws={full_ws} [{which}]", last_slot_clear=False)


def test_tile_kernel_constant_video_ties(cuda, oracle):
    """a constant video makes ALL distances equal: the kept candidates must be the first K in candidate order"""
    T, H, W = 3, 34, 35
    q = th.ones(1, T, 16, H, W)
    k = th.ones(1, T, 16, H, W)
    flows = th.zeros(1, 1, T, 2, 2, H, W)
    kw = dict(nheads=2, stride0=1, stride1=1.0, dist_type="l2", self_action=None, itype="float")
    d_ref, i_ref = oracle.nls_forward(q, k, flows, 3, 1, 3, 7, **kw)
    search = stnls_b200.search.NonLocalSearch(3, 1, 3, 7, **kw)
    for (d, i), which in zip(run_both(search, q, k, flows, cuda), ("tile", "table")):
        assert th.equal(i.cpu(), i_ref), f"tie order differs from the oracle [{which}]"
        assert th.equal(d.cpu(), d_ref)


def test_tile_kernel_non_finite_flows(cuda, oracle):
    """NaN / inf flows: those queries take the generic path; every output slot is written, the rest is unaffected"""
    q, k, flows = inputs(16, 40, 40, 3, 1, 1, 1.0, seed=9)
    bad = flows.clone()
    bad[0, 0, 1, 0, 0, 5, 7] = float("nan")
    bad[0, 0, 2, 1, 1, 20, 33] = float("inf")
    kw = dict(nheads=2, stride0=1, stride1=0.5, dist_type="l2", self_action="anchor", itype="float")
    search = stnls_b200.search.NonLocalSearch(5, 1, 3, 6, **kw)
    d0, i0 = search(q.to(cuda), k.to(cuda), flows.to(cuda))
    d1, i1 = search(q.to(cuda), k.to(cuda), bad.to(cuda))
    same = th.ones(d0.shape[2:5], dtype=th.bool, device=cuda)
    same[1, 5, 7] = False
    same[2, 20, 33] = False
    assert th.equal(d0[:, :, same], d1[:, :, same]) and th.equal(i0[:, :, same], i1[:, :, same])
    d_ref, i_ref = oracle.nls_forward(q, k, bad, 5, 1, 3, 6, **kw)
    fin = th.isfinite(d_ref)
    assert bool((th.isfinite(d1.cpu()) == fin).all())
    assert th.allclose(d1.cpu()[fin], d_ref[fin], atol=1e-4, rtol=1e-4)


REFINE_CASES = [
    # name, F, HD, H, W, T, Ks, wr, ps, K, stride0, dist
    ("c5like_fp32_l2", 32, 1, 40, 44, 3, 30, 3, 3, 30, 2, "l2"),
    ("fp16_prod", 32, 2, 33, 37, 3, 12, 3, 3, 9, 1, "prod"),
    ("fp8_l2", 16, 2, 36, 34, 4, 7, 3, 3, 32, 1, "l2"),
    ("fp16_ps1", 16, 1, 35, 35, 3, 9, 3, 1, 5, 1, "l2"),
    ("fp16_wr1", 32, 2, 34, 36, 3, 20, 1, 3, 6, 1, "prod"),
]


@pytest.mark.parametrize("c", REFINE_CASES, ids=[c[0] for c in REFINE_CASES])
def test_refine_channel_kernel_vs_oracle(c, cuda, oracle):
    """RefineSearch forward: the channel-parallel kernel (csrc/search_fwd_refine_chan.cuh) and the table kernel against
    the CPU oracle -- offsets that leave the frame / point at its border take the generic path inside the launch"""
    name, F, HD, H, W, T, Ks, wr, ps, K, s0, dist = c
    g = th.Generator().manual_seed(hash(name) % 1000)
    q, k = [th.randn(1, T, F, H, W, generator=g) for _ in range(2)]
    nH, nW = (H - 1) // s0 + 1, (W - 1) // s0 + 1
    fl = th.zeros(1, HD, T, nH, nW, Ks, 3)
    fl[..., 0] = th.randint(-1, 2, fl[..., 0].shape, generator=g).float()
    fl[..., 1:] = th.randint(-6, 7, fl[..., 1:].shape, generator=g).float() + th.rand(fl[..., 1:].shape, generator=g) / 2 + 0.2
    kw = dict(nheads=HD, stride0=s0, stride1=1, dist_type=dist, self_action=None, itype="float")
    dr, ir, ksr, rfr = oracle.refine_forward(q, k, fl, 3, 1, wr, K, -1, ps, **kw)
    search = stnls_b200.search.RefineSearch(3, 1, wr, K, -1, ps, **kw)
    for chan in ("1", "0"):
        os.environ["STNLS_REFINE_CHAN"] = chan
        try:
            d, i = search(q.to(cuda), k.to(cuda), fl.to(cuda))
        finally:
            os.environ.pop("STNLS_REFINE_CHAN", None)
        assert_topk_equal(d, i, dr, ir, exact=False, what=f"{name} [chan={chan}]", last_slot_clear=False)
        nsw = int(((i.cpu() - ir).abs().amax(-1) > 1e-3).sum())
        assert nsw <= 4 + 1e-4 * ir[..., 0].numel(), f"{name} [chan={chan}]: {nsw} index swaps against the oracle"
