"""GPU: the one-node step (stnls_b200/fused.py: 8 launches) against the module chain it replaces
(NonLocalSearch -> softmax_scaled -> WeightedPatchSum through autograd: 14 launches).  The same kernels run on the
same values and the folded normalise steps repeat the arithmetic of csrc/normz.cu in the same order, so everything
that is not accumulated with atomics must agree bit for bit; the scattered gradients to 1e-5 of their scale."""
import pytest
import torch as th

import stnls_b200
from stnls_b200 import ops, parallel

pytestmark = pytest.mark.gpu


def _inputs(cuda, seed, B, T, C, Cv, H, W, wt, stride0, HD_f=1):
    g = th.Generator().manual_seed(seed)
    q = th.randn(B, T, C, H, W, generator=g).to(cuda)
    k = th.randn(B, T, C, H, W, generator=g).to(cuda)
    v = th.randn(B, T, Cv, H, W, generator=g).to(cuda)
    nH, nW = (H - 1) // stride0 + 1, (W - 1) // stride0 + 1
    flows = (th.randn(B, HD_f, T, 2 * wt, 2, nH, nW, generator=g) * 1.7).to(cuda)
    return q, k, v, flows


def _run(model, q, k, v, flows, gout_seed=5):
    q, k, v, flows = (x.detach().clone().requires_grad_(True) for x in (q, k, v, flows))
    n0 = ops.LAUNCHES
    out, dists, inds = model(q, k, v, flows)
    gout = th.randn(out.shape, generator=th.Generator().manual_seed(gout_seed)).to(out.device)
    grads = th.autograd.grad(out, (q, k, v, flows), grad_outputs=gout)
    return (out.detach(), dists.detach(), inds.detach()) + tuple(grads), ops.LAUNCHES - n0


CASES = {
    # name: (B, T, C, Cv, H, W, HD, ws, wt, ps, K, stride0, stride1, dist, self_action, agg_ps, expected launches)
    "c2_like": (1, 5, 16, 16, 64, 48, 2, 5, 2, 3, 10, 1, 0.5, "l2", "anchor", 3, 8),
    "prod_other_v_channels": (2, 3, 8, 12, 33, 29, 1, 3, 1, 3, 7, 1, 1.0, "prod", None, 3, 8),
    "stride0_2": (1, 3, 16, 16, 40, 36, 2, 5, 1, 3, 9, 2, 1.0, "l2", None, 3, 8),
    "agg_ps1_two_launch_normalise": (1, 3, 8, 8, 24, 20, 1, 3, 1, 1, 5, 1, 1.0, "prod", None, 1, 9),
}


@pytest.mark.parametrize("name", sorted(CASES))
def test_fused_step_equals_the_module_chain(name, cuda, monkeypatch):
    B, T, C, Cv, H, W, HD, ws, wt, ps, K, s0, s1, dist, act, aps, nlaunch = CASES[name]
    q, k, v, flows = _inputs(cuda, 3, B, T, C, Cv, H, W, wt, s0)
    search = stnls_b200.search.NonLocalSearch(ws, wt, ps, K, nheads=HD, stride0=s0, stride1=s1, dist_type=dist,
                                              self_action=act, itype="float")
    agg = stnls_b200.agg.WeightedPatchSum(aps, s0, itype="float")
    model = parallel.ShardedSearchAgg(search, agg, parallel.ShardPlan(T, wt, 1, 0), scale=10.0)
    monkeypatch.setenv("STNLS_FUSED_STEP", "0")
    ref, n_ref = _run(model, q, k, v, flows)
    monkeypatch.setenv("STNLS_FUSED_STEP", "1")
    got, n_got = _run(model, q, k, v, flows)
    assert n_ref == 14 and n_got == nlaunch, f"launches: chain {n_ref}, fused {n_got}"
    names = ["out", "dists", "inds", "grad_q", "grad_k", "grad_v", "grad_flows"]
    for nm, a, b in zip(names, ref, got):
        assert a.shape == b.shape, nm
        if nm in ("dists", "inds"):
            assert th.equal(a, b), nm
        elif nm == "out":   # the weights are the same bits, the sum over K runs in the same order
            assert th.equal(a, b), f"out differs: max {float((a - b).abs().max()):.3g}"
        else:               # accumulated with atomics: order-dependent rounding only
            tol = 1e-5 * max(1.0, float(a.abs().max()))
            assert float((a - b).abs().max()) <= tol, f"{nm}: max diff {float((a - b).abs().max()):.3g} > {tol:.3g}"


def test_fused_step_on_a_frame_shard(cuda, monkeypatch):
    """owned frames [2, 5) of a 7-frame clip with the halo frames resident: tq0 / tk0 index math of the fused node"""
    T, wt = 7, 1
    q, k, v, flows = _inputs(cuda, 9, 1, T, 16, 16, 32, 32, wt, 1)
    search = stnls_b200.search.NonLocalSearch(5, wt, 3, 8, nheads=2, stride1=1.0, dist_type="l2", itype="float")
    agg = stnls_b200.agg.WeightedPatchSum(3, 1, itype="float")
    plan = parallel.ShardPlan(T, wt, 1, 0)
    plan.a, plan.b = 2, 5
    plan.fa, plan.fb = parallel.needed_frames(2, 5, T, wt)
    model = parallel.ShardedSearchAgg(search, agg, plan, scale=10.0)
    args = (q[:, 2:5], k[:, plan.fa:plan.fb], v[:, plan.fa:plan.fb], flows[:, :, 2:5])
    monkeypatch.setenv("STNLS_FUSED_STEP", "0")
    ref, _ = _run(model, *args)
    monkeypatch.setenv("STNLS_FUSED_STEP", "1")
    got, n = _run(model, *args)
    assert n == 8
    for a, b in zip(ref[:3], got[:3]):
        assert th.equal(a, b)
    for a, b in zip(ref[3:], got[3:]):
        assert float((a - b).abs().max()) <= 1e-5 * max(1.0, float(a.abs().max()))


def test_multi_pack_unpack_equal_the_single_calls(cuda):
    g = th.Generator().manual_seed(1)
    a = th.randn(2, 3, 12, 19, 37, generator=g).to(cuda)       # F = 6 per head -> padded to 8
    b = th.randn(2, 5, 12, 19, 37, generator=g).to(cuda)
    c = th.randn(2, 5, 24, 19, 37, generator=g).to(cuda)       # other channel count
    singles = [ops.pack_cl(x, 2) for x in (a, b, c)]
    multi = ops.pack_cl_multi([a, b, c], 2)
    for s_, m in zip(singles, multi):
        assert th.equal(s_, m)
    back = ops.unpack_cl_multi(multi, [a.shape, b.shape, c.shape], 2)
    for x, y in zip((a, b, c), back):
        assert th.equal(x, y)
