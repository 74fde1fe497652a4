"""GPU: the configurations BASELINE.json names, checked NUMERICALLY against the reference's own kernels
(oracle/_ref, driven by oracle/ref_cuda.py the way lib/stnls/search/impl/*.py and lib/stnls/agg/gather_add.py
drive them) -- forward (dists, indices), aggregation and every gradient:

  C2  README shape at full size (128x128, stride1=0.5, l2, anchor): forward AND backward
  C3  RVRT-style, full resolution 256x256 x 64 channels, a 4-frame slice
  C4  1080p operator arguments (ws=9, wt=3, ps=3, K=20, HD=4, F=64) on a 256x256 crop, 7 frames
  C5  RefineSearch(ws=3,wt=1,wr=3,k=30,kr=-1,ps=3,stride0=2,l2,float) with Ks=30 input offsets
      + WeightedPatchSum(ps=3,stride0=2), forward and backward
      (lib/csrc/search/refinement_bilin2d_kernel.cu:18-185,359-491)

The reference kernels materialise 16 bytes per candidate, which is what bounds the sizes
here: the C4 crop needs ~17 GB of reference intermediates, the full 1080p frame 4.8 TB.
Tolerances: dists rtol=atol=1e-4 (indices equal wherever distances are separated by more than 4 tol);
gradients rtol=atol=1e-3 against values of order 1-100 (2e-3 for flow gradients) -- the backward uses atomics.
"""
import pytest
import torch as th

import stnls_b200
from util import assert_close, assert_topk_equal

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def refc(cuda):
    from oracle import ref_cuda
    if not ref_cuda.available():
        pytest.skip("oracle/_ref/stnls_cuda_ref*.so not present")
    ref_cuda.mod()
    return ref_cuda


def nls_inputs(cuda, B, T, F, H, W, wt, seed):
    g = th.Generator(device="cuda").manual_seed(seed)
    q, k, v = [th.randn(B, T, F, H, W, device=cuda, generator=g) for _ in range(3)]
    shape = (B, 1, T, 2 * wt, 2, H, W)
    flows = th.rand(shape, device=cuda, generator=g) / 2 + th.randint(-3, 3, shape, device=cuda, generator=g) + 0.2
    return q, k, v, flows, g


def to5(g):
    B, HD, T, F, H, W = g.shape
    return g.permute(0, 2, 1, 3, 4, 5).reshape(B, T, HD * F, H, W)


def swaps(i, i_ref):
    """number of kept slots whose offset differs from the reference's (near-tie swaps), for the log"""
    return int(((i.float() - i_ref.float()).abs().amax(-1) > 1e-3).sum())


def nls_case(cuda, refc, name, B, T, F, H, W, HD, ws, wt, ps, K, stride1, dist_type, self_action, scale=10.0,
             agg=True, per_head=False):
    """search fwd -> (reference indices) -> search bwd (+ aggregation fwd/bwd) against the reference kernels"""
    q, k, v, flows, g = nls_inputs(cuda, B, T, F, H, W, wt, seed=11)
    kw = dict(nheads=HD, stride0=1, stride1=stride1, dist_type=dist_type, self_action=self_action, itype="float")
    search = stnls_b200.search.NonLocalSearch(ws, wt, ps, K, **kw)
    qg, kg, fg = q.clone().requires_grad_(True), k.clone().requires_grad_(True), flows.clone().requires_grad_(True)
    d, i = search(qg, kg, fg)
    if per_head:
        # the reference kernels index with 32-bit accessors (numel < 2^31): run them one head at a time -- heads are
        # independent (lib/csrc/search/non_local_search_int_kernel.cu:56-57)
        Fh = F // HD
        parts = []
        for h in range(HD):
            parts.append(refc.nls_forward(q[:, :, h * Fh:(h + 1) * Fh].contiguous(), k[:, :, h * Fh:(h + 1) * Fh].contiguous(),
                                          flows, ws, wt, ps, K, **dict(kw, nheads=1)))
            th.cuda.empty_cache()
        dr, ir = th.cat([p[0] for p in parts], 1), th.cat([p[1] for p in parts], 1)
    else:
        dr, ir = refc.nls_forward(q, k, flows, ws, wt, ps, K, **kw)
    th.cuda.empty_cache()
    assert_topk_equal(d, i, dr, ir, exact=False, what=name + " forward", last_slot_clear=False)
    nsw = swaps(i.detach(), ir)
    print(f"[{name}] {nsw} of {ir[..., 0].numel()} kept slots differ from the reference (near-tie swaps)")
    assert nsw <= 2e-4 * ir[..., 0].numel() + 8, f"{name}: {nsw} index swaps"
    # backward of the search at OUR kept indices (the reference backward takes inds as an input)
    gd = th.randn(d.shape, device=cuda, generator=g)
    gi = th.randn(i.shape, device=cuda, generator=g)
    gd = th.where(th.isfinite(d.detach()), gd, th.zeros_like(gd))
    ((d * gd).sum() + (i * gi).sum()).backward()
    r0, r1, rf = refc.search_backward(q, k, flows, gd, gi, i.detach(), wt, ps, nheads=HD, stride0=1,
                                      dist_type=dist_type, itype="float", flows_grad=True, dists=d.detach())
    assert_close(qg.grad, to5(r0), 1e-3, 1e-3, name + " grad_q")
    assert_close(kg.grad, to5(r1), 1e-3, 1e-3, name + " grad_k")
    assert_close(fg.grad, rf, 2e-3, 2e-3, name + " grad_flows")
    del r0, r1, rf
    if not agg:
        return
    # aggregation forward / backward on the same (weights, indices)
    sign = -1.0 if dist_type == "l2" else 1.0
    w = th.softmax(sign * scale * d.detach(), -1)
    aggm = stnls_b200.agg.WeightedPatchSum(ps, 1, itype="float")
    vg, wg, ig = v.clone().requires_grad_(True), w.clone().requires_grad_(True), i.detach().clone().requires_grad_(True)
    out = aggm(vg, wg, ig)
    outr, cr = refc.gather_add_forward(v, w, i.detach(), ps, 1, itype="float")
    assert_close(out, outr, 1e-4, 1e-4, name + " aggregated video")
    go = th.randn(out.shape, device=cuda, generator=g)
    (out * go).sum().backward()
    gv, gw, gf = refc.gather_add_backward(go, v, w, i.detach(), cr, ps, 1, itype="float")
    assert_close(vg.grad, to5(gv), 1e-3, 1e-3, name + " grad_v")
    assert_close(wg.grad, gw, 1e-3, 1e-3, name + " grad_weights")
    assert_close(ig.grad, gf, 2e-3, 2e-3, name + " grad_inds")


def test_c2_full_size_forward_backward(cuda, refc):
    """configs[1] at its full size: B=1,T=5,F=16,128x128, ws=5,wt=2,ps=3,K=10,HD=2, stride1=0.5, float, l2, anchor"""
    nls_case(cuda, refc, "C2", 1, 5, 16, 128, 128, 2, 5, 2, 3, 10, 0.5, "l2", "anchor")


def test_c2_prod_full_size(cuda, refc):
    nls_case(cuda, refc, "C2-prod", 1, 5, 16, 128, 128, 2, 5, 2, 3, 10, 0.5, "prod", "anchor", agg=False)


def test_c3_full_resolution_slice(cuda, refc):
    """configs[2] (RVRT-style): F=64, 256x256, ws=9, wt=1, ps=1, K=9, HD=4, prod -- a 4-frame slice of one clip"""
    nls_case(cuda, refc, "C3", 1, 4, 64, 256, 256, 4, 9, 1, 1, 9, 1.0, "prod", None)


def test_c4_crop(cuda, refc):
    """configs[3] (1080p) operator arguments on a 256x256 crop of 7 frames: ws=9, wt=3, ps=3, K=20, HD=4, F=64"""
    free, _ = th.cuda.mem_get_info()
    if free < 60e9:
        pytest.skip("the reference intermediates need ~40 GB of device memory")
    nls_case(cuda, refc, "C4-crop", 1, 7, 64, 256, 256, 4, 9, 3, 3, 20, 1.0, "prod", None, per_head=True)
    th.cuda.empty_cache()


def test_c5_shape_refine_agg(cuda, refc):
    """configs[4] shape: RefineSearch on Ks=30 input offsets, wr=3, stride0=2, l2 + WeightedPatchSum(ps=3, stride0=2)"""
    B, T, F, H, W, HD = 1, 4, 32, 96, 128, 1
    Ks, wr, K, ps, s0, wt, ws = 30, 3, 30, 3, 2, 1, 3
    g = th.Generator(device="cuda").manual_seed(21)
    q, k, v = [th.randn(B, T, F, H, W, device=cuda, generator=g) for _ in range(3)]
    nH, nW = (H - 1) // s0 + 1, (W - 1) // s0 + 1
    fl = th.zeros(B, HD, T, nH, nW, Ks, 3, device=cuda)
    dt = th.randint(-1, 2, fl[..., 0].shape, device=cuda, generator=g)
    tq = th.arange(T, device=cuda).view(1, 1, T, 1, 1, 1)
    fl[..., 0] = ((tq + dt).clamp(0, T - 1) - tq).float()
    fl[..., 1:] = th.randint(-8, 8, fl[..., 1:].shape, device=cuda, generator=g).float() \
        + th.rand(fl[..., 1:].shape, device=cuda, generator=g) / 2 + 0.2
    kw = dict(nheads=HD, stride0=s0, stride1=1, dist_type="l2", self_action=None, itype="float")
    search = stnls_b200.search.RefineSearch(ws, wt, wr, K, -1, ps, **kw)
    qg, kg, fg = q.clone().requires_grad_(True), k.clone().requires_grad_(True), fl.clone().requires_grad_(True)
    d, i = search(qg, kg, fg)
    assert tuple(d.shape) == (B, HD, T, nH, nW, K)
    dr, ir, ksr, rfr = refc.refine_forward(q, k, fl, ws, wt, wr, K, -1, ps, **kw)
    assert_topk_equal(d, i, dr, ir, exact=False, what="C5 refine forward", last_slot_clear=False)
    nsw = swaps(i.detach(), ir)
    print(f"[C5] {nsw} of {ir[..., 0].numel()} kept slots differ from the reference (near-tie swaps)")
    # refine backward needs kselect / reflect of the same selection: compare where the selections agree exactly
    same = bool(((i.detach() - ir).abs().amax(-1) < 1e-4).all())
    gd = th.randn(d.shape, device=cuda, generator=g)
    gi = th.randn(i.shape, device=cuda, generator=g)
    # aggregation on our own output
    w = th.softmax(-10.0 * d, -1)
    aggm = stnls_b200.agg.WeightedPatchSum(ps, s0, itype="float")
    vg = v.clone().requires_grad_(True)
    out = aggm(vg, w, i)
    outr, cr = refc.gather_add_forward(v, w.detach(), i.detach(), ps, s0, itype="float")
    assert_close(out, outr, 1e-4, 1e-4, "C5 aggregated video")
    go = th.randn(out.shape, device=cuda, generator=g)
    # one backward through aggregation -> softmax -> refine search (plus a direct term on dists / inds)
    ((out * go).sum() + (d * gd).sum() + (i * gi).sum()).backward()
    # reference chain: gather_add backward -> softmax backward (torch) -> refine backward
    gv, gw, gf = refc.gather_add_backward(go, v, w.detach(), i.detach(), cr, ps, s0, itype="float")
    assert_close(vg.grad, to5(gv), 1e-3, 1e-3, "C5 grad_v")
    d_leaf = d.detach().clone().requires_grad_(True)
    th.softmax(-10.0 * d_leaf, -1).backward(gw)
    gd_tot, gi_tot = gd + d_leaf.grad, gi + gf
    if same:
        r0, r1, rf = refc.search_backward(q, k, fl, gd_tot, gi_tot, ir, wt, ps, nheads=HD, stride0=s0, dist_type="l2",
                                          itype="float", flows_grad=True, kselect=ksr, reflect=rfr, refine=True)
        assert_close(qg.grad, to5(r0), 2e-3, 2e-3, "C5 grad_q")
        assert_close(kg.grad, to5(r1), 2e-3, 2e-3, "C5 grad_k")
        assert_close(fg.grad, rf, 4e-3, 4e-3, "C5 grad_flows")
    else:
        pytest.skip(f"C5: {nsw} near-tie swaps against the reference selection; forward and aggregation checked")


# ---- 1-GPU shard emulation: the tq0 / tk0 index math of every kernel ---------------------------------------------

def _shard_run(cuda, T, wt, parts, refine):
    """Run search -> softmax -> aggregation fwd+bwd on frame shards (shard=dict(T,tq0,tk0), k/v = the frames a rank
    would hold) one after the other on one GPU, and unsharded; returns both result sets."""
    from stnls_b200.parallel import ShardPlan
    B, F, H, W, HD, ws, ps, K = 1, 8, 24, 20, 2, 3, 3, 6
    g = th.Generator().manual_seed(5)
    q, k, v = [th.randn(B, T, F, H, W, generator=g).to(cuda) for _ in range(3)]
    shape = (B, 1, T, 2 * wt, 2, H, W)
    flows = (th.rand(shape, generator=g) / 2 + th.randint(-2, 3, shape, generator=g) + 0.2).to(cuda)
    gout = th.randn(B, HD, T, F // HD, H, W, generator=g).to(cuda)
    search = stnls_b200.search.NonLocalSearch(ws, wt, ps, K, nheads=HD, stride1=0.5, dist_type="l2",
                                              self_action="anchor", itype="float")
    refs = stnls_b200.search.RefineSearch(ws, wt, 3, K, -1, ps, nheads=HD, stride0=1, stride1=0.5, dist_type="l2",
                                          itype="float")
    agg = stnls_b200.agg.WeightedPatchSum(ps=ps, stride0=1, itype="float")

    def run(plan):
        a, b, fa, fb = plan.a, plan.b, plan.fa, plan.fb
        qs = q[:, a:b].clone().requires_grad_(True)
        ks = k[:, fa:fb].clone().requires_grad_(True)
        vs = v[:, fa:fb].clone().requires_grad_(True)
        fs = flows[:, :, a:b].clone().requires_grad_(True)
        sh = plan.shard if plan.world > 1 else None
        d, i = search(qs, ks, fs, shard=sh)
        if refine:
            d, i = refs(qs, ks, i, shard=sh)
        w = th.softmax(-10.0 * d, -1)
        out = agg(vs, w, i, shard=sh)
        (out * gout[:, :, a:b]).sum().backward()
        return dict(out=out.detach(), dists=d.detach(), inds=i.detach(), gq=qs.grad, gk=ks.grad, gv=vs.grad,
                    gf=fs.grad)

    full = run(ShardPlan(T, wt, 1, 0))
    acc = dict(gk=th.zeros_like(full["gk"]), gv=th.zeros_like(full["gv"]))
    pieces = []
    for r in range(parts):
        plan = ShardPlan(T, wt, parts, r)
        res = run(plan)
        acc["gk"][:, plan.fa:plan.fb] += res["gk"]
        acc["gv"][:, plan.fa:plan.fb] += res["gv"]
        pieces.append((plan, res))
    return full, acc, pieces


@pytest.mark.parametrize("refine", [False, True], ids=["nls", "nls+refine"])
@pytest.mark.parametrize("T,wt,parts", [(9, 2, 2), (10, 1, 3), (7, 3, 2)])
def test_shard_emulation_one_gpu(T, wt, parts, refine, cuda):
    full, acc, pieces = _shard_run(cuda, T, wt, parts, refine)
    for plan, res in pieces:
        a, b = plan.a, plan.b
        assert th.equal(res["inds"], full["inds"][:, :, a:b]), "sharded indices must be bit-identical"
        assert th.equal(res["dists"], full["dists"][:, :, a:b]), "sharded dists must be bit-identical"
        assert_close(res["out"], full["out"][:, :, a:b], 1e-5, 1e-5, "sharded aggregated video")
        assert_close(res["gq"], full["gq"][:, a:b], 1e-4, 1e-4, "sharded grad_q")
        assert_close(res["gf"], full["gf"][:, :, a:b], 1e-3, 1e-3, "sharded grad_flows")
    assert_close(acc["gk"], full["gk"], 1e-4, 1e-4, "halo-summed grad_k")
    assert_close(acc["gv"], full["gv"], 1e-4, 1e-4, "halo-summed grad_v")


def test_shard_rejects_missing_halo(cuda):
    """the resident key frames must cover the temporal windows of the owned frames (host check in the C ABI)"""
    T, wt = 9, 2
    q = th.randn(1, 4, 8, 16, 16, device=cuda)
    k = th.randn(1, 4, 8, 16, 16, device=cuda)   # frames [0,4) only: the halo [4,6) is missing
    fl = th.zeros(1, 1, 4, 2 * wt, 2, 16, 16, device=cuda)
    search = stnls_b200.search.NonLocalSearch(3, wt, 3, 4, nheads=2, itype="float")
    with pytest.raises(ValueError, match="do not cover"):
        search(q, k, fl, shard=dict(T=T, tq0=0, tk0=0))


def test_shard_out_of_window_offsets_are_skipped(cuda):
    """caller-supplied offsets (RefineSearch / WeightedPatchSum / NonLocalGather) may point at a frame of the clip
    that is NOT resident on this rank: such neighbours are skipped -- never read or written out of bounds.  Equal to
    the unsharded run with those neighbours' weights set to zero."""
    T, Tq, tk0, Tk = 12, 3, 2, 6          # owned frames [4,7), resident [2,8)
    tq0 = 4
    B, F, H, W, HD, K, ps = 1, 8, 16, 12, 2, 5, 3
    g = th.Generator().manual_seed(9)
    v = th.randn(B, T, F, H, W, generator=g).to(cuda)
    w = th.rand(B, HD, Tq, H, W, K, generator=g).to(cuda)
    inds = th.zeros(B, HD, Tq, H, W, K, 3)
    inds[..., 0] = th.randint(-6, 7, inds[..., 0].shape, generator=g).float()      # reaches well outside [2,8)
    inds[..., 1:] = th.randint(-3, 4, inds[..., 1:].shape, generator=g).float() + th.rand(inds[..., 1:].shape, generator=g)
    inds = inds.to(cuda)
    tq = th.arange(tq0, tq0 + Tq, device=cuda).view(1, 1, Tq, 1, 1, 1)
    tn = tq + inds[..., 0]
    tn = th.where(tn < 0, -tn, tn)
    tn = th.where(tn > T - 1, 2 * (T - 1) - tn, tn)                                   # reflect once, as the kernels do
    res = (tn >= tk0) & (tn <= tk0 + Tk - 1)
    assert bool((~res).any()) and bool(res.any())
    for itype in ("float", "int"):
        ii = inds if itype == "float" else inds.round().int()
        agg = stnls_b200.agg.WeightedPatchSum(ps, 1, itype=itype)
        vs = v[:, tk0:tk0 + Tk].clone().requires_grad_(True)
        ws_ = w.clone().requires_grad_(True)
        out = agg(vs, ws_, ii, shard=dict(T=T, tq0=tq0, tk0=tk0))
        out.sum().backward()
        # unsharded twin: all T frames resident, the non-resident neighbours masked out through their weights
        vf = v.clone().requires_grad_(True)
        wf = th.zeros(B, HD, T, H, W, K, device=cuda)
        wf[:, :, tq0:tq0 + Tq] = w * res
        i_full = th.zeros((B, HD, T, H, W, K, 3), device=cuda, dtype=ii.dtype)
        i_full[:, :, tq0:tq0 + Tq] = ii
        out_f = agg(vf, wf, i_full)
        out_f[:, :, tq0:tq0 + Tq].sum().backward()
        assert_close(out, out_f[:, :, tq0:tq0 + Tq], 1e-5, 1e-5, "out-of-window neighbours must be skipped")
        assert_close(vs.grad, vf.grad[:, tk0:tk0 + Tk], 1e-4, 1e-4, "grad_v with skipped neighbours")
        assert float(vf.grad[:, :tk0].abs().max()) == 0 and float(vf.grad[:, tk0 + Tk:].abs().max()) == 0
        assert float((ws_.grad * (~res)).abs().max()) == 0, "skipped neighbours get no weight gradient"
