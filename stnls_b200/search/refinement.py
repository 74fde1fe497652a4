"""stnls_b200.search.RefineSearch -- mirrors lib/stnls/search/refinement.py:29-181 and
lib/stnls/search/impl/refinement.py:18-181 on the fused sm_100a kernels.

flows: (B,HD,T,nH,nW,Ks,3) (or (B,HD,Q,Ks,3)) relative offsets (dt,dh,dw), e.g. the `inds` returned by
NonLocalSearch.  Output dists (B,HD,T,nH,nW,K), inds (...,K,3) with K candidates out of Ks*wr*wr.
"""
import torch as th

from .. import ops
from .._C import SearchOpts
from ..utils import extract_pairs
from .non_local_search import DIST_MENU, check_common, normz_bwd, shard_of, topk_each, topk_fallback


def filter_k(inds, kr):
    """lib/stnls/search/utils.py:115-125"""
    K = inds.shape[-2]
    kr = K if kr is None else kr
    if kr <= 0:
        return inds
    if isinstance(kr, float):
        assert 0 < kr <= 1
        Ks = int(K * kr)
    else:
        Ks = int(kr)
    return inds[..., :Ks, :].contiguous()


class RefineSearchFunction(th.autograd.Function):

    @staticmethod
    def forward(ctx, vid0, vid1, flows, ws, wt, wr, k, kr=-1, ps=1, nheads=1, stride0=4, stride1=1,
                strideQ=None, dilation=1, pt=1, dist_type="l2", restricted_radius=False,
                reflect_bounds=True, full_ws=True, topk_mode="all", self_action=None, use_adj=False,
                normalize_bwd=False, k_agg=-1, off_Hq=0, off_Wq=0, itype="float", shard=None):
        each = topk_mode == "each"  # top-k inside every group of wr*wr candidates (impl/refinement.py:99-106)
        if self_action == "anchor_each":  # impl/refinement.py:76-84: any "anchor*" anchors every group
            self_action = "anchor"
        check_common(pt, "all" if each else topk_mode, self_action, normalize_bwd, dist_type, itype)
        if each and shard is not None:
            raise NotImplementedError("stnls_b200: frame shards support topk_mode='all' only")
        assert isinstance(ps, int) and (ps > 0), f"patch size is invalid [{ps}]"
        q_cl = ops.pack_cl(vid0, nheads)
        k_cl = ops.pack_cl(vid1, nheads)
        B, HD = vid0.shape[0], nheads
        Tq, qH, qW = q_cl.shape[1:4]
        Tk, kH, kW = k_cl.shape[1:4]
        F = (vid0.shape[2] // nheads) if vid0.dim() == 5 else vid0.shape[3]
        T, tq0, tk0 = shard_of(shard, Tq)
        nH, nW = (kH - 1) // stride0 + 1, (kW - 1) // stride0 + 1
        flows_in_shape = tuple(flows.shape)
        if flows.dim() == 5:  # (B,HD,Q,K,3), lib/stnls/search/utils.py:159-175
            flows = flows.reshape(flows.shape[0], flows.shape[1], Tq, nH, nW, flows.shape[-2], 3)
        assert flows.dim() == 7
        assert flows.shape[1] == HD
        Ks_in = flows.shape[-2]
        flows = filter_k(flows, kr)
        Ks = flows.shape[-2]
        dist_type_i, descending = DIST_MENU[dist_type]
        anchor = self_action is not None
        if itype == "int":
            flows_k = flows.round().int() if flows.is_floating_point() else flows.int()
            stride1_k = float(max(1, int(stride1)))
            strideQ_k = stride0 if strideQ is None else strideQ
            # flag permutation of the reference launcher (refinement_int_kernel.cu:25 vs :232-233):
            # inside the int kernel reflect_bounds := restricted_radius and full_ws := reflect_bounds
            k_reflect, k_full = bool(restricted_radius), bool(reflect_bounds)
        else:
            flows_k = flows.float()
            stride1_k = float(stride1)
            strideQ_k = stride0
            k_reflect, k_full = bool(reflect_bounds), bool(full_ws)
        flows_k = flows_k.contiguous()
        L = Ks * wr * wr
        fused = not each and ((k <= 0) or (min(k, L) - (1 if anchor else 0) <= ops.MAX_FUSED_K))
        geom = ops.make_geom(B, HD, T, F, qH, qW, kH, kW, stride0, tq0, Tq, tk0, Tk)
        opts = SearchOpts(ws, wt, Ks, wr, ps, k if fused else -1, stride0, strideQ_k, dilation, stride1_k,
                          int(k_reflect), int(k_full), 0 if use_adj else -(ps // 2), off_Hq, off_Wq,
                          dist_type_i, int(itype == "int"), int(anchor), flows_k.shape[1], 0)
        dists, inds, kselect, reflect = ops.search_forward(q_cl, k_cl, flows_k, geom, opts)
        if each:
            lead = dists.shape[:-1]
            dists, inds = topk_each(dists.view(lead + (Ks, wr * wr)), inds.view(lead + (Ks, wr * wr, 3)), k,
                                    descending, anchor_self=anchor)
            if kselect is not None:  # every entry of a group carries the group index
                kselect = kselect.view(lead + (Ks, wr * wr))[..., :dists.shape[-1]].reshape(lead + (-1,)).contiguous()
            dists, inds = dists.reshape(lead + (-1,)).contiguous(), inds.reshape(lead + (-1, 3)).contiguous()
        elif not fused:
            dists, inds, kselect = topk_fallback(dists, inds, k, anchor, descending, kselect)

        # the backward honours the *module's* reflect_bounds (impl/refinement.py:147-167)
        bopts = SearchOpts.from_buffer_copy(opts)
        bopts.reflect_bounds = int(reflect_bounds)
        ctx.save_for_backward(q_cl, k_cl, inds, kselect, reflect)
        if itype == "int":
            ctx.mark_non_differentiable(inds)
        ctx.geom, ctx.opts = geom, bopts
        ctx.vid_shapes = (tuple(vid0.shape), tuple(vid1.shape))
        ctx.nheads, ctx.k_agg, ctx.itype = nheads, k_agg, itype
        ctx.flows_shape = tuple(flows_k.shape)
        ctx.flows_in_shape, ctx.Ks_in = flows_in_shape, Ks_in
        ctx.normz = (ps, stride0, stride1) if normalize_bwd else None  # impl/refinement.py:175-176
        return dists, inds

    @staticmethod
    def backward(ctx, grad_dists, grad_inds):
        q_cl, k_cl, inds, kselect, reflect = ctx.saved_tensors
        flows_grad = ctx.itype == "float" and ctx.needs_input_grad[2]
        if ctx.k_agg > 0:  # impl/refinement.py:128-131
            grad_dists = grad_dists[..., :ctx.k_agg]
            inds = inds[..., :ctx.k_agg, :]
            if grad_inds is not None:
                grad_inds = grad_inds[..., :ctx.k_agg, :]
            if kselect is not None:
                kselect = kselect[..., :ctx.k_agg].contiguous()
        gflows_like = None
        if flows_grad:
            gflows_like = th.empty(ctx.flows_shape, device=q_cl.device, dtype=th.float32)
            if grad_inds is None:
                grad_inds = th.zeros_like(inds, dtype=th.float32)
        gq, gk, gflows = ops.search_backward(q_cl, k_cl, gflows_like, grad_dists, grad_inds, inds, kselect,
                                             reflect, ctx.geom, ctx.opts, flows_grad)
        g0 = ops.unpack_cl(gq, ctx.vid_shapes[0], ctx.nheads)
        g1 = ops.unpack_cl(gk, ctx.vid_shapes[1], ctx.nheads)
        if ctx.normz is not None:
            normz_bwd(g0, g1, *ctx.normz)
        if gflows is not None:
            Ks = gflows.shape[-2]
            if Ks != ctx.Ks_in:  # offsets dropped by `kr` get zero gradient
                pad = th.zeros(gflows.shape[:-2] + (ctx.Ks_in - Ks, 3), device=gflows.device)
                gflows = th.cat([gflows, pad], -2)
            gflows = gflows.reshape(ctx.flows_in_shape)
        return (g0, g1, gflows) + (None,) * 25


class RefineSearch(th.nn.Module):
    """Constructor signature identical to lib/stnls/search/refinement.py:116-122."""

    def __init__(self, ws, wt, wr, k, kr, ps, nheads=1, stride0=4, stride1=1, strideQ=None, dilation=1, pt=1,
                 dist_type="l2", restricted_radius=True, reflect_bounds=True, full_ws=True, topk_mode="all",
                 self_action=None, use_adj=False, normalize_bwd=False, k_agg=-1, off_Hq=0, off_Wq=0,
                 itype="float"):
        super().__init__()
        self.ws, self.wt, self.ps, self.k, self.wr, self.kr, self.nheads = ws, wt, ps, k, wr, kr, nheads
        self.stride0, self.stride1, self.strideQ, self.dilation, self.pt = stride0, stride1, strideQ, dilation, pt
        self.dist_type, self.off_Hq, self.off_Wq = dist_type, off_Hq, off_Wq
        self.restricted_radius, self.reflect_bounds, self.full_ws = restricted_radius, reflect_bounds, full_ws
        self.topk_mode, self.self_action, self.itype, self.use_adj = topk_mode, self_action, itype, use_adj
        self.normalize_bwd, self.k_agg = normalize_bwd, k_agg

    def forward(self, vid0, vid1, flows, shard=None):
        return RefineSearchFunction.apply(vid0, vid1, flows, self.ws, self.wt, self.wr, self.k, self.kr, self.ps,
                                          self.nheads, self.stride0, self.stride1, self.strideQ, self.dilation,
                                          self.pt, self.dist_type, self.restricted_radius, self.reflect_bounds,
                                          self.full_ws, self.topk_mode, self.self_action, self.use_adj,
                                          self.normalize_bwd, self.k_agg, self.off_Hq, self.off_Wq, self.itype,
                                          shard)

    def flops(self, T, F, H, W):
        return 0

    def radius(self, H, W):
        return self.ws


def extract_config(cfg, restrict=True):
    """lib/stnls/search/refinement.py (extract_config)"""
    pairs = {"ws": -1, "wt": -1, "wr": 1, "k": -1, "kr": -1, "ps": 1, "nheads": 1, "stride0": 4, "stride1": 1,
             "strideQ": None, "dilation": 1, "pt": 1, "dist_type": "l2", "restricted_radius": True,
             "reflect_bounds": True, "full_ws": True, "topk_mode": "all", "self_action": None,
             "use_adj": False, "normalize_bwd": False, "k_agg": -1, "off_Hq": 0, "off_Wq": 0, "itype": "float"}
    return extract_pairs(cfg, pairs, restrict=restrict)


def init(cfg):
    cfg = extract_config(cfg, False)
    return RefineSearch(cfg.ws, cfg.wt, cfg.wr, cfg.k, cfg.kr, cfg.ps, nheads=cfg.nheads, stride0=cfg.stride0,
                        stride1=cfg.stride1, strideQ=cfg.strideQ, dilation=cfg.dilation, pt=cfg.pt,
                        dist_type=cfg.dist_type, restricted_radius=cfg.restricted_radius,
                        reflect_bounds=cfg.reflect_bounds, full_ws=cfg.full_ws, topk_mode=cfg.topk_mode,
                        self_action=cfg.self_action, use_adj=cfg.use_adj, normalize_bwd=cfg.normalize_bwd,
                        k_agg=cfg.k_agg, off_Hq=cfg.off_Hq, off_Wq=cfg.off_Wq, itype=cfg.itype)
