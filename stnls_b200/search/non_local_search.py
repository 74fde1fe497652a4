"""stnls_b200.search.NonLocalSearch -- same constructor arguments, input/output shapes and autograd
contract as the reference module (lib/stnls/search/non_local_search.py:30-218 and
lib/stnls/search/impl/non_local_search.py:23-212), running the fused sm_100a kernels.

Differences that are visible to a caller (all documented in DESIGN.md):
  * the candidate tensor (B,HD,Q,W_t,ws,ws) is never materialised; anchor_self and the top-K are
    fused into the forward kernel (K <= 32; larger K runs the kernel in keep-all mode and sorts
    with a stable torch sort);
  * ties between *exactly equal* distances break towards the lower candidate index (the reference
    uses an unstable argsort, lib/stnls/nn/topk.py:209);
  * the other members of the reference's menus -- self_action "anchor_each" / "remove" / "remove_ref_frame" /
    "anchor_and_remove_ref_frame", topk_mode "each" / "none" -- run the forward kernel in keep-all mode and
    then follow impl/non_local_search.py:88-139 step by step (per-frame anchoring is the kernel in
    csrc/anchor.cu, the selections are stable torch sorts, as in lib/stnls/nn/topk.py);
  * pt != 1 and ws_interior != ws raise NotImplementedError.
"""
import torch as th

from .. import ops
from .._C import SearchOpts
from ..utils import extract_pairs

DIST_MENU = {"prod": (0, True), "l2": (1, False)}  # lib/stnls/search/utils.py:216-235


def topk_fallback(dists, inds, k, anchor, descending, extra=None):
    """Stable top-K over the last candidate axis; used only when K exceeds the fused limit.
    Mirrors lib/stnls/nn/topk.py:107-123,196-220."""
    def std(d, i, K, e):
        order = th.sort(d, dim=-1, descending=descending, stable=True)[1][..., :K]
        d = th.gather(d, -1, order)
        i = th.gather(i, -2, order[..., None].expand(*order.shape, 3))
        if e is not None:
            e = th.gather(e, -1, order)
        return d, i, e
    if anchor:
        d, i, e = std(dists[..., 1:], inds[..., 1:, :], k - 1, None if extra is None else extra[..., 1:])
        d = th.cat([dists[..., :1], d], -1)
        i = th.cat([inds[..., :1, :], i], -2)
        if e is not None:
            e = th.cat([extra[..., :1], e], -1)
        return d.contiguous(), i.contiguous(), (None if e is None else e.contiguous())
    d, i, e = std(dists, inds, k, extra)
    return d.contiguous(), i.contiguous(), (None if e is None else e.contiguous())


SELF_ACTIONS = (None, "anchor", "anchor_self", "anchor_each", "remove", "remove_ref_frame",
                "anchor_and_remove_ref_frame")  # impl/non_local_search.py:90-91
FUSED_ACTIONS = (None, "anchor", "anchor_self")


def topk_each(dists, inds, k, descending, anchor_self=False):
    """Per-frame-group top-K over (...,G,S) / (...,G,S,3); lib/stnls/nn/topk.py:226-269 (stable sort)."""
    def std(d, i, K):
        order = th.sort(d, dim=-1, descending=descending, stable=True)[1][..., :K]
        return th.gather(d, -1, order), th.gather(i, -2, order[..., None].expand(*order.shape, 3))
    if k <= 0:
        return dists, inds
    if anchor_self:
        d0, i0 = dists[..., :1], inds[..., :1, :]
        if k > 1 and dists.shape[-1] > 1:
            d, i = std(dists[..., 1:], inds[..., 1:, :], k - 1)
            return th.cat([d0, d], -1), th.cat([i0, i], -2)
        return d0, i0
    return std(dists, inds, k)


def select_unfused(dists, inds, flows_k, k, ws, wt, stride0, sizes, self_action, topk_mode, descending):
    """impl/non_local_search.py:88-139 on the keep-all output of the forward kernel (dists (B,HD,T,nH,nW,L) with
    the whole-list anchor swap already applied for "anchor" / "anchor_self" / "remove")."""
    qH, qW, kH, kW = sizes
    B, HD, T, nH, nW, L = dists.shape
    Q, S = T * nH * nW, ws * ws
    anchor = self_action is not None and "anchor" in self_action
    dists = dists.view(B, HD, Q, L // S, S)
    inds = inds.view(B, HD, Q, L // S, S, 3)
    if self_action == "anchor_each":
        ops.anchor_self_time(dists, inds, flows_k, wt, stride0, qH, qW, kH, kW)
    elif self_action == "remove":
        dists = dists.view(B, HD, Q, L)[..., 1:].contiguous()
        inds = inds.view(B, HD, Q, L, 3)[..., 1:, :].contiguous()
    elif self_action in ("remove_ref_frame", "anchor_and_remove_ref_frame"):
        assert wt > 0, "Cannot remove ref frame if not searching across time."
        dists = dists[..., 1:, :].contiguous()
        inds = inds[..., 1:, :, :].contiguous()
        if self_action == "anchor_and_remove_ref_frame":
            ops.anchor_self_time(dists, inds, flows_k, wt, stride0, qH, qW, kH, kW)
    if topk_mode == "all":
        if k > 0:
            dists, inds, _ = topk_fallback(dists.reshape(B, HD, Q, -1), inds.reshape(B, HD, Q, -1, 3), k, anchor,
                                           descending)
    elif topk_mode == "each":
        if dists.dim() != 5:
            raise ValueError("topk_mode='each' needs the per-frame candidate layout (not self_action='remove')")
        dists, inds = topk_each(dists, inds, k, descending, anchor_self=anchor)
    else:  # "none"
        assert k <= 0, "If topk_mode is 'none' then k must be <= 0"
    return dists.reshape(B, HD, T, nH, nW, -1).contiguous(), inds.reshape(B, HD, T, nH, nW, -1, 3).contiguous()


def check_common(pt, topk_mode, self_action, normalize_bwd, dist_type, itype, nls=False):
    if pt != 1:
        raise NotImplementedError("stnls_b200: pt must be 1 (the reference backward asserts the same)")
    if nls:
        if topk_mode not in ("all", "each", "none"):
            raise ValueError(f"Unknown topk_mode [{topk_mode}]")
        if self_action not in SELF_ACTIONS:
            raise ValueError(f"Uknown option for self_action [{self_action}]")
    elif topk_mode != "all":
        raise NotImplementedError("stnls_b200: only topk_mode='all' is implemented")
    elif self_action not in FUSED_ACTIONS:
        raise NotImplementedError(f"stnls_b200: self_action [{self_action}] is not implemented")
    if dist_type not in DIST_MENU:
        raise KeyError(dist_type)
    if itype not in ("int", "float"):
        raise ValueError(f"Uknown itype [{itype}]")


def normz_bwd(grad_vid0, grad_vid1, ps, stride0, stride1):
    """normalize_bwd=True (lib/stnls/search/shared.py:59-97): the video gradients are divided by the number of
    ps x ps patches, on stride-`stride0` (grad_vid0) / stride-`stride1` (grad_vid1) grids, that cover each pixel --
    the reference folds a tensor of ones; the counts are separable, so two device vectors from
    stnls_overlap_counts do the same."""
    assert int(stride1) == stride1, "stride1 must be an int."
    assert grad_vid0.dim() == 5, "normalize_bwd expects (B,T,C,H,W) videos (shared.py:68)"
    H, W = grad_vid0.shape[-2:]
    pad = (ps - 1) // 2
    for g, stride in ((grad_vid0, int(stride0)), (grad_vid1, int(stride1))):
        nH, nW = (H - 1) // stride + 1, (W - 1) // stride + 1
        # F.fold's own shape check: the ones tensor has nH*nW blocks, the padded plane (H+2*pad) must hold exactly that
        if (H + 2 * pad - ps) // stride + 1 != nH or (W + 2 * pad - ps) // stride + 1 != nW:
            raise RuntimeError("normalize_bwd: fold block count does not match (ps must be odd)")
        ch, _ = ops.overlap_counts(H, nH, ps, stride, 1, -pad, g.device)
        cw, _ = ops.overlap_counts(W, nW, ps, stride, 1, -pad, g.device)
        g /= (ch.float()[:, None] * cw.float()[None, :])


def shard_of(shard, T):
    """shard = None or dict(T=global clip length, tq0=first owned frame, tk0=first resident k frame)."""
    if shard is None:
        return T, 0, 0
    return shard["T"], shard["tq0"], shard["tk0"]


class NonLocalSearchFunction(th.autograd.Function):

    @staticmethod
    def forward(ctx, vid0, vid1, flows, ws, wt, ps, k, nheads=1, stride0=1, stride1=1, strideQ=None,
                dist_type="l2", dilation=1, pt=1, topk_mode="all", self_action=None, ws_interior=0,
                reflect_bounds=True, full_ws=True, use_adj=False, normalize_bwd=False, k_agg=-1,
                off_Hq=0, off_Wq=0, itype="float", shard=None):
        check_common(pt, topk_mode, self_action, normalize_bwd, dist_type, itype, nls=True)
        if ws_interior > 0 and ws_interior != ws:
            raise NotImplementedError("stnls_b200: ws_interior != ws is not implemented")
        in_dim, flow_ndim = vid0.dim(), flows.dim()
        q_cl = ops.pack_cl(vid0, nheads)
        k_cl = ops.pack_cl(vid1, nheads)
        B = vid0.shape[0]
        HD = nheads
        Tq, qH, qW = q_cl.shape[1:4]
        Tk, kH, kW = k_cl.shape[1:4]
        F = (vid0.shape[2] // nheads) if in_dim == 5 else vid0.shape[3]
        T, tq0, tk0 = shard_of(shard, Tq)
        W_t = 2 * wt + 1
        assert T >= W_t, f"Num Frames [{T}] must be >= Temporal Window [{W_t}]"
        if flows.dim() == 6:
            flows = flows[:, None]
        elif flows.dim() != 7:
            raise ValueError(f"Input flows are wrong dimension. Must be 6 or 7 but is [{flows.dim()}]")
        nH, nW = (kH - 1) // stride0 + 1, (kW - 1) // stride0 + 1
        assert flows.shape[-2] == nH and flows.shape[-1] == nW
        assert flows.shape[2] == Tq
        msg = "flows.shape: " + str(flows.shape) + " W_t: " + str(W_t)
        assert flows.shape[3] in [W_t - 1, W_t], msg
        dist_type_i, descending = DIST_MENU[dist_type]
        plain = topk_mode == "all" and self_action in FUSED_ACTIONS
        if not plain and shard is not None:
            raise NotImplementedError("stnls_b200: frame shards support topk_mode='all' with self_action None/'anchor'")
        anchor = self_action in ("anchor", "anchor_self", "remove")  # whole-list swap (nn.anchor_self), in the kernel
        if itype == "int":
            flows_k = flows.round().int() if flows.dtype != th.int32 else flows
            stride1_k = float(max(1, int(stride1)))
            strideQ_k = stride0 if strideQ is None else strideQ
        else:
            flows_k = flows.float()
            stride1_k = float(stride1)
            strideQ_k = stride0
        flows_k = flows_k.contiguous()
        L = W_t * ws * ws
        fused = plain and ((k <= 0) or (min(k, L) - (1 if anchor else 0) <= ops.MAX_FUSED_K))
        geom = ops.make_geom(B, HD, T, F, qH, qW, kH, kW, stride0, tq0, Tq, tk0, Tk)
        opts = SearchOpts(ws, wt, 0, 0, ps, k if fused else -1, stride0, strideQ_k, dilation, stride1_k,
                          int(reflect_bounds), int(full_ws), 0 if use_adj else -(ps // 2), off_Hq, off_Wq,
                          dist_type_i, int(itype == "int"), int(anchor), flows_k.shape[1],
                          W_t - flows_k.shape[3])
        dists, inds, _, _ = ops.search_forward(q_cl, k_cl, flows_k, geom, opts)
        if not plain:
            dists, inds = select_unfused(dists, inds, flows_k, k, ws, wt, stride0, (qH, qW, kH, kW), self_action,
                                         topk_mode, descending)
        elif not fused:
            dists, inds, _ = topk_fallback(dists, inds, k, anchor, descending)

        ctx.save_for_backward(q_cl, k_cl, inds, flows_k if itype == "float" else None)
        if itype == "int":
            ctx.mark_non_differentiable(inds)
        ctx.geom, ctx.opts = geom, opts
        ctx.vid_shapes = (tuple(vid0.shape), tuple(vid1.shape))
        ctx.nheads, ctx.k_agg, ctx.itype, ctx.flow_ndim = nheads, k_agg, itype, flow_ndim
        ctx.normz = (ps, stride0, stride1) if normalize_bwd else None
        return dists, inds

    @staticmethod
    def backward(ctx, grad_dists, grad_inds):
        q_cl, k_cl, inds, flows_k = ctx.saved_tensors
        flows_grad = ctx.itype == "float" and ctx.needs_input_grad[2]
        if ctx.k_agg > 0:  # impl/non_local_search.py:154-160
            grad_dists = grad_dists[..., :ctx.k_agg]
            inds = inds[..., :ctx.k_agg, :]
            if grad_inds is not None:
                grad_inds = grad_inds[..., :ctx.k_agg, :]
        if flows_grad and grad_inds is None:
            grad_inds = th.zeros_like(inds, dtype=th.float32)
        gq, gk, gflows = ops.search_backward(q_cl, k_cl, flows_k, grad_dists, grad_inds, inds, None, None,
                                             ctx.geom, ctx.opts, flows_grad)
        g0 = ops.unpack_cl(gq, ctx.vid_shapes[0], ctx.nheads)
        g1 = ops.unpack_cl(gk, ctx.vid_shapes[1], ctx.nheads)
        if ctx.normz is not None:
            normz_bwd(g0, g1, *ctx.normz)
        if gflows is not None and ctx.flow_ndim == 6:
            gflows = gflows[:, 0].contiguous()
        return (g0, g1, gflows) + (None,) * 23


class NonLocalSearch(th.nn.Module):
    """Constructor signature identical to lib/stnls/search/non_local_search.py:124-129."""

    def __init__(self, ws, wt, ps=1, k=-1, nheads=1, stride0=1, stride1=1, dist_type="l2",
                 dilation=1, pt=1, self_action=None, topk_mode="all", ws_interior=0,
                 reflect_bounds=True, full_ws=True, use_adj=False, normalize_bwd=False, k_agg=-1,
                 off_Hq=0, off_Wq=0, strideQ=None, itype="float"):
        super().__init__()
        self.ws, self.wt, self.ps, self.k, self.nheads = ws, wt, ps, k, nheads
        self.dist_type, self.stride0, self.stride1, self.strideQ = dist_type, stride0, stride1, strideQ
        self.dilation, self.ws_interior, self.pt = dilation, ws_interior, pt
        self.off_Hq, self.off_Wq, self.itype = off_Hq, off_Wq, itype
        self.reflect_bounds, self.full_ws = reflect_bounds, full_ws
        self.topk_mode, self.self_action, self.use_adj = topk_mode, self_action, use_adj
        self.normalize_bwd, self.k_agg = normalize_bwd, k_agg

    def forward(self, *args, shard=None):
        assert self.ws > 0, "Must have nonzero spatial search window"
        assert self.wt >= 0, "Must have nonnegative time search window"
        vid0, vid1 = args[:2]
        if len(args) == 4:
            from ..nn import search_flow  # lib/stnls/search/non_local_search.py:173-175
            flows = search_flow(args[2], args[3], self.wt, self.stride0)
        elif len(args) == 3:
            flows = args[2]
        else:
            W_t = 2 * self.wt + 1
            B = vid0.shape[0]
            T = vid0.shape[1] if vid0.dim() == 5 else vid0.shape[2]
            H, W = vid0.shape[-2:]
            nH, nW = (H - 1) // self.stride0 + 1, (W - 1) // self.stride0 + 1
            flows = th.zeros((B, self.nheads, T, W_t, 2, nH, nW), device=vid0.device)
        return NonLocalSearchFunction.apply(vid0, vid1, flows, self.ws, self.wt, self.ps, self.k, self.nheads,
                                            self.stride0, self.stride1, self.strideQ, self.dist_type,
                                            self.dilation, self.pt, self.topk_mode, self.self_action,
                                            self.ws_interior, self.reflect_bounds, self.full_ws, self.use_adj,
                                            self.normalize_bwd, self.k_agg, self.off_Hq, self.off_Wq,
                                            self.itype, shard)

    def flops(self, T, F, H, W):
        return 0

    def radius(self, H, W):
        return self.ws


def extract_config(cfg, restrict=True):
    """lib/stnls/search/non_local_search.py:254-262"""
    pairs = {"ws": -1, "wt": -1, "ps": 1, "k": -1, "nheads": 1, "dist_type": "l2", "stride0": 1, "stride1": 1,
             "dilation": 1, "pt": 1, "ws_interior": 0, "reflect_bounds": True, "full_ws": True,
             "self_action": None, "use_adj": False, "normalize_bwd": False, "k_agg": -1, "topk_mode": "all",
             "off_Hq": 0, "off_Wq": 0, "strideQ": None, "itype": "float"}
    return extract_pairs(cfg, pairs, restrict=restrict)


def init(cfg):
    """lib/stnls/search/non_local_search.py:265-276"""
    cfg = extract_config(cfg, False)
    return NonLocalSearch(cfg.ws, cfg.wt, cfg.ps, cfg.k, nheads=cfg.nheads, stride0=cfg.stride0,
                          stride1=cfg.stride1, dist_type=cfg.dist_type, dilation=cfg.dilation, pt=cfg.pt,
                          self_action=cfg.self_action, topk_mode=cfg.topk_mode, ws_interior=cfg.ws_interior,
                          reflect_bounds=cfg.reflect_bounds, full_ws=cfg.full_ws, use_adj=cfg.use_adj,
                          normalize_bwd=cfg.normalize_bwd, k_agg=cfg.k_agg, off_Hq=cfg.off_Hq,
                          off_Wq=cfg.off_Wq, strideQ=cfg.strideQ, itype=cfg.itype)
