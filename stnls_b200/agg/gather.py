"""stnls_b200.agg.NonLocalGather -- K-stacked aggregation.

Mirrors lib/stnls/agg/gather.py:70-276 (kernels lib/csrc/agg/gather_{int,bilin2d}_kernel.cu):

  stack[b,hd,k,t,c,h,w] = 1/count[h,w] * sum_{q,(pi,pj): ref_q+p == (h,w)} weights[q,k] * v[t_nl, c, nl_{q,k}+p]

i.e. WeightedPatchSum without the sum over the K neighbours; `stack.sum(2)` equals the weighted
patch sum (the reference's own ground truth, tests/agg/test_gather_add.py:106-128).  The device code is
the weighted-patch-sum gather kernel in "stack" mode (csrc/wpsum.cu): slab n*K + k of the
channels-last output holds neighbour k, so no atomics and no separate divide pass are needed.
"""
import torch as th

from .. import ops
from .._C import AggOpts
from ..search.non_local_search import shard_of
from ..utils import extract_pairs
from .gather_add import get_inds


class non_local_gather(th.autograd.Function):
    """Argument order of lib/stnls/agg/gather.py:79-82."""

    @staticmethod
    def forward(ctx, vid, weights, inds, ps=7, stride0=4, pt=1, reflect_bounds=True, dilation=1,
                use_adj=False, itype="int", shard=None):
        if pt != 1:
            raise NotImplementedError("stnls_b200: pt must be 1")
        if inds.dim() == 6:
            inds = inds[:, None]
        B, HD = inds.shape[:2]
        K = inds.shape[-2]
        if vid.dim() == 5 and (vid.shape[2] % HD) != 0:
            raise NotImplementedError("stnls_b200: channels must be divisible by the number of heads")
        v_cl = ops.pack_cl(vid, HD)
        Tk, H, W = v_cl.shape[1:4]
        F = (vid.shape[2] // HD) if vid.dim() == 5 else vid.shape[3]
        nH, nW = (H - 1) // stride0 + 1, (W - 1) // stride0 + 1
        wshape = tuple(weights.shape)
        Tq = weights.numel() // (B * HD * nH * nW * K)
        weights_k = weights.reshape(B, HD, Tq, nH, nW, K).contiguous().float()
        inds_k = get_inds(inds.reshape(B, HD, Tq, nH, nW, K, 3), itype)
        T, tq0, tk0 = shard_of(shard, Tq)
        patch_offset = 0 if use_adj else -(ps // 2)
        ch, mh = ops.overlap_counts(H, nH, ps, stride0, dilation, patch_offset, vid.device)
        cw, mw = ops.overlap_counts(W, nW, ps, stride0, dilation, patch_offset, vid.device)
        assert mh > 0 and mw > 0, "every output pixel must be covered by a patch (reference: counts > 0)"
        geom = ops.make_geom(B, HD, T, F, H, W, H, W, stride0, tq0, Tq, tk0, Tk)
        opts = AggOpts(ps, K, stride0, dilation, int(reflect_bounds), patch_offset, int(inds_k.dtype == th.int32))
        stack_cl = ops.gather_forward(v_cl, weights_k, inds_k, (ch, cw), geom, opts)
        # slab n*K + k  ->  (B, HD, K, T, F, H, W): K plays the part of extra heads of the layout kernel
        stack = ops.unpack_cl(stack_cl, (B, HD * K, Tq, F, H, W), HD * K).view(B, HD, K, Tq, F, H, W)
        ctx.save_for_backward(weights_k, inds_k, v_cl, ch, cw)
        ctx.geom, ctx.opts = geom, opts
        ctx.vid_shape, ctx.HD, ctx.K, ctx.itype, ctx.wshape = tuple(vid.shape), HD, K, itype, wshape
        return stack

    @staticmethod
    def backward(ctx, grad_stack):
        weights, inds, v_cl, ch, cw = ctx.saved_tensors
        B, HD, K, Tq, F, H, W = grad_stack.shape
        gs = grad_stack.contiguous().view(B, HD * K, Tq, F, H, W)
        gstack_cl = ops.pack_cl(gs, HD * K, counts=(ch, cw))  # fused grad_stack / counts (gather.py:231)
        gv_cl, gw, gi = ops.gather_backward(gstack_cl, v_cl, weights, inds, ctx.geom, ctx.opts)
        gvid = ops.unpack_cl(gv_cl, ctx.vid_shape, ctx.HD)
        gw = gw.reshape(ctx.wshape)
        gi = gi.reshape(ctx.wshape + (3,)) if (ctx.itype != "int" and gi is not None) else None
        return (gvid, gw, gi) + (None,) * 8


class NonLocalGather(th.nn.Module):
    """Constructor identical to lib/stnls/agg/gather.py:258-265."""

    def __init__(self, ps=1, stride0=1, pt=1, dilation=1, reflect_bounds=True, use_adj=False, itype="float"):
        super().__init__()
        self.ps, self.stride0, self.pt, self.dilation = ps, stride0, pt, dilation
        self.reflect_bounds, self.use_adj, self.itype = reflect_bounds, use_adj, itype

    def forward(self, vid, weights, inds, shard=None):
        return non_local_gather.apply(vid, weights, inds, self.ps, self.stride0, self.pt, self.reflect_bounds,
                                      self.dilation, self.use_adj, self.itype, shard)

    def flops(self, nrefs, chnls_per_head, nheads, k):
        return 2 * (chnls_per_head * self.ps * self.ps * self.pt) * k * nrefs * nheads


def extract_config(cfg, restrict=True):
    """lib/stnls/agg/gather.py:296-300"""
    pairs = {"ps": 3, "stride0": 1, "pt": 1, "dilation": 1, "reflect_bounds": True, "use_adj": False,
             "itype": "float"}
    return extract_pairs(cfg, pairs, restrict=restrict)


def init(cfg):
    cfg = extract_config(cfg, False)
    return NonLocalGather(cfg.ps, cfg.stride0, pt=cfg.pt, dilation=cfg.dilation,
                          reflect_bounds=cfg.reflect_bounds, use_adj=cfg.use_adj, itype=cfg.itype)
