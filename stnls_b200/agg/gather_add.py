"""stnls_b200.agg.WeightedPatchSum / NonLocalGatherAdd.

`WeightedPatchSum(ps, stride0, ..., itype)` is the name the reference README uses (README.md:54);
in the reference snapshot its implementation is `NonLocalGatherAdd(ps, strideIn, strideOut, ...)`
(lib/stnls/agg/gather_add.py:34-221, kernels lib/csrc/agg/gather_add_kernel.cu) -- SURVEY.md 0.2.

  out[b,hd,t,c,h,w] = 1/count[h,w] * sum_{q,k,(pi,pj): ref_q+p == (h,w)} weights[q,k] * v[t_nl, c, nl_{q,k}+p]

Returns (B,HD,T,F_head,H,W) like the reference (gather_add.py:78,114).
"""
import torch as th

from .. import ops
from .._C import AggOpts
from ..search.non_local_search import shard_of
from ..utils import extract_pairs


def get_inds(inds, itype):
    """lib/stnls/agg/gather_add.py:25-32"""
    inds = inds.contiguous()
    if itype == "int" and th.is_floating_point(inds):
        return inds.round().int()
    elif itype in ["float", "2d", "3d"] and not th.is_floating_point(inds):
        return inds.float()
    return inds


class NonLocalGatherAddFunction(th.autograd.Function):

    @staticmethod
    def forward(ctx, vid, weights, flows, ps, strideIn, strideOut, outH=0, outW=0, pt=1, dilation=1,
                reflect_bounds=True, use_adj=False, itype="float", shard=None):
        if pt != 1:
            raise NotImplementedError("stnls_b200: pt must be 1")
        if strideIn != strideOut:
            raise NotImplementedError("stnls_b200: strideIn must equal strideOut (the reference kernel reads the "
                                      "output size from the input video, gather_add_kernel.cu:32-33)")
        B, HD, Tq, nH, nW, K = weights.shape
        vid_in_dim = vid.dim()
        if vid.dim() == 5 and (vid.shape[2] % HD) != 0:
            raise NotImplementedError("stnls_b200: channels must be divisible by the number of heads")
        v_cl = ops.pack_cl(vid, HD)
        Tk, H, W = v_cl.shape[1:4]
        F = (vid.shape[2] // HD) if vid.dim() == 5 else vid.shape[3]
        if (outH not in (0, H)) or (outW not in (0, W)):
            raise NotImplementedError("stnls_b200: output size must equal the input size")
        if flows.dim() == 6:
            flows = flows[:, None]
        flows_k = get_inds(flows, itype)
        assert nH == (H - 1) // strideIn + 1 and nW == (W - 1) // strideIn + 1
        assert tuple(flows_k.shape) == (B, HD, Tq, nH, nW, K, 3), "flows must be (B,HD,T,nH,nW,K,3)"
        T, tq0, tk0 = shard_of(shard, Tq)
        patch_offset = 0 if use_adj else -(ps // 2)
        ch, mh = ops.overlap_counts(H, nH, ps, strideIn, dilation, patch_offset, vid.device)
        cw, mw = ops.overlap_counts(W, nW, ps, strideIn, dilation, patch_offset, vid.device)
        assert mh > 0 and mw > 0, "every output pixel must be covered by a patch (reference: counts>1e-3)"
        geom = ops.make_geom(B, HD, T, F, H, W, H, W, strideIn, tq0, Tq, tk0, Tk)
        opts = AggOpts(ps, K, strideIn, dilation, int(reflect_bounds), patch_offset, int(flows_k.dtype == th.int32))
        weights_k = weights.contiguous().float()
        out_cl = ops.wpsum_forward(v_cl, weights_k, flows_k, (ch, cw), geom, opts)
        out = ops.unpack_cl(out_cl, (B, HD, Tq, F, H, W), HD)
        ctx.save_for_backward(weights_k, flows_k, v_cl, ch, cw)
        ctx.geom, ctx.opts = geom, opts
        ctx.vid_shape, ctx.vid_in_dim, ctx.HD, ctx.itype = tuple(vid.shape), vid_in_dim, HD, itype
        ctx.wshape = tuple(weights.shape)
        return out

    @staticmethod
    def backward(ctx, grad_out):
        weights, flows, v_cl, ch, cw = ctx.saved_tensors
        gout_cl = ops.pack_cl(grad_out, ctx.HD, counts=(ch, cw))  # fused grad_out / counts
        gv_cl, gw, gf = ops.wpsum_backward(gout_cl, v_cl, weights, flows, ctx.geom, ctx.opts)
        gvid = ops.unpack_cl(gv_cl, ctx.vid_shape, ctx.HD)
        gw = gw.reshape(ctx.wshape)
        if ctx.itype == "float" and gf is not None:
            gf = gf.reshape(ctx.wshape + (3,))
        else:
            gf = None
        return (gvid, gw, gf) + (None,) * 11


class NonLocalGatherAdd(th.nn.Module):
    """Constructor identical to lib/stnls/agg/gather_add.py:193-194."""

    def __init__(self, ps, strideIn, strideOut, outH=0, outW=0, pt=1, dilation=1, reflect_bounds=True,
                 use_adj=False, itype="float"):
        super().__init__()
        self.ps, self.strideIn, self.strideOut, self.outH, self.outW = ps, strideIn, strideOut, outH, outW
        self.pt, self.dilation, self.reflect_bounds, self.use_adj, self.itype = pt, dilation, reflect_bounds, use_adj, itype

    def forward(self, vid, weights, flows, shard=None):
        return NonLocalGatherAddFunction.apply(vid, weights, flows, self.ps, self.strideIn, self.strideOut,
                                               self.outH, self.outW, self.pt, self.dilation, self.reflect_bounds,
                                               self.use_adj, self.itype, shard)

    def flops(self, nrefs, chnls_per_head, nheads, k):
        return 2 * (chnls_per_head * self.ps * self.ps * self.pt) * k * nrefs * nheads


class WeightedPatchSum(NonLocalGatherAdd):
    """README form (README.md:54): WeightedPatchSum(ps, stride0, itype) == NonLocalGatherAdd(ps, stride0, stride0)."""

    def __init__(self, ps, stride0=1, pt=1, dilation=1, reflect_bounds=True, use_adj=False, itype="float"):
        super().__init__(ps, stride0, stride0, pt=pt, dilation=dilation, reflect_bounds=reflect_bounds,
                         use_adj=use_adj, itype=itype)
        self.stride0 = stride0


def extract_config(cfg, restrict=True):
    """lib/stnls/agg/gather_add.py:243-247"""
    pairs = {"ps": 3, "strideIn": 1, "strideOut": 1, "outH": 0, "outW": 0, "pt": 1, "dilation": 1,
             "reflect_bounds": True, "use_adj": False, "itype": "float"}
    return extract_pairs(cfg, pairs, restrict=restrict)


def init(cfg):
    cfg = extract_config(cfg, False)
    return NonLocalGatherAdd(cfg.ps, cfg.strideIn, cfg.strideOut, outH=cfg.outH, outW=cfg.outW, pt=cfg.pt,
                             dilation=cfg.dilation, reflect_bounds=cfg.reflect_bounds, use_adj=cfg.use_adj,
                             itype=cfg.itype)
