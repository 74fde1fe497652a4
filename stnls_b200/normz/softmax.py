"""SoftmaxNormalize(scale, drop_rate, dist_type, k_agg): weights = softmax(+-scale * dists[..., :k_agg], -1).

Same constructor and call contract as lib/stnls/normz/softmax.py:12-47 -- `normz(dists, inds)` returns
(weights, inds[..., :k_agg, :]) -- but slice, sign flip, scale and softmax run as ONE hand-written kernel per
direction (csrc/normz.cu) instead of four torch ops.  `softmax_scaled(dists, scale)` is the README form
`th.nn.functional.softmax(10*dists,-1)` (README.md:51).
"""
import torch as th

from .. import ops
from ..utils import extract_pairs


class _SoftmaxScaled(th.autograd.Function):

    @staticmethod
    def forward(ctx, dists, scale, K):
        dists = dists.contiguous().float()
        w = ops.softmax_forward(dists, K, scale)
        ctx.save_for_backward(w)
        ctx.scale, ctx.Kin = scale, dists.shape[-1]
        return w

    @staticmethod
    def backward(ctx, grad_w):
        (w,) = ctx.saved_tensors
        return ops.softmax_backward(grad_w, w, ctx.Kin, ctx.scale), None, None


def softmax_scaled(dists, scale, k_agg=-1):
    """softmax(scale * dists[..., :k_agg], -1); k_agg <= 0 keeps every neighbour."""
    Kin = dists.shape[-1]
    K = min(k_agg, Kin) if k_agg > 0 else Kin
    if K > ops.MAX_SOFTMAX_K:
        return th.softmax(scale * dists[..., :K], -1)   # beyond the register budget of the fused kernel
    return _SoftmaxScaled.apply(dists, float(scale), K)


class SoftmaxNormalize(th.nn.Module):

    def __init__(self, scale, drop_rate=0., dist_type="l2", k_agg=-1):
        super().__init__()
        self.scale, self.drop_rate, self.dist_type, self.k_agg = scale, drop_rate, dist_type, k_agg
        self.drop = th.nn.Dropout(drop_rate)

    def __call__(self, dists, inds):
        if self.k_agg > 0:
            inds = inds[..., :self.k_agg, :].contiguous()
        else:
            inds = inds.contiguous()
        sign = -1.0 if self.dist_type == "l2" else 1.0
        weights = softmax_scaled(dists, sign * self.scale, self.k_agg)
        if self.drop_rate > 0:
            weights = self.drop(weights)
        return weights, inds


def default_pairs():
    """lib/stnls/normz/api.py:42-48"""
    return {"normz_scale": 10, "normz_name": "softmax", "normz_drop_rate": 0., "dist_type": "l2", "k_agg": -1}


def extract_config(cfg, restrict=True):
    return extract_pairs(cfg, default_pairs(), restrict=restrict)


def init(cfg):
    """lib/stnls/normz/api.py:24-40 + softmax.py:7-8"""
    cfg = extract_config(cfg, False)
    if cfg.normz_name != "softmax":
        raise NotImplementedError(f"stnls_b200: normalisation [{cfg.normz_name}] (the reference menu has softmax only)")
    return SoftmaxNormalize(cfg.normz_scale, cfg.normz_drop_rate, cfg.dist_type, cfg.k_agg)
