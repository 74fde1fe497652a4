"""flows = stnls_b200.nn.search_flow(fflow, bflow, wt, stride0)      

Mirrors lib/stnls/nn/search_flow.py:24-96: optical flows between consecutive frames (B,T,2,H,W) are
chained through the temporal window into per-query accumulated offsets (B,T,W_t-1,2,nH,nW),
W_t = min(2*wt+1, T) -- the tensor NonLocalSearch takes as `flows`.  Forward and backward run
hand-written kernels behind stnls_search_flow_forward / _backward (csrc/search_flow.cu).
"""
import torch as th

from .. import ops


def init():
    return run


def empty_flows(fflow, wt, stride0):
    """lib/stnls/nn/search_flow.py:36-43"""
    B, T, _, H, W = fflow.shape
    nH, nW = (H - 1) // stride0 + 1, (W - 1) // stride0 + 1
    W_t = min(2 * wt + 1, T)
    return th.zeros((B, T, W_t - 1, 2, nH, nW), device=fflow.device, dtype=fflow.dtype)


def run(fflow, bflow, wt, stride0=1):
    if wt > 0:
        return search_flow_th.apply(fflow.contiguous(), bflow.contiguous(), wt, stride0)
    return empty_flows(fflow, wt, stride0)


class search_flow_th(th.autograd.Function):

    @staticmethod
    def forward(ctx, fflow, bflow, wt, stride0):
        flows = ops.search_flow_forward(fflow, bflow, wt, stride0)
        ctx.save_for_backward(fflow, bflow, flows)
        ctx.wt, ctx.stride0 = wt, stride0
        return flows

    @staticmethod
    def backward(ctx, grad_flows):
        fflow, bflow, flows = ctx.saved_tensors
        gf, gb = ops.search_flow_backward(grad_flows, fflow, bflow, flows, ctx.wt, ctx.stride0)
        return gf, gb, None, None
