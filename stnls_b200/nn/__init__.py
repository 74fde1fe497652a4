"""stnls_b200.nn -- the callers either side of the hot path: search_flow (N1), NonLocalAttention (N4)."""
from . import search_flow as _search_flow_mod  # noqa: F401
from .search_flow import run as search_flow  # noqa: F401
from .search_flow import search_flow_th  # noqa: F401
from . import non_local_attn  # noqa: F401
from .non_local_attn import NonLocalAttention, NonLocalAttentionStack  # noqa: F401


def anchor_self_time(dists, inds, flows, wt, stride0, qH, qW, kH, kW):
    """stnls.nn.anchor_self_time (lib/stnls/nn/anchor_self.py:69-79): in place on dists (B,HD,Q,W_t,ws,ws) /
    inds (...,3) (or the flattened (B,HD,Q,W_t,ws*ws) form), flows (B,HD_f,T,W_t[-1],2,nH,nW)."""
    from .. import ops
    B, HD, Q, G = dists.shape[:4]
    ops.anchor_self_time(dists.view(B, HD, Q, G, -1), inds.view(B, HD, Q, G, -1, 3), flows.contiguous(), wt, stride0,
                         qH, qW, kH, kW)


def topk_each(dists, inds, k, descending, anchor_self=False):
    """stnls.nn.topk_each (lib/stnls/nn/topk.py:226-269), stable."""
    from ..search.non_local_search import topk_each as impl
    return impl(dists, inds, k, descending, anchor_self=anchor_self)


def topk(dists, inds, k, dim=1, anchor=False, descending=True, unique=False, return_order=False):
    """stnls.nn.topk (lib/stnls/nn/topk.py:65-95) for the layouts the searches produce: the candidate axis `dim` of
    dists is moved last, selected with a stable sort (anchored: slot 0 stays), and moved back."""
    if unique:
        raise NotImplementedError("stnls_b200: unique top-k is outside the hot path")
    if not (k > 0):
        return (dists, inds, None) if return_order else (dists, inds)
    from ..search.non_local_search import topk_fallback
    import torch as th
    d = dists.movedim(dim, -1)
    i = inds.movedim(dim, -2)
    pos = th.arange(d.shape[-1], device=d.device).expand(d.shape)
    dk, ik, order = topk_fallback(d.contiguous(), i.contiguous(), k, anchor, descending, pos.contiguous())
    dk, ik, order = dk.movedim(-1, dim), ik.movedim(-2, dim), order.movedim(-1, dim)
    return (dk, ik, order) if return_order else (dk, ik)
