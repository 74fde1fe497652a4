"""NonLocalAttention block: the caller layer around the hot path.

Mirrors lib/stnls/nn/non_local_attn.py:52-262 -- 1x1 ConvQKV -> search -> normalise -> aggregate -> 1x1
projection -- built from stnls_b200.search.init / normz.init / agg.init, i.e. every heavy step runs the
hand-written kernels behind the C ABI; only the two 1x1 convolutions are cuDNN library calls, as in the
reference.  The reference's timers (ExpTimer) are dropped; `state` handling for the refine search is kept.
"""
import torch as th
import torch.nn as nn
import torch.nn.functional as F

from .. import agg as _agg
from .. import normz as _normz
from .. import search as _search
from ..utils import AttrDict, extract_pairs


def default_pairs():
    """lib/stnls/nn/non_local_attn.py:37-46"""
    return {"nheads": 1, "inner_mult": 1, "embed_dim": 16, "qk_frac": 1., "qkv_bias": True, "qkv_ngroups": 1,
            "use_attn_projection": True, "drop_rate_proj": 0., "attn_timer": False, "use_attn_flow": True,
            "use_norm_layer": False}


def extract_config(cfg, restrict=True):
    return extract_pairs(cfg, default_pairs(), restrict=restrict)


def rescale_flows(flows_og, H, W):
    """lib/stnls/nn/utils.py:6-44: resize (fflow, bflow) to the feature resolution, scaling the vectors."""
    if flows_og is None:
        return None
    fshape = list(flows_og.fflow.shape)
    _H, _W = fshape[-2:]
    if _H == H:
        return flows_og
    fshape[-2], fshape[-1] = H, W
    scale = th.tensor([_W / W, _H / H], device=flows_og.fflow.device).view(1, 2, 1, 1)
    out = AttrDict()
    for name in ("fflow", "bflow"):
        f = getattr(flows_og, name).reshape(-1, 2, _H, _W)
        out[name] = F.interpolate(f / scale, size=(H, W), mode="bilinear", align_corners=True).view(*fshape)
    return out


class ConvQKV(nn.Module):
    """lib/stnls/nn/non_local_attn.py:265-300"""

    def __init__(self, input_dim, heads=8, dim_head=64, qk_frac=1., kernel_size=1, q_stride=1, k_stride=1,
                 v_stride=1, ngroups=1, bias=True):
        super().__init__()
        inner_dim = dim_head * heads
        inner_dim_qk = max(int(qk_frac * dim_head), 1) * heads
        self.heads = heads
        pad = (kernel_size - q_stride) // 2
        kw = dict(kernel_size=kernel_size, padding=pad, bias=bias, groups=ngroups, padding_mode="reflect")
        self.to_q = nn.Conv2d(input_dim, inner_dim_qk, stride=q_stride, **kw)
        self.to_k = nn.Conv2d(input_dim, inner_dim_qk, stride=k_stride, **kw)
        self.to_v = nn.Conv2d(input_dim, inner_dim, stride=v_stride, **kw)

    def forward(self, x, attn_kv=None):
        attn_kv = x if attn_kv is None else attn_kv
        return self.to_q(x), self.to_k(attn_kv), self.to_v(attn_kv)


class LayerNorm2D(nn.LayerNorm):
    """channel LayerNorm of a (B,T,C,H,W) video (lib/stnls/nn/non_local_attn.py:386-399)"""

    def __init__(self, num_channels, eps=1e-6, affine=True):
        super().__init__(num_channels, eps=eps, elementwise_affine=affine)

    def forward(self, vid):
        B, T, C, H, W = vid.shape
        x = vid.reshape(B * T, C, H, W).permute(0, 2, 3, 1)
        x = F.layer_norm(x, self.normalized_shape, self.weight, self.bias, self.eps).permute(0, 3, 1, 2)
        return x.reshape(B, T, C, H, W).contiguous()


class NonLocalAttention(nn.Module):

    def __init__(self, attn_cfg, search_cfg, normz_cfg, agg_cfg):
        super().__init__()
        attn_cfg = extract_pairs(attn_cfg, default_pairs(), restrict=False)
        search_cfg, normz_cfg, agg_cfg = AttrDict(search_cfg), AttrDict(normz_cfg), AttrDict(agg_cfg)
        nheads = attn_cfg.nheads
        inner_mult = attn_cfg.inner_mult
        embed_dim = attn_cfg.embed_dim * inner_mult
        io_dim = attn_cfg.embed_dim * nheads
        self.dim = io_dim
        self.attn_cfg, self.search_cfg, self.normz_cfg, self.agg_cfg = attn_cfg, search_cfg, normz_cfg, agg_cfg
        self.search = _search.init(search_cfg)
        self.normz = _normz.init(normz_cfg)
        self.agg = _agg.init(agg_cfg)
        self.use_flow = attn_cfg.use_attn_flow
        self.use_state_update = search_cfg.get("use_state_update", False)
        self.search_name = search_cfg.get("search_name", "nls")
        self.stride0 = search_cfg.get("stride0", 1)
        self.qkv = ConvQKV(io_dim, nheads, embed_dim, attn_cfg.qk_frac, bias=attn_cfg.qkv_bias,
                           ngroups=attn_cfg.qkv_ngroups)
        if attn_cfg.use_attn_projection:
            self.proj = nn.Conv2d(io_dim * inner_mult, io_dim, (1, 1), stride=1, padding="same", groups=1)
            self.proj_drop = nn.Dropout(attn_cfg.drop_rate_proj)
        else:
            self.proj, self.proj_drop = nn.Identity(), nn.Identity()
        self.norm_layer = LayerNorm2D(io_dim) if attn_cfg.use_norm_layer else nn.Identity()

    def forward(self, vid, flows=None, state=None):
        B, T, C, H, W = vid.shape
        if self.use_flow:
            flows = rescale_flows(flows, H, W)
        vid = self.norm_layer(vid)
        q_vid, k_vid, v_vid = self.get_qkv(vid)
        dists, inds = self.run_search(q_vid, k_vid, flows, state)
        weights, inds = self.normz(dists, inds)
        out = self.agg(v_vid, weights, inds)                    # (B,HD,T,F_head,H,W)
        return self.run_projection(out)

    def get_qkv(self, vid):
        B, T, C, H, W = vid.shape
        q, k, v = self.qkv(vid.reshape(B * T, C, H, W), None)
        return q.view(B, T, -1, H, W), k.view(B, T, -1, H, W), v.view(B, T, -1, H, W)

    def run_search(self, q_vid, k_vid, flows, state):
        if self.search_name in ("refine", "refinement"):
            dists, inds = self.search(q_vid, k_vid, state[0])
        elif flows is None:
            dists, inds = self.search(q_vid, k_vid)
        else:
            dists, inds = self.search(q_vid, k_vid, flows.fflow, flows.bflow)
        if self.use_state_update and state is not None:
            state[1] = state[0]
            state[0] = inds.detach()
        return dists, inds

    def run_projection(self, vid):
        if vid.dim() == 6:  # (B,HD,T,F,H,W) -> (B,T,HD*F,H,W), the layout the reference's agg returns to the block
            B, HD, T, Fh, H, W = vid.shape
            vid = vid.permute(0, 2, 1, 3, 4, 5).reshape(B, T, HD * Fh, H, W)
        B, T, C, H, W = vid.shape
        x = self.proj_drop(self.proj(vid.reshape(B * T, C, H, W)))
        return x.view(B, T, -1, H, W)


class NonLocalAttentionStack(NonLocalAttention):
    """lib/stnls/nn/non_local_attn_stack.py:53-147: the same block without the output projection -- forward returns
    what the aggregator returns (with agg_name="gather" / "nlstack": the K-stacked video (B,HD,K,T,F_head,H,W) of
    NonLocalGather), so that a caller can fuse the K neighbours itself."""

    def __init__(self, attn_cfg, search_cfg, normz_cfg, agg_cfg):
        attn_cfg = dict(attn_cfg)
        attn_cfg["use_attn_projection"] = False  # the stack block has no projection (non_local_attn_stack.py:107-117)
        super().__init__(attn_cfg, search_cfg, normz_cfg, agg_cfg)

    def forward(self, vid, flows=None, state=None):
        B, T, C, H, W = vid.shape
        if self.use_flow:
            flows = rescale_flows(flows, H, W)
        vid = self.norm_layer(vid)
        q_vid, k_vid, v_vid = self.get_qkv(vid)
        dists, inds = self.run_search(q_vid, k_vid, flows, state)
        weights, inds = self.normz(dists, inds)
        return self.agg(v_vid, weights, inds)


def init(cfg):
    """Split one flat config into the four the block takes (lib/stnls/nn/non_local_attn.py + api usage)."""
    cfg = AttrDict(cfg)
    return NonLocalAttention(extract_config(cfg, False), cfg, cfg, cfg)
