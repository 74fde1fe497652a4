"""GPU: NonLocalAttention block == the same pipeline assembled from the CPU oracle
(1x1 convs in torch, search_flow / search / weighted patch sum from oracle/), forward 1e-3, and gradients reach
the convolution weights and the input video."""
import pytest
import torch as th
import torch.nn.functional as F

import stnls_b200
from stnls_b200.utils import AttrDict
from util import assert_close

pytestmark = pytest.mark.gpu


def make_block(cuda, search_name="nls", itype="float"):
    cfg = AttrDict(nheads=2, embed_dim=4, search_name=search_name, ws=3, wt=1, ps=3, k=5, stride0=1, stride1=1.0,
                   dist_type="l2", self_action="anchor", itype=itype, normz_scale=10, k_agg=-1, agg_name="wpsum",
                   strideIn=1, strideOut=1)
    th.manual_seed(3)
    return stnls_b200.nn.non_local_attn.init(cfg).to(cuda), cfg


def test_attention_matches_oracle_pipeline(cuda, oracle):
    block, cfg = make_block(cuda)
    g = th.Generator().manual_seed(4)
    B, T, C, H, W = 1, 3, 8, 16, 16
    vid = th.randn(B, T, C, H, W, generator=g)
    flows = AttrDict(fflow=th.randn(B, T, 2, H, W, generator=g), bflow=th.randn(B, T, 2, H, W, generator=g))
    vg = vid.to(cuda).requires_grad_(True)
    out = block(vg, AttrDict(fflow=flows.fflow.to(cuda), bflow=flows.bflow.to(cuda)))
    assert tuple(out.shape) == (B, T, C, H, W)
    # the same pipeline from the oracle
    cpu = lambda t: t.detach().cpu()
    x = vid.reshape(B * T, C, H, W)
    q = F.conv2d(x, cpu(block.qkv.to_q.weight), cpu(block.qkv.to_q.bias)).view(B, T, -1, H, W)
    k = F.conv2d(x, cpu(block.qkv.to_k.weight), cpu(block.qkv.to_k.bias)).view(B, T, -1, H, W)
    v = F.conv2d(x, cpu(block.qkv.to_v.weight), cpu(block.qkv.to_v.bias)).view(B, T, -1, H, W)
    fl = oracle.search_flow_forward(flows.fflow, flows.bflow, cfg.wt, cfg.stride0)
    d, i = oracle.nls_forward(q, k, fl, cfg.ws, cfg.wt, cfg.ps, cfg.k, nheads=cfg.nheads, stride0=1, stride1=1.0,
                              dist_type="l2", self_action="anchor", itype="float")
    w = th.softmax(-10.0 * d, -1)
    o, _ = oracle.gather_add_forward(v, w, i, cfg.ps, 1, itype="float")
    o = o.permute(0, 2, 1, 3, 4, 5).reshape(B * T, -1, H, W)
    ref = F.conv2d(o, cpu(block.proj.weight), cpu(block.proj.bias)).view(B, T, C, H, W)
    assert_close(out, ref, 2e-3, 2e-3, "attention output")
    out.square().mean().backward()
    for p in (block.qkv.to_q.weight, block.qkv.to_k.weight, block.qkv.to_v.weight, block.proj.weight):
        assert p.grad is not None and bool(th.isfinite(p.grad).all()) and float(p.grad.abs().sum()) > 0
    assert vg.grad is not None and bool(th.isfinite(vg.grad).all())


def test_attention_without_flows_and_int(cuda):
    block, cfg = make_block(cuda, itype="int")
    vid = th.randn(1, 3, 8, 12, 12, device=cuda)
    out = block(vid, None)
    assert tuple(out.shape) == tuple(vid.shape) and bool(th.isfinite(out).all())


def test_menus(cuda):
    assert isinstance(stnls_b200.agg.init(dict(agg_name="gather", ps=3, stride0=1)), stnls_b200.agg.NonLocalGather)
    assert isinstance(stnls_b200.normz.init(dict(normz_scale=5.0)), stnls_b200.normz.SoftmaxNormalize)
    with pytest.raises(NotImplementedError):
        stnls_b200.search.init(dict(search_name="n3mm", ws=3, wt=0))


def test_attention_stack_sums_to_attention(cuda):
    """NonLocalAttentionStack (lib/stnls/nn/non_local_attn_stack.py) with the K-stacked aggregator: summing its
    output over K gives what the projection-free NonLocalAttention block with the weighted patch sum returns
    (the reference's own ground truth for gather vs gather_add, tests/agg/test_gather_add.py:106-128)."""
    from stnls_b200.nn.non_local_attn import NonLocalAttentionStack, NonLocalAttention, extract_config
    base = dict(nheads=2, embed_dim=4, search_name="nls", ws=3, wt=1, ps=3, k=5, stride0=1, stride1=1.0,
                dist_type="l2", self_action="anchor", itype="float", normz_scale=10, k_agg=-1, strideIn=1,
                strideOut=1, use_attn_projection=False)
    blocks = []
    for cls, agg_name in ((NonLocalAttentionStack, "gather"), (NonLocalAttention, "wpsum")):
        cfg = AttrDict(dict(base, agg_name=agg_name))
        th.manual_seed(9)
        blocks.append(cls(extract_config(cfg, False), cfg, cfg, cfg).to(cuda))
    g = th.Generator().manual_seed(10)
    B, T, C, H, W = 1, 3, 8, 16, 16
    vid = th.randn(B, T, C, H, W, generator=g).to(cuda)
    flows = AttrDict(fflow=th.randn(B, T, 2, H, W, generator=g).to(cuda), bflow=th.randn(B, T, 2, H, W, generator=g).to(cuda))
    vg = vid.clone().requires_grad_(True)
    stack = blocks[0](vg, flows)
    assert tuple(stack.shape) == (B, 2, 5, T, C // 2, H, W)
    summed = blocks[1](vid, flows)                                  # (B,T,C,H,W)
    got = stack.sum(2).permute(0, 2, 1, 3, 4, 5).reshape(B, T, C, H, W)
    assert_close(got, summed, 1e-4, 1e-4, "stack.sum(K) vs weighted patch sum")
    stack.square().mean().backward()
    assert vg.grad is not None and bool(th.isfinite(vg.grad).all()) and float(vg.grad.abs().sum()) > 0
