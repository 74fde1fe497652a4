"""Host-side selection glue of stnls_b200.nn (torch only, no kernels): topk / topk_each follow
lib/stnls/nn/topk.py:65-123,226-269 with a stable sort."""
import torch as th

import stnls_b200
from oracle import ref


def test_topk_matches_oracle_selection():
    g = th.Generator().manual_seed(0)
    d, i = th.randn(2, 3, 11, generator=g), th.randn(2, 3, 11, 3, generator=g)
    for anchor in (False, True):
        for desc in (False, True):
            dk, ik, order = stnls_b200.nn.topk(d, i, 4, dim=2, anchor=anchor, descending=desc, return_order=True)
            do, io, oo = ref.topk(d, i, 4, anchor, desc)
            assert th.equal(dk, do) and th.equal(ik, io) and th.equal(order, oo)
            # any candidate axis
            dk2, ik2 = stnls_b200.nn.topk(d.permute(0, 2, 1).contiguous(), i.permute(0, 2, 1, 3).contiguous(), 4,
                                          dim=1, anchor=anchor, descending=desc)
            assert th.equal(dk2.permute(0, 2, 1), dk) and th.equal(ik2.permute(0, 2, 1, 3), ik)
    assert stnls_b200.nn.topk(d, i, -1)[0] is d


def test_topk_each_matches_oracle_selection():
    g = th.Generator().manual_seed(1)
    d, i = th.randn(2, 5, 3, 9, generator=g), th.randn(2, 5, 3, 9, 3, generator=g)
    for anchor in (False, True):
        for k in (1, 4, 20):
            dk, ik = stnls_b200.nn.topk_each(d, i, k, False, anchor_self=anchor)
            do, io = ref.topk_each(d, i, k, False, anchor_self=anchor)
            assert th.equal(dk, do) and th.equal(ik, io)
