"""GPU: the fused normalise step against the torch op chain the reference runs
(lib/stnls/normz/softmax.py:21-42): slice to k_agg, negate for l2, scale, softmax.  fp32, tolerance 1e-6."""
import pytest
import torch as th

import stnls_b200
from util import assert_close

pytestmark = pytest.mark.gpu


def reference_normz(dists, inds, scale, dist_type, k_agg):
    if k_agg > 0:
        dists, inds = dists[..., :k_agg].contiguous(), inds[..., :k_agg, :].contiguous()
    if dist_type == "l2":
        dists = -dists
    return th.softmax(scale * dists, -1), inds


@pytest.mark.parametrize("K,k_agg", [(10, -1), (9, -1), (20, 7), (33, -1), (5, 5), (64, 40)])
@pytest.mark.parametrize("dist_type", ["l2", "prod"])
def test_softmax_normalize(K, k_agg, dist_type, cuda):
    g = th.Generator().manual_seed(5)
    d = (th.randn(2, 2, 3, 7, 5, K, generator=g) * 3).to(cuda)
    bad = -float("inf") if dist_type == "prod" else float("inf")
    d[0, 0, 0, 0, 0, -1] = bad            # an invalid candidate gets weight 0
    inds = th.randn(2, 2, 3, 7, 5, K, 3, generator=g).to(cuda)
    normz = stnls_b200.normz.SoftmaxNormalize(10.0, 0.0, dist_type, k_agg)
    dg = d.clone().requires_grad_(True)
    w, i = normz(dg, inds)
    dr = d.clone().requires_grad_(True)
    w_ref, i_ref = reference_normz(dr, inds, 10.0, dist_type, k_agg)
    assert w.shape == w_ref.shape and th.equal(i, i_ref)
    assert_close(w, w_ref, 1e-6, 1e-5, "weights")
    go = th.randn(w_ref.shape, generator=g).to(cuda)
    (w * go).sum().backward()
    (w_ref * go).sum().backward()
    assert dg.grad.shape == d.shape
    assert_close(dg.grad, th.nan_to_num(dr.grad), 1e-5, 1e-4, "grad_dists")


def test_softmax_rejects_cpu():
    with pytest.raises(RuntimeError):
        stnls_b200.normz.softmax_scaled(th.zeros(4, 10), 10.0)
