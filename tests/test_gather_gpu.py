"""GPU parity of NonLocalGather (K-stacked aggregation): CUDA path vs the CPU oracle, vs the
reference's own gather kernels (oracle/_ref) and the reference's property test
`gather_add == sum_K gather` (tests/agg/test_gather_add.py:106-128 of the reference)."""
import pytest
import torch as th

import stnls_b200
from util import assert_close

pytestmark = pytest.mark.gpu

CASES = [
    dict(name="gather_int_ps3", itype="int", ps=3, stride0=1, K=4, F=8, HD=2),
    dict(name="gather_float_ps3", itype="float", ps=3, stride0=1, K=5, F=8, HD=2),
    dict(name="gather_float_ps1", itype="float", ps=1, stride0=1, K=3, F=6, HD=1),
    dict(name="gather_float_ps5_s2", itype="float", ps=5, stride0=2, K=3, F=4, HD=1),
    dict(name="gather_int_noreflect", itype="int", ps=3, stride0=2, K=3, F=4, HD=1, reflect_bounds=False),
    dict(name="gather_float_intvalued", itype="float", ps=3, stride0=1, K=3, F=4, HD=1, integer_valued=True),
]
B, T, H, W = 1, 4, 16, 14


def inputs(c):
    gen = th.Generator().manual_seed(91)
    v = th.randn(B, T, c["F"] * c["HD"], H, W, generator=gen)
    nH, nW = (H - 1) // c["stride0"] + 1, (W - 1) // c["stride0"] + 1
    w = th.rand(B, c["HD"], T, nH, nW, c["K"], generator=gen)
    fl = th.zeros(B, c["HD"], T, nH, nW, c["K"], 3)
    fl[..., 0] = th.randint(-2, 3, fl[..., 0].shape, generator=gen).float()
    fl[..., 1:] = th.randint(-5, 6, fl[..., 1:].shape, generator=gen).float()
    if c["itype"] == "float" and not c.get("integer_valued"):
        fl[..., 1:] += th.rand(fl[..., 1:].shape, generator=gen) - 0.5
    return v, w, fl


def run_cuda(c, v, w, fl, cuda, go=None):
    kw = dict(dilation=c.get("dilation", 1), reflect_bounds=c.get("reflect_bounds", True), itype=c["itype"])
    agg = stnls_b200.agg.NonLocalGather(c["ps"], c["stride0"], **kw)
    flt = c["itype"] == "float"
    vg, wg = v.to(cuda).requires_grad_(True), w.to(cuda).requires_grad_(True)
    fg = fl.to(cuda).requires_grad_(flt)
    stack = agg(vg, wg, fg)
    if go is not None:
        (stack * go.to(cuda)).sum().backward()
    return stack, vg, wg, fg, kw


@pytest.mark.parametrize("c", CASES, ids=[c["name"] for c in CASES])
def test_gather_vs_oracle(c, cuda, oracle):
    v, w, fl = inputs(c)
    K, F, HD = c["K"], c["F"], c["HD"]
    go = th.randn(B, HD, K, T, F, H, W, generator=th.Generator().manual_seed(92))
    stack, vg, wg, fg, kw = run_cuda(c, v, w, fl, cuda, go)
    st_ref, counts = oracle.gather_forward(v, w, fl, c["ps"], c["stride0"], **kw)
    assert tuple(stack.shape) == (B, HD, K, T, F, H, W)
    assert_close(stack, st_ref, 1e-4, 1e-4, c["name"] + " stack")
    gv, gw, gi = oracle.gather_backward(go, v, w, fl, counts, c["ps"], c["stride0"], **kw)
    assert_close(vg.grad, gv.permute(0, 2, 1, 3, 4, 5).reshape(v.shape), 1e-3, 1e-3, c["name"] + " grad_v")
    assert_close(wg.grad, gw, 1e-3, 1e-3, c["name"] + " grad_w")
    if c["itype"] == "float":
        assert_close(fg.grad[..., 1:], gi[..., 1:], 2e-3, 2e-3, c["name"] + " grad_inds")
        assert float(fg.grad[..., 0].abs().max()) == 0.0


@pytest.mark.parametrize("itype", ["int", "float"])
def test_gather_add_is_sum_of_gather(itype, cuda):
    c = dict(itype=itype, ps=3, stride0=1, K=4, F=8, HD=2)
    v, w, fl = inputs(c)
    stack, *_ = run_cuda(c, v, w, fl, cuda)
    out = stnls_b200.agg.WeightedPatchSum(3, 1, itype=itype)(v.to(cuda), w.to(cuda), fl.to(cuda))
    assert_close(stack.sum(2), out, 1e-5, 1e-5, "sum_K gather vs weighted patch sum")


@pytest.mark.parametrize("c", CASES[:4] + CASES[5:], ids=[c["name"] for c in CASES[:4] + CASES[5:]])
def test_gather_vs_reference_kernels(c, cuda, oracle):
    from oracle import ref_cuda
    if not ref_cuda.available() or not hasattr(ref_cuda.mod(), "gather_int_forward"):
        pytest.skip("oracle/_ref without the gather kernels")
    v, w, fl = inputs(c)
    K, F, HD = c["K"], c["F"], c["HD"]
    go = th.randn(B, HD, K, T, F, H, W, generator=th.Generator().manual_seed(93))
    stack, vg, wg, fg, kw = run_cuda(c, v, w, fl, cuda, go)
    st_r, counts_r = ref_cuda.gather_forward(v.to(cuda), w.to(cuda), fl.to(cuda), c["ps"], c["stride0"], **kw)
    st_o, counts_o = oracle.gather_forward(v, w, fl, c["ps"], c["stride0"], **kw)
    assert th.equal(counts_r[0, 0].cpu(), counts_o), "oracle counts != reference counts"
    assert_close(st_o, st_r, 1e-4, 1e-4, "oracle vs reference stack")
    assert_close(stack, st_r, 1e-4, 1e-4, "cuda vs reference stack")
    gv, gw, gi = ref_cuda.gather_backward(go.to(cuda), v.to(cuda), w.to(cuda), fl.to(cuda), counts_r, c["ps"],
                                          c["stride0"], **kw)
    gvo, gwo, gio = oracle.gather_backward(go, v, w, fl, counts_o, c["ps"], c["stride0"], **kw)
    assert_close(gvo, gv, 1e-3, 1e-3, "oracle vs reference grad_v")
    assert_close(gwo, gw, 1e-3, 1e-3, "oracle vs reference grad_w")
    assert_close(vg.grad, gv.permute(0, 2, 1, 3, 4, 5).reshape(v.shape), 1e-3, 1e-3, "cuda vs reference grad_v")
    assert_close(wg.grad, gw, 1e-3, 1e-3, "cuda vs reference grad_w")
    if c["itype"] == "float":
        assert_close(gio, gi, 2e-3, 2e-3, "oracle vs reference grad_inds")
        assert_close(fg.grad, gi, 2e-3, 2e-3, "cuda vs reference grad_inds")
