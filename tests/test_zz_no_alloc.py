"""GPU: the allocation-free entry point `stnls_search_forward_planar_noalloc` (include/stnls_b200.h) against the
ordinary route pack_cl + pack_cl + stnls_search_forward, both through the C ABI.  The same kernels run on the same
values, so distances and indices must agree exactly (checked bit for bit; the tolerance of the float path, 1e-4,
is the fallback statement in the assert message).  Two calls in a row with DIFFERENT shapes share the built-in
scratch: the second must not see what the first left behind.  Collected last on purpose (the round's end runs -x)."""
import ctypes

import pytest
import torch as th

from stnls_b200 import _C, ops

pytestmark = pytest.mark.gpu
MAX_BYTES = 32 << 20   # STNLS_NOALLOC_MAX_BYTES


def _p(t):
    return ctypes.c_void_p(t.data_ptr())


def _case(L, cuda, seed, first, B, HD, T, F, H, W, ws, wt, ps, K, stride1, itype, dist):
    g = th.Generator().manual_seed(seed)
    q = th.randn(B, T, F, H, W, generator=g).to(cuda).contiguous()
    k = th.randn(B, T, F, H, W, generator=g).to(cuda).contiguous()
    isint = itype == "int"
    shape = (B, 1, T, 2 * wt, 2, H, W)
    fl = th.rand(shape, generator=g) / 2 + th.randint(-2, 3, shape, generator=g) + 0.2
    fl = (fl.round().int() if isint else fl).to(cuda).contiguous()
    geom = ops.make_geom(B, HD, T, F // HD, H, W, H, W, 1)
    opts = _C.SearchOpts(ws, wt, 0, 0, ps, K, 1, 1, 1, stride1, 1, 1, -(ps // 2), 0, 0, dist, int(isint), 1, 1, 1)
    nq = B * HD * T * H * W
    it = th.int32 if isint else th.float32
    st = ctypes.c_void_p(th.cuda.current_stream().cuda_stream)
    # ordinary route: caller-allocated channels-last copies
    q_cl, k_cl = ops.pack_cl(q, HD), ops.pack_cl(k, HD)
    d0 = th.full((nq * K,), 7.0, device=cuda)
    i0 = th.full((nq * K * 3,), -7, device=cuda, dtype=it)
    _C.check(L.stnls_search_forward(_p(q_cl), _p(k_cl), _p(fl), _p(d0), _p(i0), None, None, ctypes.byref(geom),
                                    ctypes.byref(opts), st))
    # allocation-free route: outputs allocated BEFORE the call, nothing allocated by it
    d1 = th.full((nq * K,), 9.0, device=cuda)
    i1 = th.full((nq * K * 3,), -9, device=cuda, dtype=it)
    th.cuda.synchronize()
    stats0 = th.cuda.memory_stats(cuda)["allocation.all.allocated"]
    free0, _ = th.cuda.mem_get_info(cuda)
    _C.check(L.stnls_search_forward_planar_noalloc(_p(q), _p(k), _p(fl), _p(d1), _p(i1), ctypes.byref(geom),
                                                   ctypes.byref(opts), st))
    th.cuda.synchronize()
    free1, _ = th.cuda.mem_get_info(cuda)
    assert th.cuda.memory_stats(cuda)["allocation.all.allocated"] == stats0, "the call went through torch's allocator"
    # the first call on a device looks the array up (module load under lazy loading may reserve it then)
    assert first or free1 >= free0, f"device memory shrank by {free0 - free1} bytes during the allocation-free call"
    assert bool(th.isfinite(d1).all())
    assert th.equal(i0, i1), "indices differ from the ordinary route"
    assert th.equal(d0, d1), f"distances differ from the ordinary route (max {float((d0 - d1).abs().max()):.3g}; tolerance 1e-4)"
    return d1, i1


def test_noalloc_matches_the_ordinary_route_twice_in_a_row(cuda):
    L = _C.lib()
    # 1: float flows, l2, tile kernel shape (Fp = 16), larger clip: fills most of what call 2 will reuse
    _case(L, cuda, 11, True, B=1, HD=2, T=3, F=32, H=64, W=48, ws=5, wt=1, ps=3, K=7, stride1=1.0, itype="float", dist=1)
    # 2: other shapes, padded channels (F/HD = 6 -> Fp = 8), prod distance: the pad channels and every pixel of the
    # smaller layout sit where call 1 left unrelated values
    _case(L, cuda, 12, False, B=2, HD=1, T=3, F=6, H=33, W=29, ws=3, wt=1, ps=3, K=5, stride1=1.0, itype="float", dist=0)
    # 3: integer offsets (bit-exact table kernel), and call 1's shape once more after call 2
    _case(L, cuda, 13, False, B=1, HD=2, T=3, F=32, H=21, W=19, ws=5, wt=1, ps=3, K=7, stride1=1.0, itype="int", dist=1)
    _case(L, cuda, 11, False, B=1, HD=2, T=3, F=32, H=64, W=48, ws=5, wt=1, ps=3, K=7, stride1=1.0, itype="float", dist=1)


def test_noalloc_rejects_sizes_beyond_the_built_in_scratch(cuda):
    L = _C.lib()
    B, HD, T, Fh, H, W = 1, 1, 4, 16, 256, 256      # q_cl + k_cl = 2 * 16 MiB + nothing to spare -> just over with T = 5
    T = 5
    geom = ops.make_geom(B, HD, T, Fh, H, W, H, W, 1)
    assert 2 * B * HD * T * H * W * geom.Fp * 4 > MAX_BYTES
    opts = _C.SearchOpts(3, 1, 0, 0, 3, 5, 1, 1, 1, 1.0, 1, 1, -1, 0, 0, 1, 0, 1, 1, 1)
    one = th.zeros(16, device=cuda)     # never touched: the size check comes first
    with pytest.raises(ValueError, match="STNLS_NOALLOC_MAX_BYTES"):
        _C.check(L.stnls_search_forward_planar_noalloc(_p(one), _p(one), _p(one), _p(one), _p(one), ctypes.byref(geom),
                                                       ctypes.byref(opts), None))
    with pytest.raises(ValueError, match="null argument"):
        _C.check(L.stnls_search_forward_planar_noalloc(None, _p(one), _p(one), _p(one), _p(one), ctypes.byref(geom),
                                                       ctypes.byref(opts), None))
