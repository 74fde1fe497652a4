"""GPU: out-of-bounds WRITE detection without compute-sanitizer (closed on this pool).  Every output / accumulator the
kernels write is a window inside a larger buffer pre-filled with a sentinel; after the call the guard bands on both
sides must be untouched.  Goes through the C ABI with raw pointers (the Python wrappers allocate exact-size tensors,
where an overrun could land in a neighbouring allocation unnoticed)."""
import ctypes

import pytest
import torch as th

from stnls_b200 import _C, ops

pytestmark = pytest.mark.gpu
GUARD = 4096
SENT = 12345.0


class Guarded:
    def __init__(self, numel, dtype, dev, fill=0):
        self.buf = th.full((numel + 2 * GUARD,), SENT, dtype=th.float32, device=dev).view(dtype) if dtype != th.float32 \
            else th.full((numel + 2 * GUARD,), SENT, dtype=th.float32, device=dev)
        self.win = self.buf[GUARD:GUARD + numel]
        self.win.fill_(fill)
        self.ref = th.full((GUARD,), SENT, dtype=th.float32, device=dev).view(dtype)

    def ptr(self):
        return ctypes.c_void_p(self.win.data_ptr())

    def check(self, what):
        assert th.equal(self.buf[:GUARD], self.ref) and th.equal(self.buf[-GUARD:], self.ref), what + ": guard band overwritten"


@pytest.mark.parametrize("itype,stride1,F,H,W", [("float", 0.5, 16, 37, 35), ("float", 1.0, 32, 34, 41), ("int", 1.0, 16, 21, 19)])
def test_no_write_outside_the_declared_buffers(itype, stride1, F, H, W, cuda):
    L = _C.lib()
    B, HD, T, ws, wt, ps, K = 1, 2, 3, 5, 1, 3, 7
    g = th.Generator().manual_seed(3)
    q, k, v = [th.randn(B, T, F, H, W, generator=g).to(cuda) for _ in range(3)]
    shape = (B, 1, T, 2 * wt, 2, H, W)
    fl = th.rand(shape, generator=g) / 2 + th.randint(-3, 3, shape, generator=g) + 0.2
    isint = itype == "int"
    fl = (fl.round().int() if isint else fl).to(cuda).contiguous()
    q_cl, k_cl, v_cl = ops.pack_cl(q, HD), ops.pack_cl(k, HD), ops.pack_cl(v, HD)
    geom = ops.make_geom(B, HD, T, F // HD, H, W, H, W, 1)
    opts = _C.SearchOpts(ws, wt, 0, 0, ps, K, 1, 1, 1, stride1, 1, 1, -(ps // 2), 0, 0, 1, int(isint), 1, 1, 1)  # st_offset = W_t - flow slots = 1
    nq = B * HD * T * H * W
    it = th.int32 if isint else th.float32
    dists, inds = Guarded(nq * K, th.float32, cuda), Guarded(nq * K * 3, it, cuda)
    st = ctypes.c_void_p(th.cuda.current_stream().cuda_stream)
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    _C.check(L.stnls_search_forward(p(q_cl), p(k_cl), p(fl), dists.ptr(), inds.ptr(), None, None, ctypes.byref(geom),
                                    ctypes.byref(opts), st))
    th.cuda.synchronize()
    dists.check("search_forward dists"); inds.check("search_forward inds")
    assert bool(th.isfinite(dists.win).all())
    # backward of the search
    gq, gk = Guarded(q_cl.numel(), th.float32, cuda), Guarded(k_cl.numel(), th.float32, cuda)
    gd = th.randn(nq * K, generator=g).to(cuda)
    gi = th.randn(nq * K * 3, generator=g).to(cuda)
    gfl = Guarded(fl.numel(), th.float32, cuda)
    _C.check(L.stnls_search_backward(gq.ptr(), gk.ptr(), None if isint else gfl.ptr(), p(q_cl), p(k_cl),
                                     None if isint else p(fl), p(gd), None if isint else p(gi), inds.ptr(), None, None, K,
                                     ctypes.byref(geom), ctypes.byref(opts), st))
    th.cuda.synchronize()
    gq.check("search_backward grad_q"); gk.check("search_backward grad_k"); gfl.check("search_backward grad_flows")
    # aggregation forward / backward on the kept offsets
    aopts = _C.AggOpts(ps, K, 1, 1, 1, -(ps // 2), int(isint))
    ch, _ = ops.overlap_counts(H, H, ps, 1, 1, -(ps // 2), cuda)
    cw, _ = ops.overlap_counts(W, W, ps, 1, 1, -(ps // 2), cuda)
    w = th.softmax(-dists.win.view(nq, K), -1).contiguous()
    out = Guarded(q_cl.numel(), th.float32, cuda)
    _C.check(L.stnls_wpsum_forward(out.ptr(), p(v_cl), p(w), inds.ptr(), p(ch), p(cw), ctypes.byref(geom),
                                   ctypes.byref(aopts), st))
    gv, gw, gf = Guarded(v_cl.numel(), th.float32, cuda), Guarded(nq * K, th.float32, cuda), Guarded(nq * K * 3, th.float32, cuda)
    go = th.randn(q_cl.shape, generator=g).to(cuda)
    _C.check(L.stnls_wpsum_backward(gv.ptr(), gw.ptr(), None if isint else gf.ptr(), p(go), p(v_cl), p(w), inds.ptr(),
                                    ctypes.byref(geom), ctypes.byref(aopts), st))
    th.cuda.synchronize()
    out.check("wpsum_forward out"); gv.check("wpsum_backward grad_v"); gw.check("wpsum_backward grad_weights")
    gf.check("wpsum_backward grad_flows")
    assert bool(th.isfinite(out.win).all()) and bool(th.isfinite(gv.win).all())


@pytest.mark.parametrize("itype,tile", [("float", "1"), ("float", "0"), ("int", "1")])
def test_nan_pixels_leave_no_output_slot_unwritten(itype, tile, cuda, monkeypatch):
    """NaN / inf pixels in q and k (a diverging training run): the reference returns a full arg-sorted list, so every slot
    of dists / inds must be WRITTEN -- the wrappers allocate them uninitialised and the backward dereferences inds.
    Through the C ABI into sentinel-filled buffers, tile kernel and table kernel (STNLS_FWD_TILE=0), float and int."""
    monkeypatch.setenv("STNLS_FWD_TILE", tile)
    L = _C.lib()
    B, HD, T, F, H, W, ws, wt, ps, K = 1, 2, 3, 16, 40, 36, 5, 1, 3, 9
    g = th.Generator().manual_seed(21)
    q, k = [th.randn(B, T, F, H, W, generator=g) for _ in range(2)]
    k[0, 1, 3, 10:14, 8:30] = float("nan")
    k[0, 2, :, 20, 20] = float("inf")
    q[0, 0, 5, 30:33, 5:9] = float("nan")
    k[0, 0] = float("nan")                       # a whole key frame
    isint = itype == "int"
    shape = (B, 1, T, 2 * wt, 2, H, W)
    fl = th.rand(shape, generator=g) * 3 - 1.5
    fl = (fl.round().int() if isint else fl).to(cuda).contiguous()
    q_cl, k_cl = ops.pack_cl(q.to(cuda), HD), ops.pack_cl(k.to(cuda), HD)
    geom = ops.make_geom(B, HD, T, F // HD, H, W, H, W, 1)
    opts = _C.SearchOpts(ws, wt, 0, 0, ps, K, 1, 1, 1, 1.0, 1, 1, -(ps // 2), 0, 0, 1, int(isint), 1, 1, 1)
    nq = B * HD * T * H * W
    dists = th.full((nq * K,), SENT, device=cuda)
    inds = th.full((nq * K * 3,), -777, dtype=th.int32, device=cuda) if isint else th.full((nq * K * 3,), SENT, device=cuda)
    st = ctypes.c_void_p(th.cuda.current_stream().cuda_stream)
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    _C.check(L.stnls_search_forward(p(q_cl), p(k_cl), p(fl), p(dists), p(inds), None, None, ctypes.byref(geom),
                                    ctypes.byref(opts), st))
    th.cuda.synchronize()
    assert int((dists == SENT).sum()) == 0, "dists slots left unwritten"
    assert int((inds == (-777 if isint else SENT)).sum()) == 0, "inds slots left unwritten"
    iv = inds.view(nq, K, 3).float()
    # the temporal window shifts at the ends of the clip (up to 2*wt frames to one side)
    assert bool((iv[..., 0].abs() <= 2 * wt).all()) and bool((iv[..., 1:].abs() < 64).all()), "indices outside any window"
