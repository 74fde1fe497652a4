"""Thin tensor-level wrappers over the C ABI (pointer extraction, shape checks, allocation).

Every function requires CUDA tensors and raises otherwise -- there is no CPU path.
"""
import ctypes

import torch as th

from . import _C
from ._C import AggOpts, Geom, SearchOpts

MAX_FUSED_K = 32
MAX_SOFTMAX_K = 64

# launch accounting: every C-ABI call below launches exactly one of our kernels.  `LAUNCHES` counts
# them; when `TIMELINE` is a list, each launch is bracketed by CUDA events on the launching stream
# (bench.py reads them for the per-kernel roofline).
LAUNCHES = 0
TIMELINE = None


def _launch(name, fn, *args):
    global LAUNCHES
    LAUNCHES += 1
    if TIMELINE is None:
        _C.check(fn(*args))
        return
    e0, e1 = th.cuda.Event(enable_timing=True), th.cuda.Event(enable_timing=True)
    e0.record()
    _C.check(fn(*args))
    e1.record()
    TIMELINE.append((name, e0, e1))


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


def _stream():
    return ctypes.c_void_p(th.cuda.current_stream().cuda_stream)


def _need_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError("stnls_b200: tensors must be CUDA tensors (there is no CPU path)")


def padded_channels(F):
    return (F + 3) // 4 * 4


def _planar_strides(vid, nheads):
    """(B,T,HD*F,H,W) or (B,HD,T,F,H,W) -> B,HD,T,F,H,W + element strides sB,sHD,sT,sC"""
    if vid.dim() not in (5, 6):
        raise ValueError("Must be 5 or 6 dims.")
    if vid.dtype != th.float32:
        raise TypeError(f"stnls_b200 supports float32 videos only (got {vid.dtype})")
    H, W = vid.shape[-2:]
    if vid.stride(-1) != 1 or vid.stride(-2) != W:
        vid = vid.contiguous()
    if vid.dim() == 5:
        B, T, Fc = vid.shape[:3]
        if Fc % nheads != 0:
            raise ValueError("must be multiple of each other.")
        F = Fc // nheads
        s = (vid.stride(0), F * vid.stride(2), vid.stride(1), vid.stride(2))
        return vid, (B, nheads, T, F, H, W), s
    B, HD, T, F = vid.shape[:4]
    if HD != nheads:
        raise ValueError(f"video has {HD} heads, expected {nheads}")
    return vid, (B, HD, T, F, H, W), (vid.stride(0), vid.stride(1), vid.stride(2), vid.stride(3))


def pack_cl(vid, nheads, t0=0, Tloc=None, counts=None):
    """planar video -> channels-last (B*HD, Tloc, H, W, Fp); optional fused divide by counts."""
    _need_cuda(vid)
    vid, (B, HD, T, F, H, W), (sB, sHD, sT, sC) = _planar_strides(vid, nheads)
    Tloc = T - t0 if Tloc is None else Tloc
    Fp = padded_channels(F)
    out = th.empty((B * HD, Tloc, H, W, Fp), device=vid.device, dtype=th.float32)
    ch, cw = (None, None) if counts is None else counts
    _launch("pack_cl", _C.lib().stnls_pack_cl, _ptr(out), _ptr(vid), B, HD, t0, Tloc, F, Fp, H, W,
                                    sB, sHD, sT, sC, _ptr(ch), _ptr(cw), _stream())
    return out


def unpack_cl(cl, shape, nheads, t0=0, out=None, accumulate=False):
    """channels-last -> planar tensor of `shape` ((B,T,HD*F,H,W) or (B,HD,T,F,H,W))."""
    _need_cuda(cl)
    if out is None:
        out = th.empty(shape, device=cl.device, dtype=th.float32)
    _, (B, HD, T, F, H, W), (sB, sHD, sT, sC) = _planar_strides(out, nheads)
    Tloc, Fp = cl.shape[1], cl.shape[4]
    assert cl.shape[0] == B * HD and cl.shape[2] == H and cl.shape[3] == W and cl.is_contiguous()
    _launch("unpack_cl", _C.lib().stnls_unpack_cl, _ptr(out), _ptr(cl), B, HD, t0, Tloc, F, Fp, H, W,
                                      sB, sHD, sT, sC, int(accumulate), _stream())
    return out


def _arr(ctype, vals):
    return (ctype * len(vals))(*vals)


def pack_cl_multi(vids, nheads, outs=None):
    """[planar video] -> [channels-last copy] in ONE launch (at most 4 videos, same head count; q, k, v of a step).
    outs[i]: None, or (buffer (N, Tpitch, H, W, Fp), first frame): the video's frames are written into the buffer starting
    at that frame (a rank packs its own frames straight into the owned + halo buffer its peers read)."""
    _need_cuda(*vids)
    assert 1 <= len(vids) <= 4
    info = [_planar_strides(v, nheads) for v in vids]
    res, B_, HD_, t0_, Tl_, F_, H_, W_, st_, tp_, tc_ = [], [], [], [], [], [], [], [], [], [], []
    for i, (vid, (B, HD, T, F, H, W), strides) in enumerate(info):
        dst = None if outs is None else outs[i]
        if dst is None:
            out, tc = th.empty((B * HD, T, H, W, padded_channels(F)), device=vid.device, dtype=th.float32), 0
        else:
            out, tc = dst
            assert out.is_contiguous() and out.dtype == th.float32 and out.shape[0] == B * HD and out.shape[1] >= tc + T
            assert tuple(out.shape[2:]) == (H, W, padded_channels(F))
        res.append(out)
        B_.append(B); HD_.append(HD); t0_.append(0); Tl_.append(T); F_.append(F); H_.append(H); W_.append(W)
        st_ += list(strides); tp_.append(out.shape[1]); tc_.append(tc)
    n = len(vids)
    _launch("pack_cl", _C.lib().stnls_pack_cl_multi, n, _arr(ctypes.c_void_p, [o.data_ptr() for o in res]),
            _arr(ctypes.c_void_p, [v.data_ptr() for v, _, _ in info]), _arr(ctypes.c_int, B_), _arr(ctypes.c_int, HD_),
            _arr(ctypes.c_int, t0_), _arr(ctypes.c_int, Tl_), _arr(ctypes.c_int, F_), _arr(ctypes.c_int, H_),
            _arr(ctypes.c_int, W_), _arr(ctypes.c_int64, st_), None, None, _arr(ctypes.c_int, tp_),
            _arr(ctypes.c_int, tc_), _stream())
    return res


def unpack_cl_multi(cls, shapes, nheads, cl_t0=None):
    """[channels-last] -> [planar tensor of shape] in ONE launch (at most 4; the three gradients of a step).
    cl_t0[i]: first frame of cls[i] to unpack (the owned slice of an owned + halo gradient); default 0."""
    _need_cuda(*cls)
    assert 1 <= len(cls) <= 4 and len(cls) == len(shapes)
    outs, B_, HD_, t0_, Tl_, F_, H_, W_, st_, tp_, tc_ = [], [], [], [], [], [], [], [], [], [], []
    for i, (cl, shape) in enumerate(zip(cls, shapes)):
        out = th.empty(shape, device=cl.device, dtype=th.float32)
        _, (B, HD, T, F, H, W), strides = _planar_strides(out, nheads)
        tc = 0 if cl_t0 is None else cl_t0[i]
        assert cl.is_contiguous() and cl.shape[0] == B * HD and cl.shape[1] >= tc + T
        assert tuple(cl.shape[2:]) == (H, W, padded_channels(F))
        outs.append(out)
        B_.append(B); HD_.append(HD); t0_.append(0); Tl_.append(T); F_.append(F); H_.append(H); W_.append(W)
        st_ += list(strides); tp_.append(cl.shape[1]); tc_.append(tc)
    n = len(cls)
    _launch("unpack_cl", _C.lib().stnls_unpack_cl_multi, n, _arr(ctypes.c_void_p, [o.data_ptr() for o in outs]),
            _arr(ctypes.c_void_p, [c.data_ptr() for c in cls]), _arr(ctypes.c_int, B_), _arr(ctypes.c_int, HD_),
            _arr(ctypes.c_int, t0_), _arr(ctypes.c_int, Tl_), _arr(ctypes.c_int, F_), _arr(ctypes.c_int, H_),
            _arr(ctypes.c_int, W_), _arr(ctypes.c_int64, st_), None, _arr(ctypes.c_int, tp_), _arr(ctypes.c_int, tc_),
            _stream())
    return outs


def make_geom(B, HD, T, F, qH, qW, kH, kW, stride0, tq0=0, Tq=None, tk0=0, Tk=None):
    Tq = T if Tq is None else Tq
    Tk = T if Tk is None else Tk
    return Geom(B, HD, T, tq0, Tq, tk0, Tk, F, padded_channels(F), qH, qW, kH, kW,
                (kH - 1) // stride0 + 1, (kW - 1) // stride0 + 1)


def search_forward(q_cl, k_cl, flows, geom, opts):
    """Fused search + anchor + top-K.  Returns dists, inds, kselect, reflect (last two refine-float only)."""
    _need_cuda(q_cl, k_cl, flows)
    g, o = geom, opts
    refine = o.wr > 0
    L = (o.Ks * o.wr * o.wr) if refine else ((2 * o.wt + 1) * o.ws * o.ws)
    Kout = min(o.K, L) if o.K > 0 else L
    dev = q_cl.device
    itype = th.int32 if o.itype_int else th.float32
    dists = th.empty((g.B, g.HD, g.Tq, g.nH, g.nW, Kout), device=dev, dtype=th.float32)
    inds = th.empty((g.B, g.HD, g.Tq, g.nH, g.nW, Kout, 3), device=dev, dtype=itype)
    kselect = reflect = None
    if refine and not o.itype_int:
        kselect = th.empty(dists.shape, device=dev, dtype=th.int32)
        reflect = th.zeros(tuple(flows.shape[:-1]) + (2,), device=dev, dtype=th.uint8)
    assert flows.is_contiguous() and flows.dtype == itype
    _launch("search_forward", _C.lib().stnls_search_forward, _ptr(q_cl), _ptr(k_cl), _ptr(flows), _ptr(dists), _ptr(inds),
                                           _ptr(kselect), _ptr(reflect), ctypes.byref(g), ctypes.byref(o),
                                           _stream())
    return dists, inds, kselect, reflect


def search_backward(q_cl, k_cl, flows, grad_dists, grad_inds, inds, kselect, reflect, geom, opts,
                    flows_grad):
    """Returns gq_cl, gk_cl, grad_flows (or None)."""
    _need_cuda(q_cl, k_cl, grad_dists, inds)
    g, o = geom, opts
    gq = th.zeros_like(q_cl)
    gk = th.zeros_like(k_cl)
    gflows = None
    grad_dists = grad_dists.contiguous().float()
    K = grad_dists.shape[-1]
    inds = inds.contiguous()
    if flows_grad:
        gflows = th.zeros_like(flows, dtype=th.float32)
        grad_inds = grad_inds.contiguous().float()
        if reflect is not None:
            reflect = reflect.to(th.uint8).contiguous()
    _launch("search_backward", _C.lib().stnls_search_backward, _ptr(gq), _ptr(gk), _ptr(gflows), _ptr(q_cl), _ptr(k_cl),
                                            _ptr(flows if flows_grad else None), _ptr(grad_dists),
                                            _ptr(grad_inds if flows_grad else None), _ptr(inds),
                                            _ptr(kselect), _ptr(reflect), K,
                                            ctypes.byref(g), ctypes.byref(o), _stream())
    return gq, gk, gflows


def search_backward_sm(q_cl, k_cl, flows, grad_weights, weights, scale, grad_inds, inds, geom, opts, flows_grad,
                       gk_out=None, gq_out=None, gflows_out=None):
    """search_backward with the normalise step's backward folded in (grad_dists computed from grad_weights, weights).
    *_out: optional PRE-ZEROED accumulators (views of one workspace, a symmetric-memory buffer, ...)."""
    _need_cuda(q_cl, k_cl, grad_weights, weights, inds)
    K = weights.shape[-1]
    assert grad_weights.is_contiguous() and weights.is_contiguous() and inds.is_contiguous()
    gq = th.zeros_like(q_cl) if gq_out is None else gq_out
    gk = th.zeros_like(k_cl) if gk_out is None else gk_out
    gflows = None
    if flows_grad:
        gflows = th.zeros_like(flows, dtype=th.float32) if gflows_out is None else gflows_out
        assert grad_inds is not None and grad_inds.is_contiguous() and grad_inds.dtype == th.float32
    _launch("search_backward", _C.lib().stnls_search_backward_sm, _ptr(gq), _ptr(gk), _ptr(gflows), _ptr(q_cl),
            _ptr(k_cl), _ptr(flows if flows_grad else None), _ptr(grad_weights), _ptr(weights), float(scale),
            _ptr(grad_inds if flows_grad else None), _ptr(inds), None, None, K, ctypes.byref(geom), ctypes.byref(opts),
            _stream())
    return gq, gk, gflows


def anchor_self_time(dists, inds, flows, wt, stride0, qH, qW, kH, kW):
    """In place on dists (B,HD,Q,G,S) / inds (B,HD,Q,G,S,3); flows (B,HD_f,T,nflow,2,nH,nW) of inds' dtype."""
    _need_cuda(dists, inds, flows)
    assert dists.is_contiguous() and inds.is_contiguous() and flows.is_contiguous() and flows.dtype == inds.dtype
    B, HD, Q, G, S = dists.shape
    _launch("anchor_self_time", _C.lib().stnls_anchor_self_time, _ptr(dists), _ptr(inds), _ptr(flows), B, HD,
            flows.shape[1], flows.shape[2], G, S, flows.shape[3], wt, stride0, qH, qW, kH, kW,
            int(inds.dtype == th.int32), _stream())
    return dists, inds


_COUNT_CACHE = {}


def overlap_counts(H, nH, ps, stride0, dilation, patch_offset, device):
    """Separable patch-overlap counts (int32 device vector) + their minimum."""
    key = (H, nH, ps, stride0, dilation, patch_offset, str(device))
    if key not in _COUNT_CACHE:
        buf = (ctypes.c_int * H)()
        mn = _C.lib().stnls_overlap_counts(buf, H, nH, ps, stride0, dilation, patch_offset)
        _COUNT_CACHE[key] = (th.tensor(list(buf), dtype=th.int32, device=device), mn)
    return _COUNT_CACHE[key]


def wpsum_forward(v_cl, weights, flows, counts, geom, opts):
    _need_cuda(v_cl, weights, flows)
    g = geom
    out = th.empty((g.B * g.HD, g.Tq, g.kH, g.kW, g.Fp), device=v_cl.device, dtype=th.float32)
    _launch("wpsum_forward", _C.lib().stnls_wpsum_forward, _ptr(out), _ptr(v_cl), _ptr(weights), _ptr(flows),
                                          _ptr(counts[0]), _ptr(counts[1]),
                                          ctypes.byref(g), ctypes.byref(opts), _stream())
    return out


def wpsum_forward_sm(v_cl, dists, scale, flows, counts, geom, opts):
    """Aggregation with the normalise step folded in: weights = softmax(scale * dists[..., :K]) inside the kernel.
    Returns out_cl and the weights (for the backward pass)."""
    global LAUNCHES
    _need_cuda(v_cl, dists, flows)
    g = geom
    assert dists.is_contiguous() and dists.dtype == th.float32
    Kin = dists.shape[-1]
    out = th.empty((g.B * g.HD, g.Tq, g.kH, g.kW, g.Fp), device=v_cl.device, dtype=th.float32)
    weights = th.empty(tuple(dists.shape[:-1]) + (opts.K,), device=v_cl.device, dtype=th.float32)
    n = ctypes.c_int(0)
    _launch("wpsum_forward", _C.lib().stnls_wpsum_forward_sm, _ptr(out), _ptr(weights), _ptr(v_cl), _ptr(dists), Kin,
            float(scale), _ptr(flows), _ptr(counts[0]), _ptr(counts[1]), ctypes.byref(g), ctypes.byref(opts), _stream(),
            ctypes.byref(n))
    LAUNCHES += max(n.value, 1) - 1     # 2 where the normalise kernel had to run on its own
    return out, weights


def wpsum_backward(gout_cl, v_cl, weights, flows, geom, opts, gv_out=None, gw_out=None, gf_out=None):
    """*_out: optional PRE-ZEROED accumulators (views of one workspace, a symmetric-memory buffer, ...)."""
    _need_cuda(gout_cl, v_cl, weights, flows)
    gv = th.zeros_like(v_cl) if gv_out is None else gv_out
    gw = th.zeros_like(weights) if gw_out is None else gw_out
    if opts.itype_int:
        gf = None
    else:
        gf = th.zeros(flows.shape, device=flows.device, dtype=th.float32) if gf_out is None else gf_out
    _launch("wpsum_backward", _C.lib().stnls_wpsum_backward, _ptr(gv), _ptr(gw), _ptr(gf), _ptr(gout_cl), _ptr(v_cl),
                                           _ptr(weights), _ptr(flows),
                                           ctypes.byref(geom), ctypes.byref(opts), _stream())
    return gv, gw, gf


def gather_forward(v_cl, weights, flows, counts, geom, opts):
    """NonLocalGather forward: (N*K, Tq, H, W, Fp), slab n*K + k = neighbour k."""
    _need_cuda(v_cl, weights, flows)
    g = geom
    out = th.empty((g.B * g.HD * opts.K, g.Tq, g.kH, g.kW, g.Fp), device=v_cl.device, dtype=th.float32)
    _launch("gather_forward", _C.lib().stnls_gather_forward, _ptr(out), _ptr(v_cl), _ptr(weights), _ptr(flows),
            _ptr(counts[0]), _ptr(counts[1]), ctypes.byref(g), ctypes.byref(opts), _stream())
    return out


def gather_backward(gstack_cl, v_cl, weights, flows, geom, opts):
    _need_cuda(gstack_cl, v_cl, weights, flows)
    gv = th.zeros_like(v_cl)
    gw = th.zeros_like(weights)
    gf = None if opts.itype_int else th.zeros(flows.shape, device=flows.device, dtype=th.float32)
    _launch("gather_backward", _C.lib().stnls_gather_backward, _ptr(gv), _ptr(gw), _ptr(gf), _ptr(gstack_cl),
            _ptr(v_cl), _ptr(weights), _ptr(flows), ctypes.byref(geom), ctypes.byref(opts), _stream())
    return gv, gw, gf


def seg_copy(pairs, accumulate=False):
    """[(dst, src)] contiguous fp32 tensors of equal numel (numel % 4 == 0, 16-byte aligned) copied -- or added with
    vector REDs -- in ONE launch (at most 8 pairs per launch).  src may live on a peer GPU (symmetric memory)."""
    pairs = [(d, s_) for d, s_ in pairs if d.numel()]
    for i in range(0, len(pairs), 8):
        chunk = pairs[i:i + 8]
        n = len(chunk)
        for d, s_ in chunk:
            assert d.is_contiguous() and s_.is_contiguous() and d.numel() == s_.numel() and d.dtype == th.float32
        dst = (ctypes.c_void_p * n)(*[d.data_ptr() for d, _ in chunk])
        src = (ctypes.c_void_p * n)(*[s_.data_ptr() for _, s_ in chunk])
        cnt = (ctypes.c_int64 * n)(*[d.numel() for d, _ in chunk])
        _launch("seg_copy", _C.lib().stnls_seg_copy, dst, src, cnt, n, int(accumulate), _stream())


def seg_copy2d(jobs, accumulate=False):
    """jobs: [(dst_ptr, src_ptr, nfloat, rows, dst_pitch, src_pitch)] raw device addresses (src may be peer memory):
    rows runs of nfloat floats each, copied -- or added with vector REDs -- in ONE launch per 8 jobs."""
    jobs = [j for j in jobs if j[2] > 0 and j[3] > 0]
    for i in range(0, len(jobs), 8):
        ch = jobs[i:i + 8]
        n = len(ch)
        _launch("seg_copy", _C.lib().stnls_seg_copy2d, _arr(ctypes.c_void_p, [j[0] for j in ch]),
                _arr(ctypes.c_void_p, [j[1] for j in ch]), _arr(ctypes.c_int64, [j[2] for j in ch]),
                _arr(ctypes.c_int, [j[3] for j in ch]), _arr(ctypes.c_int64, [j[4] for j in ch]),
                _arr(ctypes.c_int64, [j[5] for j in ch]), n, int(accumulate), _stream())


def seg_copy_ok(*ts):
    return all(t.is_cuda and t.dtype == th.float32 and t.is_contiguous() and t.numel() % 4 == 0 and t.data_ptr() % 16 == 0
               for t in ts)


def _flow_check(fflow, bflow):
    _need_cuda(fflow, bflow)
    if fflow.dim() != 5 or fflow.shape[2] != 2 or fflow.shape != bflow.shape:
        raise ValueError("search_flow: fflow and bflow must both be (B,T,2,H,W)")
    if fflow.dtype != th.float32 or bflow.dtype != th.float32:
        raise TypeError("stnls_b200 supports float32 flows only")


def search_flow_forward(fflow, bflow, wt, stride0):
    """(B,T,2,H,W) x2 -> accumulated flows (B,T,W_t-1,2,nH,nW), W_t = min(2wt+1,T)."""
    _flow_check(fflow, bflow)
    fflow, bflow = fflow.contiguous(), bflow.contiguous()
    B, T, _, H, W = fflow.shape
    nH, nW = (H - 1) // stride0 + 1, (W - 1) // stride0 + 1
    W_t = min(2 * wt + 1, T)
    flows = th.empty((B, T, W_t - 1, 2, nH, nW), device=fflow.device, dtype=th.float32)
    if flows.numel():
        _launch("search_flow_forward", _C.lib().stnls_search_flow_forward, _ptr(flows), _ptr(fflow), _ptr(bflow),
                B, T, H, W, wt, stride0, _stream())
    return flows


def search_flow_backward(grad_flows, fflow, bflow, flows, wt, stride0):
    _flow_check(fflow, bflow)
    B, T, _, H, W = fflow.shape
    gf, gb = th.zeros_like(fflow), th.zeros_like(bflow)
    grad_flows = grad_flows.contiguous().float()
    if flows.numel():
        _launch("search_flow_backward", _C.lib().stnls_search_flow_backward, _ptr(gf), _ptr(gb), _ptr(grad_flows),
                _ptr(fflow.contiguous()), _ptr(bflow.contiguous()), _ptr(flows.contiguous()),
                B, T, H, W, wt, stride0, _stream())
    return gf, gb


def softmax_forward(dists, K, scale):
    """weights = softmax(scale * dists[..., :K], -1) in one kernel (dists contiguous, last dim Kin >= K)."""
    _need_cuda(dists)
    assert dists.is_contiguous() and dists.dtype == th.float32
    Kin = dists.shape[-1]
    rows = dists.numel() // Kin
    out = th.empty(tuple(dists.shape[:-1]) + (K,), device=dists.device, dtype=th.float32)
    _launch("softmax_forward", _C.lib().stnls_softmax_forward, _ptr(out), _ptr(dists), rows, K, Kin,
            float(scale), _stream())
    return out


def softmax_backward(grad_weights, weights, Kin, scale):
    _need_cuda(grad_weights, weights)
    K = weights.shape[-1]
    rows = weights.numel() // K
    grad_weights = grad_weights.contiguous().float()
    gd = th.empty(tuple(weights.shape[:-1]) + (Kin,), device=weights.device, dtype=th.float32)
    _launch("softmax_backward", _C.lib().stnls_softmax_backward, _ptr(gd), _ptr(grad_weights), _ptr(weights),
            rows, K, Kin, float(scale), _stream())
    return gd


def add_inplace(dst, src):
    assert dst.is_contiguous() and src.is_contiguous() and dst.numel() == src.numel()
    _launch("add_inplace", _C.lib().stnls_add_inplace, _ptr(dst), _ptr(src), dst.numel(), _stream())
    return dst


def ffma_peak_tflops(device, iters=4096, reps=5):
    """Measured fp32 FMA-pipe throughput of `device` in TFLOP/s (stnls_ffma_peak, CUDA events, best of `reps`)."""
    sms = th.cuda.get_device_properties(device).multi_processor_count
    blocks = sms * 8
    sink = th.zeros(4, device=device)
    flops = 2.0 * 16 * iters * 256 * blocks
    best = None
    for i in range(reps + 2):
        e0, e1 = th.cuda.Event(enable_timing=True), th.cuda.Event(enable_timing=True)
        e0.record()
        _C.check(_C.lib().stnls_ffma_peak(_ptr(sink), blocks, iters, _stream()))
        e1.record()
        e1.synchronize()
        ms = e0.elapsed_time(e1)
        if i >= 2:
            best = ms if best is None else min(best, ms)
    return flops / (best * 1e-3) / 1e12
