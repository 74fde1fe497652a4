"""
oracle/ref_cuda.py -- TEST INFRASTRUCTURE.  Drives the REFERENCE's own compiled kernels
(oracle/_ref/stnls_cuda_ref*.so, built by oracle/build_ref.py from /root/reference/lib/csrc) exactly
the way the reference's Python does, without importing the (broken) `stnls` package:

  lib/stnls/search/impl/non_local_search.py:23-212   lib/stnls/search/impl/refinement.py:18-181
  lib/stnls/nn/anchor_self.py:18-67                  lib/stnls/nn/topk.py (via oracle.ref.topk, stable)
  lib/stnls/agg/gather_add.py:36-188

Needs a GPU.  Same call signatures as oracle/ref.py so tests can swap one for the other.
"""
import glob
import importlib.util
import os

import torch as th

from . import ref as cref

_HERE = os.path.dirname(os.path.abspath(__file__))
_MOD = None


def available():
    return bool(glob.glob(os.path.join(_HERE, "_ref", "stnls_cuda_ref*.so"))) and th.cuda.is_available()


def mod():
    global _MOD
    if _MOD is None:
        so = glob.glob(os.path.join(_HERE, "_ref", "stnls_cuda_ref*.so"))[0]
        spec = importlib.util.spec_from_file_location("stnls_cuda_ref", so)
        _MOD = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(_MOD)
    return _MOD


def _shape(vid, nheads):
    return cref.shape_vid(vid, nheads)


def nls_forward(vid0, vid1, flows, ws, wt, ps, k, nheads=1, stride0=1, stride1=1, dist_type="l2",
                dilation=1, self_action=None, ws_interior=0, reflect_bounds=True, full_ws=True,
                use_adj=False, off_Hq=0, off_Wq=0, strideQ=None, itype="float", return_full=False,
                topk_mode="all"):
    vid0, vid1 = _shape(vid0, nheads), _shape(vid1, nheads)
    dev = vid0.device
    B, HD, T, F, qH, qW = vid0.shape
    kH, kW = vid1.shape[-2:]
    if flows.ndim == 6:
        flows = flows[:, None]
    W_t = 2 * wt + 1
    nH, nW = (kH - 1) // stride0 + 1, (kW - 1) // stride0 + 1
    Q = T * nH * nW
    dist_type_i, descending, dval = cref.dist_menu(dist_type)
    patch_offset = 0 if use_adj else -(ps // 2)
    dists = th.full((B, HD, Q, W_t, ws, ws), dval, dtype=th.float32, device=dev)
    if itype == "int":
        flows_k = (flows.round().int() if flows.dtype != th.int32 else flows).contiguous()
        inds = th.full((B, HD, Q, W_t, ws, ws, 3), int(-1e8), dtype=th.int32, device=dev)
        mod().non_local_search_int_forward(vid0, vid1, flows_k, dists, inds, ps, k, stride0,
                                           max(1, int(stride1)), stride0 if strideQ is None else strideQ,
                                           dilation, 1, ws if ws_interior <= 0 else ws_interior,
                                           reflect_bounds, full_ws, patch_offset, off_Hq, off_Wq, dist_type_i)
    else:
        flows_k = flows.float().contiguous()
        inds = th.full((B, HD, Q, W_t, ws, ws, 3), -1e8, dtype=th.float32, device=dev)
        mod().non_local_search_bilin2d_forward(vid0, vid1, flows_k, dists, inds, ps, k, stride0,
                                               float(stride1), dilation, 1, reflect_bounds, full_ws,
                                               patch_offset, off_Hq, off_Wq, dist_type_i)
    full = (dists.clone(), inds.clone()) if return_full else None
    def anchor_all(d, i):
        d3, i3 = d.view(B * HD, Q, -1), i.view(B * HD, Q, -1, 3)
        order = th.zeros((B * HD, Q), dtype=th.int32, device=dev)
        mod().anchor_self(d3, i3, order, stride0, nH, nW)

    def anchor_time(d, i, f):
        mod().anchor_self_time(d, i, f, wt, stride0, qH, qW, kH, kW)

    dists, inds = cref.select(dists, inds, flows_k, k, wt, stride0, (T, nH, nW, qH, qW, kH, kW), self_action,
                              topk_mode, descending, anchor_all, anchor_time)
    if return_full:
        return dists, inds, full
    return dists, inds


def search_backward(vid0, vid1, flows, grad_dists, grad_inds, inds, wt, ps, nheads=1, stride0=1,
                    dist_type="l2", dilation=1, reflect_bounds=True, use_adj=False, off_Hq=0, off_Wq=0,
                    strideQ=None, itype="float", flows_grad=False, kselect=None, reflect=None, refine=False,
                    dists=None):
    vid0, vid1 = _shape(vid0, nheads), _shape(vid1, nheads)
    dist_type_i = cref.dist_menu(dist_type)[0]
    patch_offset = 0 if use_adj else -(ps // 2)
    g0, g1 = th.zeros_like(vid0), th.zeros_like(vid1)
    grad_dists = grad_dists.contiguous().float()
    gflows = None
    if itype == "int":
        inds_k = (inds.round().int() if inds.is_floating_point() else inds).contiguous()
        mod().non_local_search_int_vid_backward(g0, g1, vid0, vid1, grad_dists, inds_k, ps, 1, stride0,
                                                stride0 if strideQ is None else strideQ, dilation,
                                                reflect_bounds, patch_offset, off_Hq, off_Wq, dist_type_i)
    elif not flows_grad:
        mod().non_local_search_bilin2d_vid_backward(g0, g1, vid0, vid1, grad_dists, inds.float().contiguous(),
                                                    wt, ps, 1, stride0, dilation, reflect_bounds, patch_offset,
                                                    off_Hq, off_Wq, dist_type_i)
    elif not refine:
        flows_k = flows.float().contiguous()
        if flows_k.ndim == 6:
            flows_k = flows_k[:, None].contiguous()
        gflows = th.zeros_like(flows_k)
        d = dists if dists is not None else th.zeros_like(grad_dists)
        mod().non_local_search_bilin2d_vidflows_backward(g0, g1, gflows, vid0, vid1, flows_k, grad_dists,
                                                         grad_inds.float().contiguous(), d.contiguous(),
                                                         inds.float().contiguous(), wt, ps, 1, stride0,
                                                         dilation, reflect_bounds, patch_offset, off_Hq, off_Wq,
                                                         dist_type_i)
    else:
        gflows = th.zeros(flows.shape, dtype=th.float32, device=vid0.device)
        mod().refinement_bilin2d_vidflows_backward(g0, g1, gflows, vid0, vid1, grad_dists,
                                                   grad_inds.float().contiguous(), inds.float().contiguous(),
                                                   kselect.int().contiguous(), reflect.bool().contiguous(),
                                                   wt, ps, 1, stride0, dilation, reflect_bounds, patch_offset,
                                                   off_Hq, off_Wq, dist_type_i)
    return g0, g1, gflows


def refine_forward(vid0, vid1, flows, ws, wt, wr, k, kr, ps, nheads=1, stride0=4, stride1=1, strideQ=None,
                   dilation=1, dist_type="l2", restricted_radius=True, reflect_bounds=True, full_ws=True,
                   self_action=None, use_adj=False, off_Hq=0, off_Wq=0, itype="float"):
    vid0, vid1 = _shape(vid0, nheads), _shape(vid1, nheads)
    dev = vid0.device
    B, HD, T, F, qH, qW = vid0.shape
    kH, kW = vid1.shape[-2:]
    nH, nW = (kH - 1) // stride0 + 1, (kW - 1) // stride0 + 1
    if kr is not None and kr > 0:
        Ksel = int(flows.shape[-2] * kr) if isinstance(kr, float) else int(kr)
        flows = flows[..., :Ksel, :]
    flows = flows.contiguous()
    Ks = flows.shape[-2]
    Q = T * nH * nW
    dist_type_i, descending, dval = cref.dist_menu(dist_type)
    patch_offset = 0 if use_adj else -(ps // 2)
    dists = th.full((B, HD, T, nH, nW, Ks, wr, wr), dval, dtype=th.float32, device=dev)
    if itype == "int":
        inds = th.full(dists.shape + (3,), int(-1e8), dtype=th.int32, device=dev)
        flows_k = (flows.round().int() if flows.is_floating_point() else flows.int()).contiguous()
        kselect = reflect = None
        mod().refinement_int_forward(vid0, vid1, flows_k, dists, inds, ws, ps, stride0, max(1, int(stride1)),
                                     stride0 if strideQ is None else strideQ, dilation, 1, restricted_radius,
                                     reflect_bounds, full_ws, patch_offset, off_Hq, off_Wq, dist_type_i)
    else:
        inds = th.full(dists.shape + (3,), -1e8, dtype=th.float32, device=dev)
        flows_k = flows.float().contiguous()
        kselect = th.ones(dists.shape, dtype=th.int32, device=dev)
        reflect = th.zeros(flows_k.shape[:-1] + (2,), dtype=th.bool, device=dev)
        mod().refinement_bilin2d_forward(vid0, vid1, flows_k, dists, inds, kselect, reflect, ws, ps, stride0,
                                         float(stride1), dilation, 1, restricted_radius, reflect_bounds,
                                         full_ws, patch_offset, off_Hq, off_Wq, dist_type_i)
    th.cuda.synchronize()
    anchor = self_action is not None and "anchor" in self_action
    if anchor:
        mod().anchor_self_refine(dists.view(B, HD, Q, Ks, wr * wr), inds.view(B, HD, Q, Ks, wr * wr, 3),
                                 flows_k.view(B, flows_k.shape[1], Q, Ks, 3), stride0, qH, qW, kH, kW)
    L = Ks * wr * wr
    dists, inds, order = cref.topk(dists.view(B, HD, Q, L), inds.view(B, HD, Q, L, 3), k, anchor, descending)
    if kselect is not None:
        kselect = kselect.view(B, HD, Q, L)
        if order is not None:
            kselect = th.gather(kselect, -1, order)
        kselect = kselect.reshape(B, HD, T, nH, nW, -1).contiguous()
    Kout = dists.shape[-1]
    return (dists.reshape(B, HD, T, nH, nW, Kout).contiguous(),
            inds.reshape(B, HD, T, nH, nW, Kout, 3).contiguous(), kselect, reflect)


def gather_add_forward(vid, weights, flows, ps, stride0, dilation=1, reflect_bounds=True, use_adj=False,
                       itype="float"):
    vid, w, f = cref._agg_prep(vid, weights, flows, itype)
    B, HD, T, F, H, W = vid.shape
    out = th.zeros_like(vid)
    counts = th.zeros((H, W), dtype=th.int32, device=vid.device)
    patch_offset = 0 if use_adj else -(ps // 2)
    mod().gather_add_forward(out, counts, vid, w, f, ps, stride0, stride0, 1, dilation, reflect_bounds,
                             patch_offset, itype == "int")
    out = out / (counts.view(1, 1, 1, 1, H, W) + 1e-10)
    return out, counts


def gather_add_backward(grad_out, vid, weights, flows, counts, ps, stride0, dilation=1, reflect_bounds=True,
                        use_adj=False, itype="float"):
    vid, w, f = cref._agg_prep(vid, weights, flows, itype)
    B, HD, T, F, H, W = vid.shape
    g = (grad_out.float() / counts.view(1, 1, 1, 1, H, W)).contiguous()
    gvid, gw = th.zeros_like(vid), th.zeros_like(w)
    patch_offset = 0 if use_adj else -(ps // 2)
    gf = None
    if itype == "int":
        mod().gather_add_int_backward(gvid, gw, g, vid, w, f, ps, stride0, stride0, 1, dilation, reflect_bounds,
                                      patch_offset)
    else:
        gf = th.zeros(f.shape, dtype=th.float32, device=vid.device)
        mod().gather_add_bilin2d_backward(gvid, gw, gf, g, vid, w, f, ps, stride0, stride0, 1, dilation,
                                          reflect_bounds, patch_offset)
        gf = gf.reshape(weights.shape + (3,))
    return gvid, gw.reshape(weights.shape), gf


def gather_forward(vid, weights, inds, ps, stride0, dilation=1, reflect_bounds=True, use_adj=False, itype="float"):
    """Drives gather_{int,bilin2d}_forward as lib/stnls/agg/gather.py:79-160 does."""
    vid, w, f = cref._agg_prep(vid, weights, inds, itype)
    B, HD, T, F, H, W = vid.shape
    K = w.shape[-1]
    stack = th.zeros((B, HD, K, T, F, H, W), device=vid.device, dtype=vid.dtype)
    counts = th.zeros((B, HD, H, W), device=vid.device, dtype=th.int32)
    patch_offset = 0 if use_adj else -(ps // 2)
    fn = mod().gather_int_forward if itype == "int" else mod().gather_bilin2d_forward
    fn(vid, w, f, stack, counts, ps, 1, dilation, stride0, reflect_bounds, patch_offset)
    stack = stack / (counts.view(B, HD, 1, 1, 1, H, W) + 1e-10)
    return stack, counts


def gather_backward(grad_stack, vid, weights, inds, counts, ps, stride0, dilation=1, reflect_bounds=True,
                    use_adj=False, itype="float"):
    """Drives gather_{int,bilin2d}_backward as lib/stnls/agg/gather.py:162-253 does."""
    vid, w, f = cref._agg_prep(vid, weights, inds, itype)
    B, HD, T, F, H, W = vid.shape
    g = (grad_stack.float() / (counts.view(B, HD, 1, 1, 1, H, W) + 1e-10)).contiguous()
    gvid, gw = th.zeros_like(vid), th.zeros_like(w)
    patch_offset = 0 if use_adj else -(ps // 2)
    stack = th.zeros_like(g)  # saved forward output; the kernels do not read it
    gi = None
    if itype == "int":
        mod().gather_int_backward(gvid, gw, g, vid, w, f, stack, counts, ps, 1, dilation, stride0,
                                  reflect_bounds, patch_offset)
    else:
        gi = th.zeros(f.shape, dtype=th.float32, device=vid.device)
        mod().gather_bilin2d_backward(gvid, gw, gi, g, vid, w, f, stack, counts, ps, 1, dilation, stride0,
                                      reflect_bounds, patch_offset)
        gi = gi.reshape(weights.shape + (3,))
    return gvid, gw.reshape(weights.shape), gi


def search_flow_forward(fflow, bflow, wt, stride0=1):
    """Drives search_flow_forward as lib/stnls/nn/search_flow.py:47-67 does."""
    fflow, bflow = fflow.contiguous(), bflow.contiguous()
    B, T, _, H, W = fflow.shape
    nH, nW = (H - 1) // stride0 + 1, (W - 1) // stride0 + 1
    W_t = min(2 * wt + 1, T)
    flows = th.zeros((B, T, W_t - 1, 2, nH, nW), device=fflow.device, dtype=fflow.dtype)
    mod().search_flow_forward(fflow, bflow, flows, wt, stride0)
    return flows


def search_flow_backward(grad_flows, fflow, bflow, flows, wt, stride0=1):
    """lib/stnls/nn/search_flow.py:69-96"""
    gf, gb = th.zeros_like(fflow), th.zeros_like(bflow)
    mod().search_flow_backward(gf, gb, grad_flows.contiguous(), fflow.contiguous(), bflow.contiguous(),
                               flows.contiguous(), wt, stride0)
    return gf, gb
