"""
oracle/ref.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Python driver around oracle/stnls_oracle.c that restates the reference's *pipeline* for the hot
path: allocate every candidate, run the "kernel", anchor, top-K, reshape.  Follows

  lib/stnls/search/impl/non_local_search.py:23-212      (forward / backward)
  lib/stnls/search/impl/refinement.py:18-181
  lib/stnls/nn/topk.py:65-123,196-220                   (anchored / standard top-K)
  lib/stnls/nn/anchor_self.py:18-67
  lib/stnls/agg/gather_add.py:34-188                    (WeightedPatchSum == NonLocalGatherAdd)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  The one declared deviation: `th.argsort` in the reference is not stable, so the
order among *exactly equal* distances is unspecified there; here (and in the CUDA path) ties break
towards the lower linear candidate index (stable sort).
"""
import ctypes
import os
import subprocess

import torch as th

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libstnls_oracle.so")
_LIB = None


def build(force=False):
    src = os.path.join(_HERE, "stnls_oracle.c")
    if (not force) and os.path.exists(_SO) and os.path.getmtime(_SO) >= os.path.getmtime(src):
        return _SO
    os.makedirs(os.path.dirname(_SO), exist_ok=True)
    tmp = _SO + f".tmp{os.getpid()}"   # built aside and moved into place: concurrent ranks never load a partial file
    cmd = ["gcc", "-O2", "-fPIC", "-shared", "-fopenmp", "-ffp-contract=off", "-std=gnu11",
           "-o", tmp, src, "-lm"]
    subprocess.check_call(cmd)
    os.replace(tmp, _SO)
    return _SO


def lib():
    global _LIB
    if _LIB is None:
        build()
        _LIB = ctypes.CDLL(_SO)
    return _LIB


class NlsArgs(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int) for n in
                ["B", "HD", "HD_f", "T", "F", "qH", "qW", "kH", "kW", "nH", "nW",
                 "ws", "wt", "ps", "stride0", "strideQ", "dilation", "ws_interior",
                 "reflect_bounds", "full_ws", "patch_offset", "off_Hq", "off_Wq",
                 "dist_type", "st_offset"]] + [("stride1", ctypes.c_float)]


def _p(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


def shape_vid(vid, nheads):
    """lib/stnls/search/utils.py:178-190"""
    if vid.ndim == 5:
        B, T, Fc, H, W = vid.shape
        assert Fc % nheads == 0
        vid = vid.reshape(B, T, nheads, Fc // nheads, H, W).permute(0, 2, 1, 3, 4, 5)
    return vid.contiguous().float()


def dist_menu(dist_type):
    """lib/stnls/search/utils.py:216-235"""
    return {"prod": (0, True, -float("inf")), "l2": (1, False, float("inf"))}[dist_type]


def _mk_args(vid0, vid1, HD_f, nH, nW, ws, wt, ps, stride0, stride1, strideQ, dilation,
             ws_interior, reflect_bounds, full_ws, patch_offset, off_Hq, off_Wq, dist_type_i,
             st_offset):
    B, HD, T, F, qH, qW = vid0.shape
    kH, kW = vid1.shape[-2:]
    return NlsArgs(B, HD, HD_f, T, F, qH, qW, kH, kW, nH, nW, ws, wt, ps, stride0, strideQ,
                   dilation, ws_interior, int(reflect_bounds), int(full_ws), patch_offset,
                   off_Hq, off_Wq, dist_type_i, st_offset, float(stride1))


# ---------------------------------------------------------------------------------------------
#   top-K  (lib/stnls/nn/topk.py)
# ---------------------------------------------------------------------------------------------

def standard_topk(dists, inds, K, descending):
    order = th.sort(dists, dim=-1, descending=descending, stable=True)[1][..., :K]
    d = th.gather(dists, -1, order)
    i = th.gather(inds, -2, order[..., None].expand(*order.shape, inds.shape[-1]))
    return d, i, order


def topk(dists, inds, k, anchor, descending):
    """dists (...,L) inds (...,L,3) -> (...,K); lib/stnls/nn/topk.py:65-123"""
    if not (k > 0):
        return dists, inds, None
    if anchor:
        d, i, o = standard_topk(dists[..., 1:], inds[..., 1:, :], k - 1, descending)
        d = th.cat([dists[..., :1], d], -1)
        i = th.cat([inds[..., :1, :], i], -2)
        o = th.cat([th.zeros_like(o[..., :1]), o + 1], -1)
        return d, i, o
    return standard_topk(dists, inds, k, descending)


def topk_each(dists, inds, k, descending, anchor_self=False):
    """per-frame-group top-K: dists (...,G,S) inds (...,G,S,3) -> (...,G,K'); lib/stnls/nn/topk.py:226-269"""
    if k <= 0:
        return dists, inds
    if anchor_self:
        d0, i0 = dists[..., :1], inds[..., :1, :]
        if k > 1 and dists.shape[-1] > 1:
            d, i, _ = standard_topk(dists[..., 1:], inds[..., 1:, :], min(k - 1, dists.shape[-1] - 1), descending)
            return th.cat([d0, d], -1), th.cat([i0, i], -2)
        return d0, i0
    d, i, _ = standard_topk(dists, inds, min(k, dists.shape[-1]), descending)
    return d, i


SELF_ACTIONS = (None, "anchor", "anchor_self", "anchor_each", "remove", "remove_ref_frame",
                "anchor_and_remove_ref_frame")


def select(dists, inds, flows_k, k, wt, stride0, shapes, self_action, topk_mode, descending,
           anchor_all, anchor_time):
    """Everything after the search kernel: lib/stnls/search/impl/non_local_search.py:88-139.
    dists (B,HD,Q,W_t,ws,ws), inds (...,3); anchor_all(dists(N,L), inds) and anchor_time(dists(B,HD,Q,G,S), inds,
    flows) work in place (the C oracle here, the reference's kernels in oracle/ref_cuda.py)."""
    B, HD, Q, W_t, ws, _ = dists.shape
    T, nH, nW, qH, qW, kH, kW = shapes
    assert self_action in SELF_ACTIONS
    anchor = self_action is not None and "anchor" in self_action
    dists = dists.reshape(B, HD, Q, W_t, ws * ws).contiguous()
    inds = inds.reshape(B, HD, Q, W_t, ws * ws, 3).contiguous()
    if self_action in ("anchor", "anchor_self", "remove"):
        anchor_all(dists.view(B * HD * Q, -1), inds.view(B * HD * Q, -1, 3))
    if self_action == "anchor_each":
        anchor_time(dists, inds, flows_k)
    elif self_action == "remove":
        dists = dists.view(B, HD, Q, -1)[..., 1:].contiguous()
        inds = inds.view(B, HD, Q, -1, 3)[..., 1:, :].contiguous()
    elif self_action in ("remove_ref_frame", "anchor_and_remove_ref_frame"):
        assert wt > 0, "Cannot remove ref frame if not searching across time."
        dists = dists[..., 1:, :].contiguous()
        inds = inds[..., 1:, :, :].contiguous()
        if self_action == "anchor_and_remove_ref_frame":
            anchor_time(dists, inds, flows_k)
    if topk_mode == "all":
        dists, inds, _ = topk(dists.reshape(B, HD, Q, -1), inds.reshape(B, HD, Q, -1, 3), k, anchor, descending)
    elif topk_mode == "each":
        assert dists.dim() == 5, "topk_mode='each' needs the per-frame layout"
        dists, inds = topk_each(dists, inds, k, descending, anchor_self=anchor)
    elif topk_mode == "none":
        assert k <= 0, "If topk_mode is 'none' then k must be <= 0"
    else:
        raise ValueError(f"Unknown topk_mode [{topk_mode}]")
    dists = dists.reshape(B, HD, T, nH, nW, -1).contiguous()
    inds = inds.reshape(B, HD, T, nH, nW, -1, 3).contiguous()
    return dists, inds


# ---------------------------------------------------------------------------------------------
#   NonLocalSearch
# ---------------------------------------------------------------------------------------------

def nls_forward(vid0, vid1, flows, ws, wt, ps, k, nheads=1, stride0=1, stride1=1,
                dist_type="l2", dilation=1, self_action=None, ws_interior=0,
                reflect_bounds=True, full_ws=True, use_adj=False, off_Hq=0, off_Wq=0,
                strideQ=None, itype="float", return_full=False, topk_mode="all"):
    """lib/stnls/search/non_local_search.py:33-105 + impl/non_local_search.py:23-140"""
    vid0, vid1 = shape_vid(vid0, nheads), shape_vid(vid1, nheads)
    B, HD, T, F, qH, qW = vid0.shape
    kH, kW = vid1.shape[-2:]
    if flows.ndim == 6:
        flows = flows[:, None]
    W_t = 2 * wt + 1
    assert T >= W_t
    nH, nW = (kH - 1) // stride0 + 1, (kW - 1) // stride0 + 1
    assert flows.shape[-2:] == (nH, nW)
    assert flows.shape[3] in (W_t - 1, W_t)
    Q = T * nH * nW
    dist_type_i, descending, dval = dist_menu(dist_type)
    patch_offset = 0 if use_adj else -(ps // 2)
    itype_int = itype == "int"
    dists = th.full((B, HD, Q, W_t, ws, ws), dval, dtype=th.float32)
    if itype_int:
        flows_k = flows.round().int().contiguous() if flows.dtype != th.int32 else flows.contiguous()
        inds = th.full((B, HD, Q, W_t, ws, ws, 3), int(-1e8), dtype=th.int32)
        stride1_k = max(1, int(stride1))
        strideQ_k = stride0 if strideQ is None else strideQ
        wsi = ws if ws_interior <= 0 else ws_interior
    else:
        flows_k = flows.float().contiguous()
        inds = th.full((B, HD, Q, W_t, ws, ws, 3), -1e8, dtype=th.float32)
        stride1_k = float(stride1)
        strideQ_k = stride0
        wsi = ws
    a = _mk_args(vid0, vid1, flows_k.shape[1], nH, nW, ws, wt, ps, stride0, stride1_k, strideQ_k,
                 dilation, wsi, reflect_bounds, full_ws, patch_offset, off_Hq, off_Wq,
                 dist_type_i, W_t - flows_k.shape[3])
    lib().oracle_nls_forward(_p(vid0), _p(vid1), _p(flows_k), _p(dists), _p(inds),
                             ctypes.byref(a), int(itype_int))
    full = (dists.clone(), inds.clone()) if return_full else None
    def anchor_all(d, i):
        lib().oracle_anchor_self(_p(d), _p(i), ctypes.c_int64(d.shape[0]), d.shape[1], int(itype_int))

    def anchor_time(d, i, f):
        assert d.is_contiguous() and i.is_contiguous() and f.is_contiguous()
        lib().oracle_anchor_self_time(_p(d), _p(i), _p(f), B, HD, f.shape[1], f.shape[2], d.shape[3], d.shape[4],
                                      f.shape[3], wt, stride0, qH, qW, kH, kW, int(itype_int))

    dists, inds = select(dists, inds, flows_k, k, wt, stride0, (T, nH, nW, qH, qW, kH, kW), self_action,
                         topk_mode, descending, anchor_all, anchor_time)
    if return_full:
        return dists, inds, full
    return dists, inds


def search_backward(vid0, vid1, flows, grad_dists, grad_inds, inds, wt, ps, nheads=1, stride0=1,
                    dist_type="l2", dilation=1, reflect_bounds=True, use_adj=False,
                    off_Hq=0, off_Wq=0, strideQ=None, itype="float", flows_grad=False,
                    kselect=None, reflect=None, refine=False):
    """impl/non_local_search.py:142-212 ; impl/refinement.py:114-181.
    Returns grads in the 6-D layout (B,HD,T,F,H,W) plus grad_flows (or None)."""
    vid0, vid1 = shape_vid(vid0, nheads), shape_vid(vid1, nheads)
    B, HD, T, F, qH, qW = vid0.shape
    nH, nW, K = inds.shape[3:6]
    dist_type_i = dist_menu(dist_type)[0]
    patch_offset = 0 if use_adj else -(ps // 2)
    g0, g1 = th.zeros_like(vid0), th.zeros_like(vid1)
    grad_dists = grad_dists.contiguous().float()
    Ks, st_offset, HD_f, gflows, flows_k = 0, 0, HD, None, None
    if itype == "int":
        mode = 0
        inds_k = inds.round().int().contiguous() if inds.is_floating_point() else inds.contiguous()
    else:
        inds_k = inds.float().contiguous()
        mode = 1
        if flows_grad:
            mode = 3 if refine else 2
            flows_k = flows.float().contiguous()
            if not refine and flows_k.ndim == 6:
                flows_k = flows_k[:, None].contiguous()
            HD_f = flows_k.shape[1]
            gflows = th.zeros_like(flows_k)
            grad_inds = grad_inds.contiguous().float()
            if refine:
                Ks = flows_k.shape[5]
                kselect = kselect.int().contiguous()
                reflect = reflect.to(th.uint8).contiguous()
            else:
                st_offset = (2 * wt + 1) - flows_k.shape[3]
    strideQ_k = stride0 if strideQ is None else strideQ
    a = _mk_args(vid0, vid1, HD_f, nH, nW, 1, wt, ps, stride0, 1.0, strideQ_k, dilation, 1,
                 reflect_bounds, True, patch_offset, off_Hq, off_Wq, dist_type_i, st_offset)
    lib().oracle_search_backward(_p(g0), _p(g1), _p(gflows), _p(vid0), _p(vid1), _p(flows_k),
                                 _p(grad_dists), _p(grad_inds if mode >= 2 else None), _p(inds_k),
                                 _p(kselect if mode == 3 else None),
                                 _p(reflect if mode == 3 else None),
                                 ctypes.byref(a), K, Ks, mode)
    return g0, g1, gflows


# ---------------------------------------------------------------------------------------------
#   RefineSearch
# ---------------------------------------------------------------------------------------------

def refine_forward(vid0, vid1, flows, ws, wt, wr, k, kr, ps, nheads=1, stride0=4, stride1=1,
                   strideQ=None, dilation=1, dist_type="l2", restricted_radius=True,
                   reflect_bounds=True, full_ws=True, self_action=None, use_adj=False,
                   off_Hq=0, off_Wq=0, itype="float", topk_mode="all"):
    """lib/stnls/search/refinement.py:32-101 + impl/refinement.py:18-112"""
    vid0, vid1 = shape_vid(vid0, nheads), shape_vid(vid1, nheads)
    B, HD, T, F, qH, qW = vid0.shape
    kH, kW = vid1.shape[-2:]
    nH, nW = (kH - 1) // stride0 + 1, (kW - 1) // stride0 + 1
    if flows.ndim == 5:
        flows = flows.reshape(B, flows.shape[1], T, nH, nW, flows.shape[-2], 3)
    assert flows.shape[1] == HD
    # filter_k: lib/stnls/search/utils.py:115-125
    if kr is not None and kr > 0:
        Ksel = int(flows.shape[-2] * kr) if isinstance(kr, float) else int(kr)
        flows = flows[..., :Ksel, :]
    flows = flows.contiguous()
    Ks = flows.shape[-2]
    Q = T * nH * nW
    dist_type_i, descending, dval = dist_menu(dist_type)
    patch_offset = 0 if use_adj else -(ps // 2)
    itype_int = itype == "int"
    dists = th.full((B, HD, T, nH, nW, Ks, wr, wr), dval, dtype=th.float32)
    if itype_int:
        inds = th.full(dists.shape + (3,), int(-1e8), dtype=th.int32)
        flows_k = flows.round().int().contiguous() if flows.is_floating_point() else flows.int().contiguous()
        kselect, reflect = None, None
        stride1_k = max(1, int(stride1))
        strideQ_k = stride0 if strideQ is None else strideQ
        # flag permutation quirk, lib/csrc/search/refinement_int_kernel.cu:25 vs :232-233
        k_reflect, k_full = bool(restricted_radius), bool(reflect_bounds)
    else:
        inds = th.full(dists.shape + (3,), -1e8, dtype=th.float32)
        flows_k = flows.float().contiguous()
        kselect = th.ones(dists.shape, dtype=th.int32)
        reflect = th.zeros(flows_k.shape[:-1] + (2,), dtype=th.uint8)
        stride1_k = float(stride1)
        strideQ_k = stride0
        k_reflect, k_full = bool(reflect_bounds), bool(full_ws)
    a = _mk_args(vid0, vid1, flows_k.shape[1], nH, nW, ws, wt, ps, stride0, stride1_k, strideQ_k,
                 dilation, ws, k_reflect, k_full, patch_offset, off_Hq, off_Wq, dist_type_i, 0)
    lib().oracle_refine_forward(_p(vid0), _p(vid1), _p(flows_k), _p(dists), _p(inds),
                                _p(kselect), _p(reflect), ctypes.byref(a), Ks, wr, int(itype_int))
    anchor = self_action is not None and "anchor" in self_action
    if anchor:
        lib().oracle_anchor_self_refine(_p(dists), _p(inds), _p(flows_k), B, HD, flows_k.shape[1],
                                        Q, Ks, wr * wr, stride0, qH, qW, kH, kW, int(itype_int))
    L = Ks * wr * wr
    assert self_action in (None, "anchor", "anchor_self", "anchor_each")  # impl/refinement.py:83
    if topk_mode == "each":  # impl/refinement.py:99-106: top-k inside every group of wr*wr candidates
        dists, inds = topk_each(dists.view(B, HD, Q, Ks, wr * wr), inds.view(B, HD, Q, Ks, wr * wr, 3), k,
                                descending, anchor_self=anchor)
        if kselect is not None:  # "all same across dim": every entry of a group carries the group index
            kselect = kselect.view(B, HD, Q, Ks, wr * wr)[..., :dists.shape[-1]]
            kselect = kselect.reshape(B, HD, T, nH, nW, -1).contiguous()
    else:
        assert topk_mode == "all", f"Unknown topk_mode [{topk_mode}]"
        dists = dists.view(B, HD, Q, L)
        inds = inds.view(B, HD, Q, L, 3)
        dists, inds, order = topk(dists, inds, k, anchor, descending)
        if kselect is not None:
            kselect = kselect.view(B, HD, Q, L)
            if order is not None:
                kselect = th.gather(kselect, -1, order)
            kselect = kselect.reshape(B, HD, T, nH, nW, -1).contiguous()
    dists, inds = dists.reshape(B, HD, Q, -1), inds.reshape(B, HD, Q, -1, 3)
    Kout = dists.shape[-1]
    dists = dists.reshape(B, HD, T, nH, nW, Kout).contiguous()
    inds = inds.reshape(B, HD, T, nH, nW, Kout, 3).contiguous()
    return dists, inds, kselect, (reflect.bool() if reflect is not None else None)


def normz_bwd(grad_vid0, grad_vid1, ps, stride0, stride1):
    """normalize_bwd=True: lib/stnls/search/shared.py:59-97 (fold of ones, cropped, divides the grads in place)"""
    import torch.nn.functional as F
    assert int(stride1) == stride1, "stride1 must be an int."
    B, T, C, H, W = grad_vid0.shape
    for g, stride in ((grad_vid0, int(stride0)), (grad_vid1, int(stride1))):
        nH, nW = (H - 1) // stride + 1, (W - 1) // stride + 1
        counts = th.ones((1, ps * ps, nH * nW))
        pad = (ps - 1) // 2
        Hp, Wp = H + 2 * pad, W + 2 * pad
        sH, sW = (Hp - H + 1) // 2, (Wp - W + 1) // 2
        counts = F.fold(counts, (Hp, Wp), [ps] * 2, 1, [0, 0], stride)
        g /= counts[..., sH:sH + H, sW:sW + W]
    return grad_vid0, grad_vid1


# ---------------------------------------------------------------------------------------------
#   WeightedPatchSum == NonLocalGatherAdd
# ---------------------------------------------------------------------------------------------

def _agg_prep(vid, weights, flows, itype):
    B, HD, T, nH, nW, K = weights.shape
    vid = shape_vid(vid, HD)
    if flows.ndim == 6:
        flows = flows[:, None]
    if itype == "int":
        flows = flows.round().int() if flows.is_floating_point() else flows.int()
    else:
        flows = flows.float()
    Q = T * nH * nW
    return vid, weights.reshape(B, HD, Q, K).contiguous().float(), flows.reshape(B, HD, Q, K, 3).contiguous()


def gather_add_forward(vid, weights, flows, ps, stride0, dilation=1, reflect_bounds=True,
                       use_adj=False, itype="float"):
    """lib/stnls/agg/gather_add.py:36-114 ; returns (out (B,HD,T,F,H,W), counts (H,W))"""
    vid, w, f = _agg_prep(vid, weights, flows, itype)
    B, HD, T, F, H, W = vid.shape
    Q, K = w.shape[2:]
    assert weights.shape[3] == (H - 1) // stride0 + 1 and weights.shape[4] == (W - 1) // stride0 + 1
    out = th.zeros_like(vid)
    counts = th.zeros((H, W), dtype=th.int32)
    patch_offset = 0 if use_adj else -(ps // 2)
    lib().oracle_gather_add_forward(_p(out), _p(counts), _p(vid), _p(w), _p(f), B, HD, T, F, H, W,
                                    Q, K, ps, stride0, dilation, int(reflect_bounds), patch_offset,
                                    int(itype == "int"))
    assert th.all(counts > 1e-3)
    out = out / (counts.view(1, 1, 1, 1, H, W) + 1e-10)
    return out, counts


def gather_add_backward(grad_out, vid, weights, flows, counts, ps, stride0, dilation=1,
                        reflect_bounds=True, use_adj=False, itype="float"):
    """lib/stnls/agg/gather_add.py:116-188 ; grad_out (B,HD,T,F,H,W)"""
    vid, w, f = _agg_prep(vid, weights, flows, itype)
    B, HD, T, F, H, W = vid.shape
    Q, K = w.shape[2:]
    g = (grad_out.float() / counts.view(1, 1, 1, 1, H, W)).contiguous()
    gvid, gw = th.zeros_like(vid), th.zeros_like(w)
    gf = th.zeros(f.shape, dtype=th.float32) if itype != "int" else None
    patch_offset = 0 if use_adj else -(ps // 2)
    lib().oracle_gather_add_backward(_p(gvid), _p(gw), _p(gf), _p(g), _p(vid), _p(w), _p(f),
                                     B, HD, T, F, H, W, Q, K, ps, stride0, dilation,
                                     int(reflect_bounds), patch_offset, int(itype == "int"))
    gw = gw.reshape(weights.shape)
    if gf is not None:
        gf = gf.reshape(weights.shape + (3,))
    return gvid, gw, gf


# ---------------------------------------------------------------------------------------------
#   NonLocalGather (K-stacked aggregation)
# ---------------------------------------------------------------------------------------------

def gather_forward(vid, weights, inds, ps, stride0, dilation=1, reflect_bounds=True, use_adj=False,
                   itype="float"):
    """lib/stnls/agg/gather.py:79-160 ; returns (stack (B,HD,K,T,F,H,W), counts (H,W))"""
    vid, w, f = _agg_prep(vid, weights, inds, itype)
    B, HD, T, F, H, W = vid.shape
    Q, K = w.shape[2:]
    stack = th.zeros((B, HD, K, T, F, H, W), dtype=th.float32)
    counts = th.zeros((H, W), dtype=th.int32)
    patch_offset = 0 if use_adj else -(ps // 2)
    lib().oracle_gather_forward(_p(stack), _p(counts), _p(vid), _p(w), _p(f), B, HD, T, F, H, W,
                                Q, K, ps, stride0, dilation, int(reflect_bounds), patch_offset,
                                int(itype == "int"))
    assert th.all(counts > 0)
    stack = stack / (counts.view(1, 1, 1, 1, 1, H, W) + 1e-10)
    return stack, counts


def gather_backward(grad_stack, vid, weights, inds, counts, ps, stride0, dilation=1, reflect_bounds=True,
                    use_adj=False, itype="float"):
    """lib/stnls/agg/gather.py:162-253 ; grad_stack (B,HD,K,T,F,H,W)"""
    vid, w, f = _agg_prep(vid, weights, inds, itype)
    B, HD, T, F, H, W = vid.shape
    Q, K = w.shape[2:]
    g = (grad_stack.float() / (counts.view(1, 1, 1, 1, 1, H, W) + 1e-10)).contiguous()
    gvid, gw = th.zeros_like(vid), th.zeros_like(w)
    gi = th.zeros(f.shape, dtype=th.float32) if itype != "int" else None
    patch_offset = 0 if use_adj else -(ps // 2)
    lib().oracle_gather_backward(_p(gvid), _p(gw), _p(gi), _p(g), _p(vid), _p(w), _p(f),
                                 B, HD, T, F, H, W, Q, K, ps, stride0, dilation,
                                 int(reflect_bounds), patch_offset, int(itype == "int"))
    gw = gw.reshape(weights.shape)
    if gi is not None:
        gi = gi.reshape(weights.shape + (3,))
    return gvid, gw, gi


# ---------------------------------------------------------------------------------------------
#   search_flow: lib/stnls/nn/search_flow.py:24-96
# ---------------------------------------------------------------------------------------------

def search_flow_forward(fflow, bflow, wt, stride0=1):
    fflow, bflow = fflow.contiguous().float(), bflow.contiguous().float()
    B, T, _, H, W = fflow.shape
    nH, nW = (H - 1) // stride0 + 1, (W - 1) // stride0 + 1
    W_t = min(2 * wt + 1, T)
    flows = th.zeros((B, T, W_t - 1, 2, nH, nW), dtype=th.float32)
    if wt > 0:
        lib().oracle_search_flow_forward(_p(flows), _p(fflow), _p(bflow), B, T, H, W, wt, stride0)
    return flows


def search_flow_backward(grad_flows, fflow, bflow, flows, wt, stride0=1):
    fflow, bflow = fflow.contiguous().float(), bflow.contiguous().float()
    B, T, _, H, W = fflow.shape
    gf, gb = th.zeros_like(fflow), th.zeros_like(bflow)
    lib().oracle_search_flow_backward(_p(gf), _p(gb), _p(grad_flows.contiguous().float()), _p(fflow), _p(bflow),
                                      _p(flows.contiguous().float()), B, T, H, W, wt, stride0)
    return gf, gb
