"""One autograd node for the whole README step: NonLocalSearch -> SoftmaxNormalize -> WeightedPatchSum.

The three reference modules (lib/stnls/search/non_local_search.py:122-218, lib/stnls/normz/softmax.py:12-47,
lib/stnls/agg/gather_add.py:34-221) chained through autograd cost 14 kernel launches per forward + backward: every
module packs its own inputs, the normalise step is a kernel of its own in both directions and every gradient is
unpacked separately.  Chained HERE the same kernels run in 8 launches:

    forward   pack(q, k, v)  ->  search  ->  aggregation (normalise folded in)  ->  unpack(out)
    backward  pack(grad_out / counts)  ->  aggregation backward  ->  search backward (normalise backward folded in)
              ->  unpack(grad_q, grad_k, grad_v)

The folded normalise steps use the arithmetic of the stand-alone kernels in the same order (csrc/normz.cu), so the
fused and the unfused routes agree bit for bit (tests/test_fused_step_gpu.py).  `dists` and `inds` are returned for
inspection but carry no gradient on this route; a caller that differentiates through them uses the modules directly.
`fusable()` says whether a (search, agg) pair is inside the supported set; everything else keeps the module chain.
"""
import os

import torch as th

from . import ops
from ._C import AggOpts, SearchOpts
from .agg.gather_add import NonLocalGatherAdd
from .search.non_local_search import DIST_MENU, FUSED_ACTIONS, NonLocalSearch, shard_of


def fusable(search, agg, q, k, v, flows):
    if os.environ.get("STNLS_FUSED_STEP", "1") == "0":
        return False
    if not isinstance(search, NonLocalSearch) or not isinstance(agg, NonLocalGatherAdd):
        return False
    s, a = search, agg
    anchor = s.self_action in ("anchor", "anchor_self")
    ok = (s.itype == "float" and s.topk_mode == "all" and s.self_action in FUSED_ACTIONS and s.pt == 1
          and s.ws_interior in (0, s.ws) and s.k > 0 and s.k - (1 if anchor else 0) <= ops.MAX_FUSED_K
          and s.k <= ops.MAX_SOFTMAX_K and s.k_agg in (-1, 0, s.k) and not s.normalize_bwd and s.strideQ is None
          and s.off_Hq == 0 and s.off_Wq == 0 and s.k <= (2 * s.wt + 1) * s.ws * s.ws)
    ok = ok and (a.itype == "float" and a.pt == 1 and a.strideIn == a.strideOut == s.stride0 and a.outH == 0 and a.outW == 0)
    ok = ok and all(x.dim() == 5 and x.dtype == th.float32 and x.is_cuda for x in (q, k, v))
    ok = ok and q.shape[-2:] == k.shape[-2:] == v.shape[-2:] and q.shape[2] == k.shape[2]
    ok = ok and q.shape[2] % s.nheads == 0 and v.shape[2] % s.nheads == 0 and k.shape[1] == v.shape[1]
    ok = ok and flows.is_floating_point() and flows.dim() in (6, 7)
    return bool(ok)


class FusedSearchAggFunction(th.autograd.Function):
    """q (B,Tq,C,H,W): owned frames; k, v (B,Tk,C,H,W): owned + halo frames; flows (B,[HD_f,]Tq,W_t[-1],2,nH,nW).
    Returns out (B,HD,Tq,Fv,H,W), dists (B,HD,Tq,nH,nW,K), inds (...,K,3)."""

    @staticmethod
    def forward(ctx, q, k, v, flows, search, agg, scale, shard, halo):
        """halo: None (k, v hold the owned + halo frames) or a parallel.PeerHaloCL (k, v hold the OWNED frames; the halo
        frames are exchanged here, in the channels-last layout)."""
        s, a = search, agg
        HD = s.nheads
        B, Tq, C, H, W = q.shape
        Tk = k.shape[1] if halo is None else halo.plan.fb - halo.plan.fa
        F, Fv = C // HD, v.shape[2] // HD
        T, tq0, tk0 = shard_of(shard, Tq)
        W_t = 2 * s.wt + 1
        assert T >= W_t, f"Num Frames [{T}] must be >= Temporal Window [{W_t}]"
        flow_ndim = flows.dim()
        if flow_ndim == 6:
            flows = flows[:, None]
        nH, nW = (H - 1) // s.stride0 + 1, (W - 1) // s.stride0 + 1
        assert flows.shape[-2] == nH and flows.shape[-1] == nW and flows.shape[2] == Tq
        assert flows.shape[3] in [W_t - 1, W_t], "flows.shape: " + str(flows.shape) + " W_t: " + str(W_t)
        flows_k = flows.float().contiguous()
        dist_type_i, _ = DIST_MENU[s.dist_type]
        anchor = s.self_action in ("anchor", "anchor_self")
        if halo is None:
            q_cl, k_cl, v_cl = ops.pack_cl_multi([q, k, v], HD)
        else:
            plan, N = halo.plan, B * HD
            Fp, Fvp = ops.padded_channels(F), ops.padded_channels(Fv)
            assert k.shape[1] == v.shape[1] == plan.b - plan.a, "k, v must hold this rank's owned frames"
            kbuf, khdl = halo.buffer("k", N, H * W * Fp, q.device)
            vbuf, vhdl = halo.buffer("v", N, H * W * Fvp, q.device)
            k_cl, v_cl = kbuf.view(N, Tk, H, W, Fp), vbuf.view(N, Tk, H, W, Fvp)
            own = plan.a - plan.fa
            q_cl = ops.pack_cl_multi([q, k, v], HD, outs=[None, (k_cl, own), (v_cl, own)])[0]
            halo.handshake(khdl)
            halo.pull([(kbuf, khdl), (vbuf, vhdl)])
        geom = ops.make_geom(B, HD, T, F, H, W, H, W, s.stride0, tq0, Tq, tk0, Tk)
        opts = SearchOpts(s.ws, s.wt, 0, 0, s.ps, s.k, s.stride0, s.stride0, s.dilation, float(s.stride1),
                          int(s.reflect_bounds), int(s.full_ws), 0 if s.use_adj else -(s.ps // 2), 0, 0,
                          dist_type_i, 0, int(anchor), flows_k.shape[1], W_t - flows_k.shape[3])
        dists, inds, _, _ = ops.search_forward(q_cl, k_cl, flows_k, geom, opts)
        K = dists.shape[-1]
        po = 0 if a.use_adj else -(a.ps // 2)
        ch, mh = ops.overlap_counts(H, nH, a.ps, a.strideIn, a.dilation, po, q.device)
        cw, mw = ops.overlap_counts(W, nW, a.ps, a.strideIn, a.dilation, po, q.device)
        assert mh > 0 and mw > 0, "every output pixel must be covered by a patch (reference: counts>1e-3)"
        ageom = ops.make_geom(B, HD, T, Fv, H, W, H, W, a.strideIn, tq0, Tq, tk0, Tk)
        aopts = AggOpts(a.ps, K, a.strideIn, a.dilation, int(a.reflect_bounds), po, 0)
        out_cl, weights = ops.wpsum_forward_sm(v_cl, dists, scale, inds, (ch, cw), ageom, aopts)
        out = ops.unpack_cl(out_cl, (B, HD, Tq, Fv, H, W), HD)
        ctx.save_for_backward(q_cl, k_cl, v_cl, inds, flows_k, weights, ch, cw)
        ctx.geom, ctx.opts, ctx.ageom, ctx.aopts = geom, opts, ageom, aopts
        ctx.shapes = (tuple(q.shape), tuple(k.shape), tuple(v.shape))
        ctx.HD, ctx.scale, ctx.flow_ndim, ctx.halo = HD, float(scale), flow_ndim, halo
        ctx.mark_non_differentiable(dists, inds)
        return out, dists, inds

    @staticmethod
    def backward(ctx, gout, _gd, _gi):
        q_cl, k_cl, v_cl, inds, flows_k, weights, ch, cw = ctx.saved_tensors
        flows_grad = ctx.needs_input_grad[3]
        gout_cl = ops.pack_cl(gout, ctx.HD, counts=(ch, cw))            # fused grad_out / counts
        halo = ctx.halo
        gk_out = gv_out = None
        if halo is not None:   # owned + halo gradient videos in symmetric memory: the neighbours read their share there
            N, Tk, H, W, Fp = k_cl.shape
            gkbuf, gkhdl = halo.buffer("gk", N, H * W * Fp, k_cl.device)
            gvbuf, gvhdl = halo.buffer("gv", N, H * W * v_cl.shape[-1], k_cl.device)
            gkbuf.zero_()
            gvbuf.zero_()
            gk_out, gv_out = gkbuf.view(k_cl.shape), gvbuf.view(v_cl.shape)
        # every other accumulator out of ONE zero-filled workspace (one fill launch instead of six)
        shapes = [q_cl.shape, weights.shape, inds.shape]
        if flows_grad:
            shapes.append(flows_k.shape)
        if halo is None:
            shapes += [k_cl.shape, v_cl.shape]
        offs, total = [], 0
        for shp in shapes:
            offs.append(total)
            total += (shp.numel() + 3) // 4 * 4          # 16-byte aligned pieces
        ws = th.zeros(total, device=q_cl.device, dtype=th.float32)
        parts = [ws[o:o + shp.numel()].view(shp) for o, shp in zip(offs, shapes)]
        gq_out, gw_out, gf_out = parts[:3]
        gflows_out = parts[3] if flows_grad else None
        if halo is None:
            gk_out, gv_out = parts[-2], parts[-1]
        gv_cl, gw, gf = ops.wpsum_backward(gout_cl, v_cl, weights, inds, ctx.ageom, ctx.aopts, gv_out=gv_out, gw_out=gw_out,
                                           gf_out=gf_out)
        gq_cl, gk_cl, gflows = ops.search_backward_sm(q_cl, k_cl, flows_k, gw, weights, ctx.scale, gf, inds, ctx.geom,
                                                      ctx.opts, flows_grad, gk_out=gk_out, gq_out=gq_out,
                                                      gflows_out=gflows_out)
        if halo is None:
            gq, gk, gv = ops.unpack_cl_multi([gq_cl, gk_cl, gv_cl], ctx.shapes, ctx.HD)
        else:
            halo.handshake(gkhdl)
            halo.pull_add([(gkbuf, gkhdl), (gvbuf, gvhdl)])
            own = halo.plan.a - halo.plan.fa
            gq, gk, gv = ops.unpack_cl_multi([gq_cl, gk_cl, gv_cl], ctx.shapes, ctx.HD, cl_t0=[0, own, own])
        if gflows is not None and ctx.flow_ndim == 6:
            gflows = gflows[:, 0].contiguous()
        return gq, gk, gv, gflows, None, None, None, None, None


def fused_search_agg(search, agg, q, k, v, flows, scale, shard=None, halo=None):
    """scale: signed softmax scale (negative for l2 distances)."""
    return FusedSearchAggFunction.apply(q, k, v, flows, search, agg, scale, shard, halo)
