"""Frame-sharded multi-GPU execution: one process per GPU, frames of a clip partitioned contiguously
across ranks, and ONE exchange step -- the temporal halo of k (search) and v (aggregation).

The reference has no distributed layer at all.  The path shards naturally: every
query only touches frames of its own clip inside its temporal window
(lib/csrc/shared_kernel.cu:300-312), so q, flows, dists, inds, weights and the output are owned and
never exchanged; k / v need frames [t_min(t), t_max(t)] for the owned t (the window is SHIFTED, not
truncated, at clip ends, so the halo is up to 2*wt frames inward there).  Forward: neighbour ranks
send halo frames (NCCL send/recv over NVLink, grouped in one batch).  Backward: the same pattern
reversed, carrying grad_k / grad_v halo slices that are SUMMED into the owner.  No all-reduce, no
all-gather of full tensors.
"""
import os

import torch as th
import torch.distributed as dist

from .normz.softmax import softmax_scaled


def shard_frames(T, world, rank):
    """Contiguous, balanced partition of T frames: returns [a, b)."""
    base, rem = divmod(T, world)
    a = rank * base + min(rank, rem)
    return a, a + base + (1 if rank < rem else 0)


def time_window(t, T, wt):
    """[t_min, t_max] searched from frame t (set_time_range_minmax, lib/csrc/shared_kernel.cu:300-305)."""
    t_shift = min(0, t - wt) + max(0, t + wt - (T - 1))
    return max(t - wt - t_shift, 0), min(T - 1, t + wt - t_shift)


def needed_frames(a, b, T, wt):
    """Frames of k/v a rank owning [a,b) must hold: union of the windows of its frames, as [fa, fb)."""
    if b <= a:
        return a, a
    lo = min(time_window(t, T, wt)[0] for t in range(a, b))
    hi = max(time_window(t, T, wt)[1] for t in range(a, b))
    return lo, hi + 1


class ShardPlan:
    """Who owns what and who needs what, computed identically on every rank (no communication)."""

    def __init__(self, T, wt, world, rank):
        self.T, self.wt, self.world, self.rank = T, wt, world, rank
        self.owned = [shard_frames(T, world, r) for r in range(world)]
        self.need = [needed_frames(a, b, T, wt) for (a, b) in self.owned]
        self.a, self.b = self.owned[rank]
        self.fa, self.fb = self.need[rank]

    @property
    def shard(self):
        """The `shard=` argument of the search / aggregation modules."""
        return dict(T=self.T, tq0=self.a, tk0=self.fa)

    def recvs(self, rank=None):
        """[(src_rank, lo, hi)]: frames [lo,hi) this rank receives from src (its halo)."""
        r = self.rank if rank is None else rank
        fa, fb = self.need[r]
        out = []
        for s, (a, b) in enumerate(self.owned):
            lo, hi = max(fa, a), min(fb, b)
            if s != r and lo < hi:
                out.append((s, lo, hi))
        return out

    def sends(self, rank=None):
        """[(dst_rank, lo, hi)]: owned frames [lo,hi) this rank sends to dst."""
        r = self.rank if rank is None else rank
        a, b = self.owned[r]
        out = []
        for d, (fa, fb) in enumerate(self.need):
            lo, hi = max(fa, a), min(fb, b)
            if d != r and lo < hi:
                out.append((d, lo, hi))
        return out

    def halo_bytes(self, frame_bytes):
        return sum(hi - lo for _, lo, hi in self.recvs()) * frame_bytes


def flat_shard(B, T, wt, world, rank):
    """Shard the flattened (B, T) frame axis contiguously over `world` ranks, never finer than a frame and never across
    a batch element (batch elements are independent: lib/csrc/search/non_local_search_int_kernel.cu:55-56).
    world >= B: world/B ranks share one batch element (C4: B=2, T=32, 8 ranks -> ranks 0-3 own batch 0, 4-7 batch 1),
    the halo exchange runs inside that sub-group.  world < B: every rank takes B/world whole clips, no halo.
    Returns (batch slice, ShardPlan inside the sub-group, ranks of the sub-group)."""
    if world >= B:
        if world % B:
            raise ValueError(f"world size {world} must be a multiple of the batch size {B}")
        rpb = world // B
        b, local = divmod(rank, rpb)
        return slice(b, b + 1), ShardPlan(T, wt, rpb, local), list(range(b * rpb, (b + 1) * rpb))
    if B % world:
        raise ValueError(f"batch size {B} must be a multiple of the world size {world}")
    per = B // world
    return slice(rank * per, (rank + 1) * per), ShardPlan(T, wt, 1, 0), [rank]


def batch_groups(B, world, rank, backend=None):
    """One process group per batch element (see flat_shard); every rank creates all of them, in the same order, as
    torch.distributed requires.  Returns this rank's group (None when it works alone)."""
    if world <= B or world // B == 1:
        return None
    rpb = world // B
    mine = None
    for b in range(B):
        ranks = list(range(b * rpb, (b + 1) * rpb))
        grp = dist.new_group(ranks, backend=backend)
        if rank in ranks:
            mine = grp
    return mine


def _run_p2p(ops_list):
    if ops_list:
        for w in dist.batch_isend_irecv(ops_list):
            w.wait()


def _time_dim(x):
    return 1 if x.dim() == 5 else 2  # (B,T,F,H,W) or (B,HD,T,F,H,W)


def _peer(group, local_rank):
    """global rank of `local_rank` of `group` (send/recv take global ranks)"""
    return local_rank if group is None else dist.get_global_rank(group, local_rank)


def _exchange_forward(x, plan, group):
    td = _time_dim(x)
    shape = list(x.shape)
    shape[td] = plan.fb - plan.fa
    ext = x.new_empty(shape)
    ext.narrow(td, plan.a - plan.fa, plan.b - plan.a).copy_(x)
    p2p, stage = [], []
    for dst, lo, hi in plan.sends():
        buf = x.narrow(td, lo - plan.a, hi - lo).contiguous()
        p2p.append(dist.P2POp(dist.isend, buf, _peer(group, dst), group))
    for src, lo, hi in plan.recvs():
        s = list(x.shape)
        s[td] = hi - lo
        buf = x.new_empty(s)
        stage.append((buf, lo, hi))
        p2p.append(dist.P2POp(dist.irecv, buf, _peer(group, src), group))
    _run_p2p(p2p)
    for buf, lo, hi in stage:
        ext.narrow(td, lo - plan.fa, hi - lo).copy_(buf)
    return ext


def _exchange_backward(gext, plan, group):
    """Reverse pattern: halo slices of the gradient travel back to their owners and are summed there."""
    td = _time_dim(gext)
    gown = gext.narrow(td, plan.a - plan.fa, plan.b - plan.a).contiguous()
    p2p, stage = [], []
    for src, lo, hi in plan.recvs():      # frames we borrowed -> give their gradient back
        buf = gext.narrow(td, lo - plan.fa, hi - lo).contiguous()
        p2p.append(dist.P2POp(dist.isend, buf, _peer(group, src), group))
    for dst, lo, hi in plan.sends():      # frames we lent -> receive their gradient
        s = list(gown.shape)
        s[td] = hi - lo
        buf = gown.new_empty(s)
        stage.append((buf, lo, hi))
        p2p.append(dist.P2POp(dist.irecv, buf, _peer(group, dst), group))
    _run_p2p(p2p)
    for buf, lo, hi in stage:
        tgt = gown.narrow(td, lo - plan.a, hi - lo)
        if tgt.is_cuda and tgt.is_contiguous():
            from . import ops
            ops.add_inplace(tgt, buf)
        else:
            tgt.add_(buf)
    return gown


class HaloExchange(th.autograd.Function):
    """owned frames [a,b) -> resident frames [fa,fb); differentiable (backward = halo gradient reduce)."""

    @staticmethod
    def forward(ctx, x, plan, group=None):
        ctx.plan, ctx.group = plan, group
        return _exchange_forward(x.contiguous(), plan, group)

    @staticmethod
    def backward(ctx, gext):
        return _exchange_backward(gext.contiguous(), ctx.plan, ctx.group), None, None


class PeerHalo:
    """Halo exchange through NVLink **peer memory** instead of NCCL send/recv.

    Every rank stages its frames in a symmetric-memory buffer (torch.distributed._symmetric_memory:
    CUDA VMM allocations mapped into every rank of the node); after ONE device-side barrier on the
    signal pads each rank pulls exactly the halo frames it needs straight out of its neighbours'
    buffers with ordinary copy kernels (P2P loads over NVLink / NVSwitch).  Backward: the extended
    gradient is staged the same way and each owner pulls-and-adds the slices it lent out.  Buffers are
    persistent and double-buffered, so an exchange costs one barrier (a few microseconds) plus the
    copies, against four NCCL P2P launches with their stream hand-offs."""

    def __init__(self, plan, group=None):
        import torch.distributed._symmetric_memory as symm_mem
        self.symm_mem = symm_mem
        self.plan = plan
        self.group = dist.group.WORLD if group is None else group
        self.bufs = {}
        self.count = {}
        self.max_own = max(b - a for a, b in plan.owned)
        self.max_ext = max(fb - fa for fa, fb in plan.need)
        # probe now (allocation + rendezvous + barrier) so an unsupported box fails here, not mid-step
        dev = th.device("cuda", th.cuda.current_device())
        probe = symm_mem.empty((16,), dtype=th.float32, device=dev)
        symm_mem.rendezvous(probe, self.group).barrier(channel=0)
        th.cuda.synchronize()

    def _buffers(self, tag, nframes, frame_shape, dtype, device):
        key = (tag, nframes, tuple(frame_shape), dtype)
        if key not in self.bufs:
            pair = []
            for _ in range(2):
                t = self.symm_mem.empty((nframes,) + tuple(frame_shape), dtype=dtype, device=device)
                pair.append((t, self.symm_mem.rendezvous(t, self.group)))
            self.bufs[key] = pair
            self.count[key] = 0
        i = self.count[key]
        self.count[key] = i + 1
        return self.bufs[key][i % 2]

    def forward(self, x, tag):
        plan, td = self.plan, _time_dim(x)
        xt = x.movedim(td, 0)                        # (T_own, ...) view, time major
        frame_shape = tuple(xt.shape[1:])
        buf, hdl = self._buffers(("f", tag), self.max_own, frame_shape, x.dtype, x.device)
        buf[:plan.b - plan.a].copy_(xt)
        hdl.barrier(channel=0)
        shape = list(x.shape)
        shape[td] = plan.fb - plan.fa
        ext = x.new_empty(shape)
        ext.narrow(td, plan.a - plan.fa, plan.b - plan.a).copy_(x)
        for src, lo, hi in plan.recvs():
            a_src = plan.owned[src][0]
            peer = hdl.get_buffer(src, (self.max_own,) + frame_shape, x.dtype)
            ext.narrow(td, lo - plan.fa, hi - lo).copy_(peer[lo - a_src:hi - a_src].movedim(0, td))
        return ext

    def backward(self, gext, tag):
        plan, td = self.plan, _time_dim(gext)
        gt = gext.movedim(td, 0)
        frame_shape = tuple(gt.shape[1:])
        buf, hdl = self._buffers(("b", tag), self.max_ext, frame_shape, gext.dtype, gext.device)
        buf[:plan.fb - plan.fa].copy_(gt)
        hdl.barrier(channel=0)
        gown = gext.narrow(td, plan.a - plan.fa, plan.b - plan.a).contiguous()
        for dst, lo, hi in plan.sends():              # frames we lent to dst: their gradient sits in dst's buffer
            fa_dst = plan.need[dst][0]
            peer = hdl.get_buffer(dst, (self.max_ext,) + frame_shape, gext.dtype)
            gown.narrow(td, lo - plan.a, hi - lo).add_(peer[lo - fa_dst:hi - fa_dst].movedim(0, td))
        return gown


    # ---- several tensors per exchange, three launches in all ------------------------------------------------------
    # stage (one segmented-copy launch) -> ONE barrier -> pull (one segmented-copy launch: own frames + the halo frames
    # of every tensor from every neighbour, NVLink peer loads).  Backward: stage, barrier, own part, then the lent
    # slices ADDED with vector REDs (two neighbours may return gradient for the same frame).  Needs the frames of a
    # tensor to be contiguous in memory: planar (1,T,F,H,W) videos, i.e. one batch element per rank.
    @staticmethod
    def fast_ok(xs):
        from . import ops
        return all(x.dim() == 5 and x.shape[0] == 1 and ops.seg_copy_ok(x) and x[0, 0].numel() % 4 == 0 for x in xs)

    def forward_multi(self, xs, tag):
        from . import ops
        plan = self.plan
        n_own, n_ext = plan.b - plan.a, plan.fb - plan.fa
        staged = []
        for i, x in enumerate(xs):
            buf, hdl = self._buffers(("f", tag, i), self.max_own, tuple(x.shape[2:]), x.dtype, x.device)
            staged.append((buf, hdl))
        ops.seg_copy([(buf[:n_own], x[0]) for (buf, _), x in zip(staged, xs)])
        staged[0][1].barrier(channel=0)
        exts, jobs = [], []
        for (buf, hdl), x in zip(staged, xs):
            ext = x.new_empty((1, n_ext) + tuple(x.shape[2:]))
            jobs.append((ext[0, plan.a - plan.fa:plan.b - plan.fa], x[0]))
            for src, lo, hi in plan.recvs():
                a_src = plan.owned[src][0]
                peer = hdl.get_buffer(src, (self.max_own,) + tuple(x.shape[2:]), x.dtype)
                jobs.append((ext[0, lo - plan.fa:hi - plan.fa], peer[lo - a_src:hi - a_src]))
            exts.append(ext)
        ops.seg_copy(jobs)
        return exts

    def backward_multi(self, gexts, tag):
        from . import ops
        plan = self.plan
        n_own, n_ext = plan.b - plan.a, plan.fb - plan.fa
        staged = []
        for i, g in enumerate(gexts):
            buf, hdl = self._buffers(("b", tag, i), self.max_ext, tuple(g.shape[2:]), g.dtype, g.device)
            staged.append((buf, hdl))
        ops.seg_copy([(buf[:n_ext], g[0]) for (buf, _), g in zip(staged, gexts)])
        staged[0][1].barrier(channel=0)
        gowns, own_jobs, add_jobs = [], [], []
        for (buf, hdl), g in zip(staged, gexts):
            gown = g.new_empty((1, n_own) + tuple(g.shape[2:]))
            own_jobs.append((gown[0], g[0, plan.a - plan.fa:plan.b - plan.fa]))
            for dst, lo, hi in plan.sends():          # frames we lent to dst: their gradient sits in dst's buffer
                fa_dst = plan.need[dst][0]
                peer = hdl.get_buffer(dst, (self.max_ext,) + tuple(g.shape[2:]), g.dtype)
                add_jobs.append((gown[0, lo - plan.a:hi - plan.a], peer[lo - fa_dst:hi - fa_dst]))
            gowns.append(gown)
        ops.seg_copy(own_jobs)
        ops.seg_copy(add_jobs, accumulate=True)
        return gowns


class PeerHaloCL:
    """Halo exchange in the kernels' own CHANNELS-LAST layout, used by the one-node step (stnls_b200/fused.py).

    A rank's owned + halo key / value videos (N, n_ext, H, W, Fp) live in symmetric memory.  The pack kernel writes the
    owned frames straight into them, a neighbour-only handshake on the allocation's signal pads (put_signal /
    wait_signal towards the ranks this rank exchanges frames with -- no group-wide barrier) makes them visible, and ONE
    segmented-copy launch pulls the halo frames of every tensor from every neighbour over NVLink.  Backward: the
    kernels accumulate into owned + halo gradient videos in symmetric memory, after the handshake ONE launch adds the
    slices lent to the neighbours (vector REDs) into the owned slice, which the unpack kernel reads in place.
    Per step: 2 copy launches and 2 handshakes (the planar exchange of PeerHalo: 5 copy launches, 2 group barriers).
    Buffers are double-buffered: a rank rewrites a buffer two exchanges after its neighbours read it, and has waited
    for their next signal in between.  The forward pass saves VIEWS of these buffers for its backward pass, so the
    backward of a step must run before the second-next forward (forward, backward, forward, backward ... or two steps
    in flight at most); every rank must make the same sequence of calls."""

    def __init__(self, plan, group=None, timeout_ms=30000):
        import torch.distributed._symmetric_memory as symm_mem
        self.symm_mem, self.plan, self.timeout_ms = symm_mem, plan, timeout_ms
        self.group = dist.group.WORLD if group is None else group
        self.nbrs = sorted({r for r, _, _ in plan.recvs()} | {r for r, _, _ in plan.sends()})
        self.max_ext = max(fb - fa for fa, fb in plan.need)
        self.bufs, self.count = {}, {}

    def buffer(self, tag, N, frame, device):
        """(view (N, n_ext, frame) of this rank's buffer, handle); every rank allocates N * max_ext * frame floats."""
        key = (tag, N, frame)
        if key not in self.bufs:
            pair = []
            for _ in range(2):
                t = self.symm_mem.empty((N * self.max_ext * frame,), dtype=th.float32, device=device)
                pair.append((t, self.symm_mem.rendezvous(t, self.group)))
            self.bufs[key] = pair
            self.count[key] = 0
        i = self.count[key]
        self.count[key] = i + 1
        t, hdl = self.bufs[key][i % 2]
        n_ext = self.plan.fb - self.plan.fa
        return t[:N * n_ext * frame].view(N, n_ext, frame), hdl

    def handshake(self, hdl):
        for r in self.nbrs:
            hdl.put_signal(r, 0, self.timeout_ms)
        for r in self.nbrs:
            hdl.wait_signal(r, 0, self.timeout_ms)

    def _peer_ptr(self, hdl, rank, N, frame):
        return hdl.get_buffer(rank, (N * self.max_ext * frame,), th.float32).data_ptr()

    def pull(self, items):
        """items: [(view, hdl)] packed owned frames in place; fetch the halo frames of each from the neighbours."""
        from . import ops
        jobs = []
        for view, hdl in items:
            N, n_ext, frame = view.shape
            for peer, doff, soff, nfloat, rows, dpitch, spitch in halo_pull_jobs(self.plan, N, frame):
                jobs.append((view.data_ptr() + doff * 4, self._peer_ptr(hdl, peer, N, frame) + soff * 4, nfloat, rows,
                             dpitch, spitch))
        ops.seg_copy2d(jobs)

    def pull_add(self, items):
        """items: [(view, hdl)] owned + halo gradients; add what the neighbours accumulated for this rank's frames."""
        from . import ops
        jobs = []
        for view, hdl in items:
            N, n_ext, frame = view.shape
            for peer, doff, soff, nfloat, rows, dpitch, spitch in halo_pull_add_jobs(self.plan, N, frame):
                jobs.append((view.data_ptr() + doff * 4, self._peer_ptr(hdl, peer, N, frame) + soff * 4, nfloat, rows,
                             dpitch, spitch))
        ops.seg_copy2d(jobs, accumulate=True)


def halo_pull_jobs(plan, N, frame):
    """Forward halo of a channels-last video (N, n_ext, frame floats) held per rank as owned + halo frames [fa, fb):
    [(peer, dst_offset, src_offset, nfloat, rows, dst_pitch, src_pitch)] in floats -- `rows` (= N) runs of nfloat floats
    copied from `peer`'s buffer (whose slabs hold ITS n_ext frames) into this rank's.  Pure index math (CPU-tested)."""
    n_ext = plan.fb - plan.fa
    jobs = []
    for src, lo, hi in plan.recvs():
        fa_s, fb_s = plan.need[src]
        jobs.append((src, (lo - plan.fa) * frame, (lo - fa_s) * frame, (hi - lo) * frame, N, n_ext * frame,
                     (fb_s - fa_s) * frame))
    return jobs


def halo_pull_add_jobs(plan, N, frame):
    """Backward: the gradient a neighbour accumulated for frames [lo, hi) this rank OWNS sits in the neighbour's owned + halo
    gradient video; same tuple layout as halo_pull_jobs, to be ADDED into this rank's buffer."""
    n_ext = plan.fb - plan.fa
    jobs = []
    for dst, lo, hi in plan.sends():
        fa_d, fb_d = plan.need[dst]
        jobs.append((dst, (lo - plan.fa) * frame, (lo - fa_d) * frame, (hi - lo) * frame, N, n_ext * frame,
                     (fb_d - fa_d) * frame))
    return jobs


class PeerHaloExchangeMulti(th.autograd.Function):
    """k and v (any number of planar videos) through ONE peer-memory exchange per direction (PeerHalo.forward_multi)."""

    @staticmethod
    def forward(ctx, peer, tag, *xs):
        ctx.peer, ctx.tag = peer, tag
        return tuple(peer.forward_multi([x.contiguous() for x in xs], tag))

    @staticmethod
    def backward(ctx, *gs):
        return (None, None) + tuple(ctx.peer.backward_multi([g.contiguous() for g in gs], ctx.tag))


class PeerHaloExchange(th.autograd.Function):
    """HaloExchange over peer memory (see PeerHalo)."""

    @staticmethod
    def forward(ctx, x, peer, tag):
        ctx.peer, ctx.tag = peer, tag
        return peer.forward(x.contiguous(), tag)

    @staticmethod
    def backward(ctx, gext):
        return ctx.peer.backward(gext.contiguous(), ctx.tag), None, None


def halo_exchange(x, plan, group=None, peer=None, tag="x"):
    if plan.world == 1:
        return x
    if peer is not None:
        return PeerHaloExchange.apply(x, peer, tag)
    return HaloExchange.apply(x, plan, group)


class ShardedSearchAgg(th.nn.Module):
    """search -> softmax -> weighted patch sum on this rank's frames; k and v halos exchanged over NCCL.

    q, k, v: the OWNED frames (B,T_own,F,H,W); flows (B,HD_f,T_own,W_t[-1],2,nH,nW) for the owned frames.
    Returns the owned frames of the aggregated video (B,HD,T_own,F_head,H,W), dists and inds.
    """

    def __init__(self, search, agg, plan, scale=10.0, group=None, halo="auto"):
        """halo: "peer" (NVLink peer memory, see PeerHalo), "nccl" (send/recv) or "auto" (peer when available)."""
        super().__init__()
        self.search, self.agg, self.plan, self.scale, self.group = search, agg, plan, scale, group
        self.peer = None
        self.peer_cl = None
        self.halo_path = None    # set on first use: which exchange the forward pass actually took
        self.halo_kind = "none" if plan.world == 1 else "nccl"
        if plan.world > 1 and halo in ("peer", "auto") and th.cuda.is_available() and dist.get_backend() == "nccl":
            err = None
            try:
                self.peer = PeerHalo(plan, group)
            except Exception as exc:   # symmetric memory not available on this box / for this group
                err = exc
            # every rank of the group must take the same path: a rank whose rendezvous failed drags the others to NCCL
            ok = th.tensor([0.0 if self.peer is None else 1.0], device="cuda")
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
            if float(ok[0]) > 0.5:
                self.halo_kind = "peer"
            else:
                self.peer = None
                if halo == "peer":
                    raise RuntimeError(f"peer-memory halo exchange unavailable on some rank: {err}")

    def forward(self, q, k, v, flows, v_ready=None):
        """v_ready: optional CUDA event after which `v` is valid (e.g. its host->device copy on a side stream);
        the current stream waits for it only when the aggregation needs v, so the copy overlaps the search."""
        if self.peer is not None and v_ready is None and os.environ.get("STNLS_HALO_CL", "1") != "0":
            from . import fused
            if fused.fusable(self.search, self.agg, q, k, v, flows):
                # one node for exchange + step: k / v packed straight into symmetric memory, halo frames pulled in the
                # channels-last layout, neighbour-only handshakes (PeerHaloCL)
                if self.peer_cl is None:
                    self.peer_cl = PeerHaloCL(self.plan, self.group)
                    self.halo_path = "peer memory, channels-last frames pulled inside the one-node step, neighbour handshakes"
                sign = -1.0 if self.search.dist_type == "l2" else 1.0
                return fused.fused_search_agg(self.search, self.agg, q, k, v, flows, sign * self.scale,
                                              shard=self.plan.shard, halo=self.peer_cl)
        joint = self.peer is not None and v_ready is None and PeerHalo.fast_ok([k, v])
        if joint:   # k and v in one exchange: one barrier and two copy launches for both (and for both gradients)
            k_ext, v_ext = PeerHaloExchangeMulti.apply(self.peer, "kv", k, v)
        else:
            k_ext = halo_exchange(k, self.plan, self.group, self.peer, "k")
        assert k_ext.shape[_time_dim(k_ext)] == self.plan.fb - self.plan.fa, "k must hold the owned frames + halo"
        sign = -1.0 if self.search.dist_type == "l2" else 1.0
        if joint or self.plan.world == 1:
            from . import fused
            if v_ready is None and fused.fusable(self.search, self.agg, q, k_ext, v if self.plan.world == 1 else v_ext, flows):
                # one autograd node, 8 launches per forward + backward instead of 14 (stnls_b200/fused.py)
                return fused.fused_search_agg(self.search, self.agg, q, k_ext, v if self.plan.world == 1 else v_ext,
                                              flows, sign * self.scale, shard=self.plan.shard)
        dists, inds = self.search(q, k_ext, flows, shard=self.plan.shard)
        weights = softmax_scaled(dists, sign * self.scale)   # fused sign/scale/softmax (normz)
        if v_ready is not None:
            th.cuda.current_stream().wait_event(v_ready)
        if not joint:
            v_ext = halo_exchange(v, self.plan, self.group, self.peer, "v")
        assert v_ext.shape[_time_dim(v_ext)] == self.plan.fb - self.plan.fa, "v must hold the owned frames + halo"
        out = self.agg(v_ext, weights, inds, shard=self.plan.shard)
        return out, dists, inds

    def chunked_step(self, q, k, v, flows, gout, frames_per_chunk, flows_grad=True):
        """Forward + backward of sum(out * gout) over the owned frames in chunks of `frames_per_chunk` frames, so that
        the intermediates of only one chunk are alive at a time (a 1080p rank holds 16 GB of indices per 8 frames).
        ONE halo exchange per tensor and direction, as in forward(): the exchanged k / v are cut from the graph, every
        chunk runs search -> softmax -> aggregation -> backward on its frames with shard=(T, tq0, tk0) of its own
        window, the halo gradients accumulate in place and travel back once.  Returns (out, grad_q, grad_k, grad_v,
        grad_flows) for the owned frames."""
        plan = self.plan
        q, k, v = (x.detach().requires_grad_(True) for x in (q, k, v))
        flows = flows.detach().requires_grad_(flows_grad and flows.is_floating_point())
        if self.peer is not None and PeerHalo.fast_ok([k, v]):
            k_ext, v_ext = PeerHaloExchangeMulti.apply(self.peer, "kv", k, v)
        else:
            k_ext = halo_exchange(k, plan, self.group, self.peer, "k")
            v_ext = halo_exchange(v, plan, self.group, self.peer, "v")
        kx, vx = k_ext.detach().requires_grad_(True), v_ext.detach().requires_grad_(True)
        sign = -1.0 if self.search.dist_type == "l2" else 1.0
        td = _time_dim(q)
        outs = []
        for a in range(plan.a, plan.b, frames_per_chunk):
            b = min(a + frames_per_chunk, plan.b)
            fa, fb = needed_frames(a, b, plan.T, plan.wt)
            sh = dict(T=plan.T, tq0=a, tk0=fa)
            qc = q.narrow(td, a - plan.a, b - a)
            fc = flows.narrow(2, a - plan.a, b - a)
            dists, inds = self.search(qc, kx.narrow(td, fa - plan.fa, fb - fa), fc, shard=sh)
            weights = softmax_scaled(dists, sign * self.scale)
            out = self.agg(vx.narrow(td, fa - plan.fa, fb - fa), weights, inds, shard=sh)
            (out * gout.narrow(2, a - plan.a, b - a)).sum().backward()
            outs.append(out.detach())
            del dists, inds, weights, out
        if plan.world > 1:
            th.autograd.backward([k_ext, v_ext], [kx.grad, vx.grad])
            gk, gv = k.grad, v.grad
        else:
            gk, gv = kx.grad, vx.grad
        return th.cat(outs, 2), q.grad, gk, gv, flows.grad


class GraphedStep:
    """Forward + backward of a search/aggregation module captured in CUDA graphs over static buffers.

    The eager front end needs about a millisecond of host time to enqueue one README-sized step (14 kernel
    launches plus the autograd bookkeeping around them); a graph replay is one launch.  `slots` independent
    buffer sets, each with its own graph, let a caller fill slot i+1 (host -> device copies on a copy stream)
    while slot i runs.  With the peer-memory halo exchange the slots must be replayed round-robin in capture
    order (each capture takes the next buffer of the exchange's double buffering) and `slots` must be even.

        step = GraphedStep(model, (q, k, v, flows, grad_out), slots=2)
        step.inputs[s][j].copy_(host_tensor, non_blocking=True)   # j: q, k, v, flows, grad_out
        step.replay(s)                                            # on the current stream
        step.outputs[s]                                           # (out, grad_q, grad_k, grad_v[, grad_flows])

    model(q, k, v, flows) -> (out, dists, inds), e.g. ShardedSearchAgg.  The gradients are those of
    sum(out * grad_out).  Shapes are frozen at construction.
    """

    def __init__(self, model, example, slots=2, flows_grad=True, warmup=2):
        assert all(t.is_cuda for t in example), "GraphedStep needs CUDA tensors (no CPU fallback)"
        peer = getattr(model, "peer", None)
        if peer is not None and slots % 2:
            raise ValueError("GraphedStep: the peer halo exchange is double-buffered, slots must be even")
        if getattr(model, "halo_kind", "none") == "nccl":
            raise NotImplementedError("GraphedStep: use the peer-memory halo exchange (halo='peer') at world size > 1")
        self.model, self.slots = model, slots
        self.inputs, self.outputs, self.graphs = [], [], []
        self.dists, self.inds = [], []
        cur, cap = th.cuda.current_stream(), th.cuda.Stream()
        for s in range(slots):
            bufs = [t.detach().clone() for t in example]
            leafs = bufs[:3] + ([bufs[3]] if flows_grad and bufs[3].is_floating_point() else [])
            for t in leafs:
                t.requires_grad_(True)

            def run(bufs=bufs, leafs=leafs):
                out, dists, inds = model(bufs[0], bufs[1], bufs[2], bufs[3])
                grads = th.autograd.grad(out, leafs, grad_outputs=bufs[4])
                return (out.detach(),) + tuple(grads), dists.detach(), inds

            if s == 0:  # warm-up off the capture: allocator, lazily built tables, symmetric-memory rendezvous
                cap.wait_stream(cur)
                with th.cuda.stream(cap):
                    for _ in range(2 * warmup):  # an even count keeps the halo double buffering in step
                        run()
                cur.wait_stream(cap)
                th.cuda.synchronize()
            g = th.cuda.CUDAGraph()
            with th.cuda.graph(g):
                res, dists, inds = run()
            self.inputs.append([t.detach() for t in bufs])  # aliases a caller may copy_ into (the leaves require grad)
            self.outputs.append(res)
            self.dists.append(dists)
            self.inds.append(inds)
            self.graphs.append(g)

    def replay(self, slot):
        self.graphs[slot].replay()
        return self.outputs[slot]
