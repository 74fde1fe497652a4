"""CPU (gloo, world_size 2 and 3): host logic of the frame-sharded multi-GPU path -- shard plan and the
differentiable halo exchange -- without any CUDA call."""
import os
import socket

import pytest
import torch as th
import torch.distributed as dist
import torch.multiprocessing as mp

from stnls_b200.parallel import ShardPlan, halo_exchange, needed_frames, shard_frames, time_window


def test_time_window_matches_reference_kat():
    # dev/trange.py:33-41: frames visited from (ti, wt, T)
    for (ti, wt, T), frames in [((0, 2, 5), [0, 1, 2, 3, 4]), ((1, 2, 5), [1, 2, 3, 4, 0]), ((3, 2, 7), [3, 4, 5, 2, 1]),
                                ((3, 3, 7), [3, 4, 5, 6, 2, 1, 0]), ((2, 1, 3), [2, 1, 0])]:
        lo, hi = time_window(ti, T, wt)
        assert (lo, hi) == (min(frames), max(frames))


def test_shard_plan_covers_everything():
    for T, wt, world in [(40, 2, 8), (64, 3, 8), (10, 2, 4), (7, 3, 2), (5, 2, 1), (9, 1, 4)]:
        plans = [ShardPlan(T, wt, world, r) for r in range(world)]
        owned = [f for p in plans for f in range(p.a, p.b)]
        assert owned == list(range(T))
        for p in plans:
            assert p.fa <= p.a and p.b <= p.fb
            for t in range(p.a, p.b):
                lo, hi = time_window(t, T, wt)
                assert p.fa <= lo and hi < p.fb
            got = sorted(f for _, lo, hi in p.recvs() for f in range(lo, hi)) + list(range(p.a, p.b))
            assert sorted(got) == list(range(p.fa, p.fb))
        # every send has a matching receive
        sends = {(r, d, lo, hi) for r, p in enumerate(plans) for d, lo, hi in p.sends()}
        recvs = {(s, r, lo, hi) for r, p in enumerate(plans) for s, lo, hi in p.recvs()}
        assert sends == recvs
    assert needed_frames(0, 5, 40, 2) == (0, 7)      # clip start: window shifted inward
    assert needed_frames(35, 40, 40, 2) == (33, 40)  # clip end
    assert shard_frames(10, 4, 0) == (0, 3) and shard_frames(10, 4, 3) == (8, 10)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, T, wt, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        plan = ShardPlan(T, wt, world, rank)
        full = th.randn(2, T, 3, 4, 5, generator=th.Generator().manual_seed(0))
        x = full[:, plan.a:plan.b].clone().requires_grad_(True)
        ext = halo_exchange(x, plan)
        ok_fwd = th.equal(ext, full[:, plan.fa:plan.fb])
        # backward: every rank weights its resident frames with its own coefficients
        coefs = [th.randn(2, p.fb - p.fa, 3, 4, 5, generator=th.Generator().manual_seed(100 + r))
                 for r, p in enumerate([ShardPlan(T, wt, world, r) for r in range(world)])]
        (ext * coefs[rank]).sum().backward()
        exp = th.zeros(2, plan.b - plan.a, 3, 4, 5)
        for r in range(world):
            p = ShardPlan(T, wt, world, r)
            lo, hi = max(p.fa, plan.a), min(p.fb, plan.b)
            if lo < hi:
                exp[:, lo - plan.a:hi - plan.a] += coefs[r][:, lo - p.fa:hi - p.fa]
        ok_bwd = th.allclose(x.grad, exp, atol=1e-6)
        ret[rank] = (ok_fwd, ok_bwd)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,T,wt", [(2, 10, 2), (3, 9, 1), (2, 6, 2)])
def test_halo_exchange_gloo(world, T, wt):
    ret = mp.Manager().dict()
    mp.spawn(_worker, args=(world, _free_port(), T, wt, ret), nprocs=world, join=True)
    assert len(ret) == world
    for r in range(world):
        assert ret[r] == (True, True), (r, ret[r])


def test_flat_shard_plan():
    """flattened (B,T) sharding: C4 = B 2, T 32 over 8 ranks -> ranks 0-3 batch 0, ranks 4-7 batch 1"""
    from stnls_b200.parallel import flat_shard
    seen = []
    for r in range(8):
        bs, plan, ranks = flat_shard(2, 32, 3, 8, r)
        assert (bs.start, bs.stop) == (r // 4, r // 4 + 1) and ranks == list(range(4 * (r // 4), 4 * (r // 4) + 4))
        assert plan.world == 4 and plan.rank == r % 4 and plan.b - plan.a == 8
        seen += [(bs.start, t) for t in range(plan.a, plan.b)]
    assert sorted(seen) == [(b, t) for b in range(2) for t in range(32)]
    bs, plan, ranks = flat_shard(4, 16, 1, 2, 1)     # fewer ranks than clips: whole clips, no halo
    assert (bs.start, bs.stop) == (2, 4) and plan.world == 1 and ranks == [1]
    with pytest.raises(ValueError):
        flat_shard(3, 8, 1, 4, 0)


def _worker_groups(rank, world, port, B, T, wt, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from stnls_b200.parallel import batch_groups, flat_shard
        bs, plan, ranks = flat_shard(B, T, wt, world, rank)
        group = batch_groups(B, world, rank)
        full = th.randn(B, T, 3, 4, 5, generator=th.Generator().manual_seed(0))
        x = full[bs, plan.a:plan.b].clone().requires_grad_(True)
        ext = halo_exchange(x, plan, group)
        ok_fwd = th.equal(ext, full[bs, plan.fa:plan.fb])
        ext.sum().backward()
        # every frame is counted once per rank of its batch group that holds it
        exp = th.zeros_like(x)
        for r in ranks:
            p = ShardPlan(T, wt, len(ranks), r - ranks[0])
            lo, hi = max(p.fa, plan.a), min(p.fb, plan.b)
            if lo < hi:
                exp[:, lo - plan.a:hi - plan.a] += 1
        ret[rank] = (ok_fwd, bool(th.equal(x.grad, exp)))
    finally:
        dist.destroy_process_group()


def test_halo_exchange_inside_batch_groups_gloo():
    """world 4, B 2: two sub-groups of two ranks, the halo exchange stays inside each"""
    ret = mp.Manager().dict()
    mp.spawn(_worker_groups, args=(4, _free_port(), 2, 8, 2, ret), nprocs=4, join=True)
    assert len(ret) == 4
    for r in range(4):
        assert ret[r] == (True, True), (r, ret[r])


@pytest.mark.parametrize("T,wt,world", [(10, 2, 2), (20, 2, 4), (9, 1, 3), (40, 2, 8), (7, 3, 2), (11, 2, 5), (6, 2, 3)])
def test_channels_last_halo_job_lists(T, wt, world):
    """Index math of the channels-last halo exchange (parallel.PeerHaloCL) without a GPU: every rank's owned + halo buffer
    (N, n_ext, frame) as a numpy array, the pull / pull-add job lists executed as strided copies.  Forward must
    reproduce the slice [fa, fb) of the clip on every rank; backward must sum, into each owner, every rank's
    gradient for the frames it lent out."""
    import numpy as np
    from stnls_b200 import parallel
    N, frame = 3, 8
    rng = np.random.default_rng(T * 100 + wt * 10 + world)
    clip = rng.standard_normal((N, T, frame)).astype(np.float32)
    plans = [parallel.ShardPlan(T, wt, world, r) for r in range(world)]

    def run(jobs_of, bufs, accumulate):
        out = [b.copy() for b in bufs]
        for r, plan in enumerate(plans):
            for peer, doff, soff, nfloat, rows, dpitch, spitch in jobs_of(plan, N, frame):
                src, dst = bufs[peer].reshape(-1), out[r].reshape(-1)
                for row in range(rows):
                    seg = src[soff + row * spitch: soff + row * spitch + nfloat]
                    if accumulate:
                        dst[doff + row * dpitch: doff + row * dpitch + nfloat] += seg
                    else:
                        dst[doff + row * dpitch: doff + row * dpitch + nfloat] = seg
        return out

    # forward: owned frames packed in place, halo frames unknown (NaN) until pulled
    bufs = []
    for plan in plans:
        b = np.full((N, plan.fb - plan.fa, frame), np.nan, dtype=np.float32)
        b[:, plan.a - plan.fa: plan.b - plan.fa] = clip[:, plan.a:plan.b]
        bufs.append(b)
    got = run(parallel.halo_pull_jobs, bufs, False)
    for plan, b in zip(plans, got):
        assert np.array_equal(b, clip[:, plan.fa:plan.fb]), f"rank {plan.rank}: halo frames differ"
    # backward: every rank holds a gradient for all of [fa, fb); owners collect the lent slices
    grads = [rng.standard_normal((N, p.fb - p.fa, frame)).astype(np.float32) for p in plans]
    want = np.zeros((N, T, frame), dtype=np.float64)
    for p, g in zip(plans, grads):
        want[:, p.fa:p.fb] += g
    got = run(parallel.halo_pull_add_jobs, grads, True)
    for p, b in zip(plans, got):
        own = b[:, p.a - p.fa: p.b - p.fa]
        assert np.allclose(own, want[:, p.a:p.b], atol=1e-5), f"rank {p.rank}: gradient of the owned frames"
