"""GPU, >= 2 devices: the frame-sharded path (NCCL and NVLink peer-memory halo exchange) must reproduce
the single-GPU result -- outputs, indices and all gradients.  Spawns one process per GPU."""
import os
import sys

import pytest
import torch as th

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, halo, port, ret):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    import stnls_b200
    from stnls_b200 import parallel
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    th.cuda.set_device(rank)
    dev = th.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    B, T, F, H, W, HD, ws, wt, ps, K = 1, 9, 8, 24, 20, 2, 3, 2, 3, 6
    g = th.Generator().manual_seed(5)
    q, k, v = [th.randn(B, T, F, H, W, generator=g) for _ in range(3)]
    flows = th.rand(B, 1, T, 2 * wt, 2, H, W, generator=g) / 2 + th.randint(-2, 3, (B, 1, T, 2 * wt, 2, H, W), generator=g) + 0.2
    gout = th.randn(B, HD, T, F // HD, H, W, generator=g)

    def run(plan, halo_kind, sl):
        search = stnls_b200.search.NonLocalSearch(ws, wt, ps, K, nheads=HD, stride1=0.5, dist_type="l2",
                                                  self_action="anchor", itype="float")
        agg = stnls_b200.agg.WeightedPatchSum(ps=ps, stride0=1, itype="float")
        model = parallel.ShardedSearchAgg(search, agg, plan, scale=10.0, halo=halo_kind)
        ins = [x[:, sl].to(dev).requires_grad_(True) for x in (q, k, v)]
        fl = flows[:, :, sl].to(dev).requires_grad_(True)
        out, dists, inds = model(*ins, fl)
        (out * gout[:, :, sl].to(dev)).sum().backward()
        return [out.detach(), dists.detach(), inds.detach()] + [x.grad for x in ins] + [fl.grad], model.halo_kind

    # "peer": planar exchange (PeerHalo) around the module chain / one-node step; "peer_cl": exchange in the channels-last
    # layout inside the one-node step (PeerHaloCL, neighbour-only handshakes) -- the default when peer memory works
    os.environ["STNLS_HALO_CL"] = "1" if halo == "peer_cl" else "0"
    halo = "peer" if halo == "peer_cl" else halo
    plan = parallel.ShardPlan(T, wt, world, rank)
    got, kind = run(plan, halo, slice(plan.a, plan.b))
    assert kind == halo, f"halo path {kind} != requested {halo}"
    full, _ = run(parallel.ShardPlan(T, wt, 1, 0), "nccl", slice(0, T))
    names = ["out", "dists", "inds", "grad_q", "grad_k", "grad_v", "grad_flows"]
    errs = {}
    for n, a, b in zip(names, got, full):
        td = 1 if n in ("grad_q", "grad_k", "grad_v") else 2
        b = b.narrow(td, plan.a, plan.b - plan.a)
        errs[n] = float((a - b).abs().max())
    if halo == "peer":
        # GraphedStep over the sharded module: two slots replayed in turn (one per halo buffer parity), three rounds
        search = stnls_b200.search.NonLocalSearch(ws, wt, ps, K, nheads=HD, stride1=0.5, dist_type="l2",
                                                  self_action="anchor", itype="float")
        agg = stnls_b200.agg.WeightedPatchSum(ps=ps, stride0=1, itype="float")
        model = parallel.ShardedSearchAgg(search, agg, plan, scale=10.0, halo="peer")
        sl = slice(plan.a, plan.b)
        ins = [x[:, sl].to(dev) for x in (q, k, v)] + [flows[:, :, sl].to(dev), gout[:, :, sl].to(dev)]
        step = parallel.GraphedStep(model, ins, slots=2)
        worst = 0.0
        for i in range(6):
            res = step.replay(i % 2)
            for a, b in zip(res, [got[0]] + got[3:]):
                worst = max(worst, float((a - b).abs().max()))
        errs["graphed"] = worst
    ret[rank] = errs
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("halo", ["nccl", "peer", "peer_cl"])
def test_sharded_matches_single(halo):
    if th.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    import torch.multiprocessing as mp
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    port = 29600 + (os.getpid() % 300) + {"nccl": 0, "peer": 1, "peer_cl": 2}[halo]
    mp.spawn(_worker, args=(world, halo, port, ret), nprocs=world, join=True)
    assert len(ret) == world
    for rank, errs in ret.items():
        for n, e in errs.items():
            tol = 0 if n == "inds" else 2e-3
            assert e <= tol, f"rank {rank} {halo}: {n} differs by {e}"


def test_graphed_step_matches_eager():
    """GraphedStep (CUDA-graph replay over static buffer slots) gives the eager module's results, slot after slot,
    also after the slot's inputs are overwritten with new data."""
    sys.path.insert(0, ROOT)
    import stnls_b200
    from stnls_b200 import parallel
    dev = th.device("cuda", 0)
    B, T, F, H, W, HD, ws, wt, ps, K = 1, 5, 8, 24, 20, 2, 3, 2, 3, 6
    search = stnls_b200.search.NonLocalSearch(ws, wt, ps, K, nheads=HD, stride1=0.5, dist_type="l2",
                                              self_action="anchor", itype="float")
    agg = stnls_b200.agg.WeightedPatchSum(ps=ps, stride0=1, itype="float")
    model = parallel.ShardedSearchAgg(search, agg, parallel.ShardPlan(T, wt, 1, 0), scale=10.0)

    def data(seed):
        g = th.Generator().manual_seed(seed)
        q, k, v = [th.randn(B, T, F, H, W, generator=g).to(dev) for _ in range(3)]
        flows = (th.rand(B, 1, T, 2 * wt + 1, 2, H, W, generator=g) * 3 - 1.5).to(dev)
        gout = th.randn(B, HD, T, F // HD, H, W, generator=g).to(dev)
        return [q, k, v, flows, gout]

    def eager(ins):
        leafs = [t.clone().requires_grad_(True) for t in ins[:4]]
        out, dists, inds = model(*leafs)
        (out * ins[4]).sum().backward()
        return [out.detach()] + [t.grad for t in leafs]

    step = parallel.GraphedStep(model, data(1), slots=2)
    for seed, slot in ((1, 0), (2, 1), (3, 0), (4, 1)):
        ins = data(seed)
        for dst, src in zip(step.inputs[slot], ins):
            dst.copy_(src)
        got = step.replay(slot)
        want = eager(ins)
        for n, a, b in zip(["out", "grad_q", "grad_k", "grad_v", "grad_flows"], got, want):
            assert a.shape == b.shape
            assert float((a - b).abs().max()) <= 1e-3 * max(1.0, float(b.abs().max())), (n, seed, slot)
