#!/usr/bin/env python
"""bench.py -- search + weighted-patch-sum forward+backward throughput (BASELINE.json metric).

One *step* = one pass of the hot path over one batch of synthetic video:
    NonLocalSearch fwd -> softmax -> WeightedPatchSum fwd -> loss -> backward of all three
(README.md:25-57 of the reference, config C2 of SURVEY.md 8d).  Metric: Gcandidates/s, a candidate
being one (query, search location) patch comparison = B*HD*T*nH*nW*(2wt+1)*ws*ws per step.

  python bench.py [--gpus N --steps K --warmup W]          product arm (CUDA kernels, C ABI)
  python bench.py --impl reference ...                    reference arm: the CPU restatement of the
                                                          reference kernels (oracle/) on the host cores
N>1 is launched by torchrun, one rank per GPU; frames of ONE clip of T = 5*N frames are sharded over
ranks (weak scaling) and the +-wt halo frames of k / v travel over NCCL (stnls_b200.parallel).

The JSON line carries `value` (inputs resident in HBM), `e2e` (host buffers in, host buffers out),
`roofline` (dominant kernel vs measured HBM copy bandwidth), `roofline_fma` (same kernel vs fp32 FMA
peak at the observed SM clock -- the search forward is FMA-bound), `cpu_baseline`.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch as th  # noqa: E402

WORKLOADS = {
    # README example on CPU and GPU (configs[0], C1): int, search + agg FORWARD only
    "c1": dict(B=1, T=5, F=16, H=64, W=64, HD=2, ws=5, wt=2, ps=3, K=10, stride0=1, stride1=1.0,
               dist_type="l2", self_action="anchor", itype="int", scale=10.0, passes="fwd"),
    # README example on 1 B200 (BASELINE.json configs[1])
    "c2": dict(B=1, T=5, F=16, H=128, W=128, HD=2, ws=5, wt=2, ps=3, K=10, stride0=1, stride1=0.5,
               dist_type="l2", self_action="anchor", itype="float", scale=10.0),
    # RVRT-style (configs[2], C3)
    "c3": dict(B=2, T=16, F=64, H=256, W=256, HD=4, ws=9, wt=1, ps=1, K=9, stride0=1, stride1=1.0,
               dist_type="prod", self_action=None, itype="float", scale=10.0),
    # quarter-resolution stand-in for c4 (tuning runs)
    "c4s": dict(B=1, T=7, F=64, H=540, W=960, HD=4, ws=9, wt=3, ps=3, K=20, stride0=1, stride1=1.0,
                dist_type="prod", self_action=None, itype="float", scale=10.0),
    # 1080p slice per rank (configs[3], C4): T frames PER RANK
    "c4": dict(B=1, T=7, F=64, H=1080, W=1920, HD=4, ws=9, wt=3, ps=3, K=20, stride0=1, stride1=1.0,
               dist_type="prod", self_action=None, itype="float", scale=10.0),
    # the 1080p block every line carries (configs[3], C4): 8 frames PER RANK of full 1080x1920; N ranks hold
    # B*T = 8N frames (N=8: B=2, T=32, ranks 0-3 batch 0, ranks 4-7 batch 1), fwd+bwd in chunks of 2 frames
    "c4_1080p": dict(B=1, T=8, F=64, H=1080, W=1920, HD=4, ws=9, wt=3, ps=3, K=20, stride0=1, stride1=1.0,
                     dist_type="prod", self_action=None, itype="float", scale=10.0, Tmax=32, chunk=2),
    # RefineSearch 4K (configs[4], C5): 4 frames PER RANK (N=8: T=32); BASELINE.json leaves ps / nheads / wt open:
    # ps=3 (README), nheads=1 (ctor default), wt=1 -- SURVEY.md 8d.  Refine forward + aggregation forward+backward
    "c5": dict(B=1, T=4, F=32, H=2160, W=3840, HD=1, ws=3, wt=1, wr=3, Ks=30, K=30, ps=3, stride0=2, stride1=1.0,
               dist_type="l2", self_action=None, itype="float", scale=10.0, Tmax=32, refine=True,
               passes="refine fwd + agg fwd+bwd"),
}


def describe(wl, name, world):
    c = dict(wl)
    c["workload"] = {"c1": "README example, int (BASELINE configs[0]): search+agg forward",
                     "c4_1080p": "1080p (BASELINE configs[3]): 8 frames per rank, search+agg fwd+bwd",
                     "c5": "RefineSearch 4K (BASELINE configs[4]): 4 frames per rank, refine fwd + agg fwd+bwd",
                     "c2": "README example (BASELINE configs[1]): search+agg fwd+bwd",
                     "c3": "RVRT-style GDA replacement (BASELINE configs[2])",
                     "c4": "1080p slice (BASELINE configs[3])",
                     "c4s": "1080p config at quarter resolution (tuning stand-in)"}[name]
    c["name"] = name
    c["T_global"] = wl["T"] * world
    c["sharding"] = "single GPU" if world == 1 else f"frames sharded over {world} ranks, halo exchange of k,v (see `halo`)"
    c["l2_policy"] = "L2 flushed (512 MiB write) before every timed step"
    return c


def candidates_per_step(wl, world=1):
    nH = (wl["H"] - 1) // wl["stride0"] + 1
    nW = (wl["W"] - 1) // wl["stride0"] + 1
    L = wl["Ks"] * wl["wr"] ** 2 if wl.get("refine") else (2 * wl["wt"] + 1) * wl["ws"] ** 2
    return wl["B"] * wl["HD"] * wl["T"] * world * nH * nW * L


def make_inputs(wl, rank, world, pin):
    """Seeded synthetic inputs for the frames this rank owns (SURVEY.md 8d recipe)."""
    g = th.Generator().manual_seed(1234 + rank)
    B, T, F, H, W = wl["B"], wl["T"], wl["F"], wl["H"], wl["W"]
    nH, nW = (H - 1) // wl["stride0"] + 1, (W - 1) // wl["stride0"] + 1
    W_t = 2 * wl["wt"] + 1
    q, k, v = [th.randn(B, T, F, H, W, generator=g) for _ in range(3)]
    fshape = (B, 1, T, W_t - 1, 2, nH, nW)
    flows = th.rand(fshape, generator=g) / 2 + th.randint(-3, 3, fshape, generator=g) + 0.2
    if wl["itype"] == "int":
        flows = flows.round()
    gout = th.randn(B, wl["HD"], T, F // wl["HD"], H, W, generator=g)
    ts = [q, k, v, flows, gout]
    if pin:
        ts = [t.pin_memory() for t in ts]
    return ts


# ------------------------------------------------------------------------------------------------
#   algorithmic bytes / flops per launch of each kernel (DESIGN.md "Kernels and rooflines")
# ------------------------------------------------------------------------------------------------

def algorithmic(wl, Tk):
    """bytes each kernel must move at minimum (every tensor touched once) and fp32 flops of the
    search forward, for ONE rank owning wl['T'] query frames and holding Tk key/value frames."""
    B, HD, T, F, H, W = wl["B"], wl["HD"], wl["T"], wl["F"], wl["H"], wl["W"]
    Fh = F // HD
    nH, nW = (H - 1) // wl["stride0"] + 1, (W - 1) // wl["stride0"] + 1
    W_t, K, ps = 2 * wl["wt"] + 1, wl["K"], wl["ps"]
    N = B * HD
    vq, vk = N * T * Fh * H * W * 4, N * Tk * Fh * H * W * 4
    dists = N * T * nH * nW * K * 4
    inds = dists * 3
    if wl.get("refine"):
        flows = N * T * nH * nW * wl["Ks"] * 3 * 4
        cands = N * T * nH * nW * wl["Ks"] * wl["wr"] ** 2
        win = wl["wr"]
    else:
        flows = B * T * (W_t - 1) * 2 * nH * nW * 4
        cands = N * T * nH * nW * W_t * wl["ws"] ** 2
        win = wl["ws"]
    per_pix = (3 if wl["dist_type"] == "l2" else 2) + (0 if wl["itype"] == "int" else 8)
    flt = wl["itype"] != "int"
    # flops the kernels actually execute per candidate: every lattice point of the window is interpolated ONCE (4 flops,
    # separable) and then feeds up to ps*ps candidates (3 flops l2 / 2 prod) -- the SURVEY figure charges 8 flops of
    # bilinear work per patch pixel of every candidate, which neither kernel performs
    half = (not wl.get("refine")) and wl["stride1"] == 0.5 and flt
    if half:
        subs = [((win + 1) // 2, (win + 1) // 2), ((win + 1) // 2, win // 2), (win // 2, (win + 1) // 2), (win // 2, win // 2)]
    else:
        subs = [(win, win)]
    lattice = sum((a + ps - 1) * (b + ps - 1) for a, b in subs)
    exe = (4 * lattice if flt else 0) + (3 if wl["dist_type"] == "l2" else 2) * ps * ps * win * win
    return {
        "search_forward": dict(bytes=vq + vk + flows + dists + inds, flops=cands * ps * ps * Fh * per_pix,
                               flops_executed=cands // (win * win) * exe * Fh),
        # reads q,k,grad_dists,inds(,grad_inds,flows); writes grad_q, grad_k(, grad_flows)
        "search_backward": dict(bytes=2 * vq + 2 * vk + dists + inds + (inds + 2 * flows if flt else 0)),
        "wpsum_forward": dict(bytes=vk + dists + inds + vq),          # v, weights, inds -> out
        # reads grad_out, v, weights, inds; writes grad_v, grad_weights(, grad_flows)
        "wpsum_backward": dict(bytes=vq + 2 * vk + 2 * dists + inds + (inds if flt else 0)),
        "pack_cl": dict(bytes=2 * vk), "unpack_cl": dict(bytes=2 * vk),
        "add_inplace": dict(bytes=0),
    }


# ------------------------------------------------------------------------------------------------
#   clocks / throttle reasons sampled DURING the timed region (pynvml; nvidia-smi fallback)
# ------------------------------------------------------------------------------------------------

class ClockSampler:
    BAD = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown"}
    NOTE = {0x4: "sw_power_cap"}

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self.period = float(os.environ.get("BENCH_CLOCK_PERIOD", "0.02"))
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _loop(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in list(self.BAD.items()) + list(self.NOTE.items()):
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(self.period)

    def __enter__(self):
        if self.nv is not None:
            self._thr = threading.Thread(target=self._loop, daemon=True)
            self._thr.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._thr is not None:
            self._thr.join()

    def summary(self):
        s = sorted(self.samples)
        med = s[len(s) // 2] if s else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


# ------------------------------------------------------------------------------------------------
#   product arm
# ------------------------------------------------------------------------------------------------

def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def load_traffic(workload, kernel):
    """DRAM bytes per launch of `kernel` from the committed `ncu --set full` capture (profiles/traffic.json)."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        with open(p) as f:
            return json.load(f).get(workload, {}).get(kernel)
    except OSError:
        return None


def bind_near_gpu(local):
    """N>1: run this rank's host thread on the CPUs next to its GPU (NVML's ideal affinity), so that the pinned
    host buffers of the e2e loop land on the GPU's NUMA node and 8 ranks do not push all copies through one socket.
    Returns the number of CPUs bound to, or None when NVML is not usable (nothing changes then)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        uuid = str(th.cuda.get_device_properties(local).uuid)
        h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid if not uuid.startswith("GPU-") else uuid).encode())
        pynvml.nvmlDeviceSetCpuAffinity(h)
        return len(os.sched_getaffinity(0))
    except Exception:
        return None


def run_product(args):
    import torch.distributed as dist
    import stnls_b200
    from stnls_b200 import ops, parallel

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not th.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product arm has no CPU path)")
    th.cuda.set_device(local)
    dev = th.device("cuda", local)
    numa = None
    if world > 1:
        numa = bind_near_gpu(local)   # before the pinned buffers are allocated (first touch decides their NUMA node)
        dist.init_process_group("nccl", device_id=dev)
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world} (launch N>1 with torchrun)"

    wl = WORKLOADS[args.workload]
    plan = parallel.ShardPlan(wl["T"] * world, wl["wt"], world, rank)
    assert plan.b - plan.a == wl["T"]
    search = stnls_b200.search.NonLocalSearch(wl["ws"], wl["wt"], wl["ps"], wl["K"], nheads=wl["HD"],
                                              stride0=wl["stride0"], stride1=wl["stride1"],
                                              dist_type=wl["dist_type"], self_action=wl["self_action"],
                                              itype=wl["itype"])
    agg = stnls_b200.agg.WeightedPatchSum(ps=wl["ps"], stride0=wl["stride0"], itype=wl["itype"])
    model = parallel.ShardedSearchAgg(search, agg, plan, scale=wl["scale"], halo=args.halo)

    host = make_inputs(wl, rank, world, pin=True)
    flt = wl["itype"] == "float"

    def step(q, k, v, flows, gout, v_ready=None, after_forward=None):
        q.requires_grad_(True); k.requires_grad_(True); v.requires_grad_(True)
        if flt:
            flows.requires_grad_(True)
        out, dists, inds = model(q, k, v, flows, v_ready=v_ready)
        if after_forward is not None:
            after_forward(out)
        loss = (out * gout).sum()
        loss.backward()
        return out, loss

    flush = th.empty(512 << 20, dtype=th.uint8, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        th.cuda.synchronize()

    def agree(flag):
        """True when `flag` holds on every rank (ranks must take the same path through the collectives below)"""
        if world == 1:
            return bool(flag)
        t = th.tensor([1.0 if flag else 0.0], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        return float(t[0]) > 0.5

    # ---- value: inputs resident in HBM -------------------------------------------------------
    resident = [t.to(dev) for t in host]
    # the timed steps' inputs are cloned BEFORE the warm-up so that the warm-up steps grow the caching
    # allocator to its steady-state size (otherwise the first timed step pays for cudaMalloc)
    fresh = [[t.detach().clone() for t in resident] for _ in range(args.steps)]
    # NCCL sets its send/recv channels up lazily, over several calls: settle them before the warm-up proper
    for _ in range(args.warmup + (8 if model.halo_kind == "nccl" else 0)):
        step(*[t.detach().clone() for t in resident])
    barrier()
    evs = []
    host_enqueue = []
    ops.TIMELINE = []
    launches0 = ops.LAUNCHES
    barrier()
    with ClockSampler(local) as clk:
        for i in range(args.steps):
            flush.zero_()
            e0, e1 = th.cuda.Event(enable_timing=True), th.cuda.Event(enable_timing=True)
            e0.record()
            h0 = time.perf_counter()
            step(*fresh[i])
            host_enqueue.append(time.perf_counter() - h0)
            e1.record()
            evs.append((e0, e1))
            for t in fresh[i]:   # the step's gradients were computed; do not keep K sets of them alive -- the caching
                t.grad = None    # allocator would have to cudaMalloc inside the timed loop once its pool is used up
        barrier()
    launches = ops.LAUNCHES - launches0
    timeline, ops.TIMELINE = ops.TIMELINE, None
    t_ms = sum(a.elapsed_time(b) for a, b in evs)
    if os.environ.get("BENCH_DEBUG"):
        print(f"[rank {rank}] host enqueue ms per step:", [round(x * 1e3, 3) for x in host_enqueue], file=sys.stderr, flush=True)
        print(f"[rank {rank}] per-step ms:", [round(a.elapsed_time(b), 3) for a, b in evs], file=sys.stderr, flush=True)
    del fresh

    # ---- value, graph form: the same step captured once in a CUDA graph and replayed ---------------
    # The eager pass above needs ~1.4 ms of host time to enqueue a 2.5 ms step; on a busy host the GPU then
    # waits for Python.  A graph replay is one launch per step, so `value` is what the kernels deliver.  The
    # eager pass stays: its per-launch CUDA events feed the kernel table / roofline.
    graph_ms, graph_note, per_graph = None, None, 0
    # N>1: the peer-memory halo exchange alternates between two buffers per tag, so the step is captured twice (one
    # graph per buffer parity) and the two graphs are replayed in turn, as GraphedStep does with its slots.
    unit = 1 if world == 1 else 2
    if args.graph != "off" and world > 1 and model.halo_kind != "peer":
        graph_note = "graph replay at N>1 is only used with the peer-memory halo exchange; eager timing reported"
    elif args.graph != "off":
        cgs = None
        try:
            static = [t.detach().clone() for t in resident]
            leafs = static[:3] + ([static[3]] if flt else [])
            for t in leafs:
                t.requires_grad_(True)

            def fwd_bwd():
                out, dists, inds = model(static[0], static[1], static[2], static[3])
                loss = (out * static[4]).sum()
                return out, th.autograd.grad(loss, leafs)

            cur, cap = th.cuda.current_stream(), th.cuda.Stream()
            cap.wait_stream(cur)
            with th.cuda.stream(cap):
                for _ in range(2 * unit):
                    fwd_bwd()
            cur.wait_stream(cap)
            th.cuda.synchronize()
            l0 = ops.LAUNCHES
            cgs, graph_out = [], None
            for _ in range(unit):   # N>1: one graph per halo buffer parity, replayed in capture order
                gph = th.cuda.CUDAGraph()
                with th.cuda.graph(gph):
                    graph_out = fwd_bwd()
                cgs.append(gph)
            per_graph = (ops.LAUNCHES - l0) // unit
        except Exception as exc:  # capture not possible here: the eager number stands
            graph_note = f"graph capture failed: {type(exc).__name__}: {exc}"[:200]
            cgs = None
            th.cuda.synchronize()
        if not agree(cgs is not None):   # every rank replays, or none does
            cgs = None
        if cgs is not None:
            for i in range(unit * -(-args.warmup // unit)):
                cgs[i % unit].replay()
            barrier()
            gevs = []
            with ClockSampler(local) as gclk:
                for i in range(args.steps):
                    flush.zero_()
                    e0, e1 = th.cuda.Event(enable_timing=True), th.cuda.Event(enable_timing=True)
                    e0.record()
                    cgs[i % unit].replay()
                    e1.record()
                    gevs.append((e0, e1))
                barrier()
            graph_ms = sum(a.elapsed_time(b) for a, b in gevs)
            assert bool(th.isfinite(graph_out[0]).all())
            if os.environ.get("BENCH_DEBUG"):
                print(f"[rank {rank}] graph per-step ms:", [round(a.elapsed_time(b), 3) for a, b in gevs], file=sys.stderr, flush=True)
            clk = gclk
            del cgs, graph_out, static

    # ---- e2e: pinned host buffers in, results back to pinned host buffers ----------------------
    # Every step uploads its own q,k,v,flows,grad_out from pinned host memory and downloads out and the four
    # gradients into pinned host memory.  Two timings of the same work:
    #   * pipelined (the e2e value): step i+1's uploads are enqueued on a copy stream before step i's kernels, as a
    #     data-parallel training loop prefetches its next batch; one synchronize after the K steps;
    #   * latency: barrier + synchronize around every single step (nothing overlaps across steps).
    h2d = sum(t.numel() * t.element_size() for t in host)
    main, up, down = th.cuda.current_stream(), th.cuda.Stream(), th.cuda.Stream()
    pinned = {}

    def to_host(name, t, after):
        """download `t` on the copy-out stream once `after` (an event on the compute stream) has passed"""
        if name not in pinned:
            pinned[name] = th.empty(t.shape, dtype=t.dtype).pin_memory()
        down.wait_event(after)
        with th.cuda.stream(down):
            pinned[name].copy_(t, non_blocking=True)
        t.record_stream(down)

    def mark():
        ev = th.cuda.Event()
        ev.record(main)
        return ev

    def upload():
        """q,k,flows first (the search waits for them), then v and grad_out (needed from the aggregation on)"""
        with th.cuda.stream(up):
            q_d, k_d, f_d = (host[j].to(dev, non_blocking=True) for j in (0, 1, 3))
            first = th.cuda.Event()
            first.record(up)
            v_d, g_d = host[2].to(dev, non_blocking=True), host[4].to(dev, non_blocking=True)
            second = th.cuda.Event()
            second.record(up)
        for t in (q_d, k_d, f_d, v_d, g_d):
            t.record_stream(main)
        return [q_d, k_d, v_d, f_d, g_d], first, second

    def e2e_step(dev_in, first, second):
        main.wait_event(first)
        dev_in[2].requires_grad_(True)
        # grad_v is complete as soon as the aggregation's backward has run: its download overlaps the search backward
        hook = dev_in[2].register_post_accumulate_grad_hook(lambda p: to_host("grad_v", p.grad, mark()))
        # the aggregated video goes back to the host while the backward runs
        out, loss = step(*dev_in, v_ready=second, after_forward=lambda o: to_host("out", o.detach(), mark()))
        hook.remove()
        done = mark()
        to_host("grad_q", dev_in[0].grad, done)
        to_host("grad_k", dev_in[1].grad, done)
        if flt:
            to_host("grad_flows", dev_in[3].grad, done)

    # measured host link: one upload of a step's inputs and one download of as many bytes, every rank at the same time
    barrier()
    t0 = time.perf_counter()
    tmp = [t.to(dev, non_blocking=True) for t in host]
    th.cuda.synchronize()
    h2d_gbs = h2d / (time.perf_counter() - t0) / 1e9
    back = [th.empty(t.shape, dtype=t.dtype).pin_memory() for t in host]
    barrier()
    t0 = time.perf_counter()
    for a_, b_ in zip(back, tmp):
        a_.copy_(b_, non_blocking=True)
    th.cuda.synchronize()
    d2h_gbs = h2d / (time.perf_counter() - t0) / 1e9
    del tmp, back

    # latency form
    lat_s = 0.0
    for i in range(args.warmup + args.steps):
        flush.zero_()
        barrier()
        t0 = time.perf_counter()
        up.wait_stream(main)
        e2e_step(*upload())
        th.cuda.synchronize()
        if i >= args.warmup:
            lat_s += time.perf_counter() - t0
    d2h = sum(t.numel() * t.element_size() for t in pinned.values())

    # pipelined form, eager front end
    def pipelined(n):
        nxt = upload()
        for i in range(n):
            cur = nxt
            if i + 1 < n:
                nxt = upload()
            e2e_step(*cur)
        th.cuda.synchronize()

    pipelined(args.warmup)
    barrier()
    t0 = time.perf_counter()
    pipelined(args.steps)
    e2e_eager_s = time.perf_counter() - t0
    assert all(bool(th.isfinite(t).all()) for t in pinned.values())
    e2e_s, e2e_from = e2e_eager_s, "eager"

    # pipelined form through stnls_b200.parallel.GraphedStep: two buffer slots, each with its own CUDA graph of the
    # step; slot s is filled from pinned host memory on the copy-in stream while the other slot runs, results leave on
    # the copy-out stream.  Per step the host enqueues the copies and one graph launch.
    if graph_ms is not None:
        gs = None
        try:
            from stnls_b200.parallel import GraphedStep
            gs = GraphedStep(model, resident, slots=2, flows_grad=flt)
        except Exception as exc:
            graph_note = f"e2e graph path failed: {type(exc).__name__}: {exc}"[:200]
            gs = None
            th.cuda.synchronize()
        if not agree(gs is not None):
            gs = None
        if gs is not None:
            names = ["out", "grad_q", "grad_k", "grad_v"] + (["grad_flows"] if flt else [])
            gpin = {n: th.empty(t.shape, dtype=t.dtype).pin_memory() for n, t in zip(names, gs.outputs[0])}
            computed = [th.cuda.Event() for _ in range(2)]
            drained = [th.cuda.Event() for _ in range(2)]
            filled = [th.cuda.Event() for _ in range(2)]

            def graphed(n):
                for ev in computed + drained:
                    ev.record(main)
                for i in range(n):
                    s = i % 2
                    up.wait_event(computed[s])       # slot s' previous step has read its inputs
                    with th.cuda.stream(up):
                        for j in (0, 1, 3, 2, 4):
                            gs.inputs[s][j].copy_(host[j], non_blocking=True)
                        filled[s].record(up)
                    main.wait_event(filled[s])
                    main.wait_event(drained[s])      # slot s' previous results have left the device
                    gs.replay(s)
                    computed[s].record(main)
                    down.wait_event(computed[s])
                    with th.cuda.stream(down):
                        for nme, t in zip(names, gs.outputs[s]):
                            gpin[nme].copy_(t, non_blocking=True)
                        drained[s].record(down)
                th.cuda.synchronize()

            graphed(2 * -(-args.warmup // 2))
            barrier()
            t0 = time.perf_counter()
            graphed(args.steps)
            e2e_graph_s = time.perf_counter() - t0
            # same inputs every step: the graphed results must be the eager ones (atomics: tolerance)
            same = all(bool(th.isfinite(gpin[n]).all()) and
                       float((gpin[n] - pinned[n]).abs().max()) <= 1e-3 * max(1.0, float(pinned[n].abs().max()))
                       for n in names)
            if not same:
                graph_note = "e2e graph path: results differ from the eager path; eager e2e reported"
            # long steps gain nothing from the graph and lose the overlap inside the step: keep the faster driver
            if agree(same) and agree(e2e_graph_s < e2e_eager_s):
                e2e_s, e2e_from = e2e_graph_s, "graph"
            del gs

    tt = th.tensor([t_ms, e2e_s * 1e3, graph_ms if graph_ms is not None else -1.0, lat_s * 1e3, e2e_eager_s * 1e3],
                   dtype=th.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ok = th.tensor([1.0 if graph_ms is not None else 0.0], device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if float(ok[0]) == 0.0:
            graph_ms = None
    t_ms, e2e_ms, lat_ms, e2e_eager_ms = float(tt[0]), float(tt[1]), float(tt[3]), float(tt[4])
    if graph_ms is not None:
        graph_ms = float(tt[2])
    return dict(world=world, rank=rank, wl=wl, plan=plan, halo=model.halo_kind, halo_path=getattr(model, "halo_path", None), t_ms=t_ms, e2e_ms=e2e_ms, timeline=timeline,
                launches=launches, clocks=clk.summary(), h2d=h2d, d2h=d2h, graph_ms=graph_ms, graph_note=graph_note,
                per_graph=per_graph, unit=unit, lat_ms=lat_ms, e2e_from=e2e_from, e2e_eager_ms=e2e_eager_ms, numa=numa,
                h2d_gbs=h2d_gbs, d2h_gbs=d2h_gbs)



# ------------------------------------------------------------------------------------------------
#   blocks: further BASELINE configurations timed beside the main workload (device-resident inputs, CUDA events,
#   L2 flushed between steps, max over ranks).  Inputs are generated ON the device (a 1080p rank holds 13 GB of video).
# ------------------------------------------------------------------------------------------------

def block_shape(wl, world):
    """(B, T) of the whole job for a per-rank workload: wl['T'] frames per rank, clips of at most wl['Tmax'] frames"""
    total = wl["T"] * world
    T = min(wl.get("Tmax", total), total)
    return total // T, T


def device_inputs(wl, B, plan, dev, rank):
    """owned frames of q,k,v,grad_out (+ flows) on the device; NLS flows as SURVEY.md 8d, refine offsets as 8d C5"""
    g = th.Generator(device=dev).manual_seed(4321 + rank)
    Tr, F, H, W, HD = plan.b - plan.a, wl["F"], wl["H"], wl["W"], wl["HD"]
    nH, nW = (H - 1) // wl["stride0"] + 1, (W - 1) // wl["stride0"] + 1
    q, k, v = [th.randn(B, Tr, F, H, W, device=dev, generator=g) for _ in range(3)]
    gout = th.randn(B, HD, Tr, F // HD, H, W, device=dev, generator=g)
    if wl.get("refine"):
        fl = th.empty(B, HD, Tr, nH, nW, wl["Ks"], 3, device=dev)
        dt = th.randint(-1, 2, fl[..., 0].shape, device=dev, generator=g)
        tq = th.arange(plan.a, plan.b, device=dev).view(1, 1, Tr, 1, 1, 1)
        fl[..., 0] = ((tq + dt).clamp(0, plan.T - 1) - tq).float()
        fl[..., 1:] = th.randint(-8, 8, fl[..., 1:].shape, device=dev, generator=g).float()
        fl[..., 1:] += th.rand(fl[..., 1:].shape, device=dev, generator=g) / 2 + 0.2
    else:
        shape = (B, 1, Tr, 2 * wl["wt"], 2, nH, nW)
        fl = th.rand(shape, device=dev, generator=g) / 2 + th.randint(-3, 3, shape, device=dev, generator=g) + 0.2
        if wl["itype"] == "int":
            fl = fl.round()
    return q, k, v, fl, gout


def run_block(name, args, world, rank, dev, steps, warmup, groups=None):
    """Time `steps` passes of workload `name` (see WORKLOADS) on this job's ranks.  Returns a dict on every rank."""
    import torch.distributed as dist
    import stnls_b200
    from stnls_b200 import ops, parallel
    wl = WORKLOADS[name]
    B, T = block_shape(wl, world)
    bslice, plan, ranks = parallel.flat_shard(B, T, wl["wt"], world, rank)
    group = None
    if len(ranks) > 1 and len(ranks) < world:
        group = groups[B] if groups and B in groups else parallel.batch_groups(B, world, rank)
        if groups is not None:
            groups[B] = group
    Bl = bslice.stop - bslice.start
    kw = dict(nheads=wl["HD"], stride0=wl["stride0"], stride1=wl["stride1"], dist_type=wl["dist_type"],
              self_action=wl["self_action"], itype=wl["itype"])
    if wl.get("refine"):
        search = stnls_b200.search.RefineSearch(wl["ws"], wl["wt"], wl["wr"], wl["K"], -1, wl["ps"], **kw)
    else:
        search = stnls_b200.search.NonLocalSearch(wl["ws"], wl["wt"], wl["ps"], wl["K"], **kw)
    agg = stnls_b200.agg.WeightedPatchSum(ps=wl["ps"], stride0=wl["stride0"], itype=wl["itype"])
    model = parallel.ShardedSearchAgg(search, agg, plan, scale=wl["scale"], group=group, halo=args.halo)
    q, k, v, fl, gout = device_inputs(wl, Bl, plan, dev, rank)
    passes = wl.get("passes", "fwd+bwd")
    sign = -1.0 if wl["dist_type"] == "l2" else 1.0

    def step():
        if passes == "fwd":
            with th.no_grad():
                return model(q, k, v, fl)[0]
        if wl.get("refine"):   # refine forward (no gradient) + aggregation forward + backward
            with th.no_grad():
                k_ext = parallel.halo_exchange(k, plan, group, model.peer, "k")
                d, i = search(q, k_ext, fl, shard=plan.shard)
                w = stnls_b200.normz.softmax.softmax_scaled(d, sign * wl["scale"])
            vg, wg = v.detach().requires_grad_(True), w.requires_grad_(True)
            v_ext = parallel.halo_exchange(vg, plan, group, model.peer, "v")
            out = agg(v_ext, wg, i, shard=plan.shard)
            (out * gout).sum().backward()
            return out.detach()
        return model.chunked_step(q, k, v, fl, gout, wl.get("chunk", plan.b - plan.a))[0]

    def barrier():
        if world > 1:
            dist.barrier()
        th.cuda.synchronize()

    flush = th.empty(512 << 20, dtype=th.uint8, device=dev)
    for _ in range(warmup):
        out = step()
    barrier()
    assert bool(th.isfinite(out).all()), name + ": non-finite output"
    del out
    ops.TIMELINE = []
    evs = []
    barrier()
    for _ in range(steps):
        flush.zero_()
        e0, e1 = th.cuda.Event(enable_timing=True), th.cuda.Event(enable_timing=True)
        e0.record()
        step()
        e1.record()
        evs.append((e0, e1))
    barrier()
    timeline, ops.TIMELINE = ops.TIMELINE, None
    ms = sum(a.elapsed_time(b) for a, b in evs) / steps
    tt = th.tensor([ms], dtype=th.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    ms = float(tt[0])
    peak_mem = th.cuda.max_memory_allocated(dev) / 2 ** 30
    del q, k, v, fl, gout, flush
    th.cuda.empty_cache()
    cands = candidates_per_step(dict(wl, B=1), world)   # wl['T'] frames per rank, `world` ranks
    kt = kernel_table(timeline, steps)
    # algorithmic bytes per LAUNCH: a launch handles one chunk of frames (its own frames + the temporal halo of k / v)
    ct = min(wl.get("chunk", plan.b - plan.a), plan.b - plan.a)
    alg = algorithmic(dict(wl, B=Bl, T=ct), min(ct + 2 * wl["wt"], plan.fb - plan.fa))
    for n_ in kt:
        if n_ in alg and alg[n_]["bytes"]:
            kt[n_]["alg_GBs"] = alg[n_]["bytes"] / (kt[n_]["ms_per_launch"] * 1e-3) / 1e9
    cfg = {k_: wl[k_] for k_ in ("F", "H", "W", "HD", "ws", "wt", "ps", "K", "stride0", "stride1", "dist_type", "itype")}
    if wl.get("refine"):
        cfg.update(wr=wl["wr"], Ks=wl["Ks"], nheads=wl["HD"])
    return {"workload": describe(wl, name, world)["workload"], "config": dict(cfg, B=B, T=T, frames_per_rank=wl["T"]),
            "passes": passes, "n_gpus": world, "steps": steps, "warmup": warmup, "ms_per_step": ms,
            "frames_per_s": B * T / (ms * 1e-3), "Gcandidates_per_s": cands / (ms * 1e-3) / 1e9,
            "halo": model.halo_kind, "peak_device_GiB": round(peak_mem, 1),
            "kernels": {n_: {"ms_per_step": round(v_["ms_per_step"], 4), "launches_per_step": v_["launches"] / steps,
                             **({"alg_GBs": round(v_["alg_GBs"], 1)} if "alg_GBs" in v_ else {})} for n_, v_ in kt.items()},
            "data": "synthetic, generated on the device", "l2_policy": "L2 flushed (512 MiB write) before every timed step"}


def reference_cuda_on_b200(dev, budget_frames=4):
    """the reference's OWN kernels (oracle/_ref, compiled from /root/reference/lib/csrc) through the
    reference's stock call path (oracle/ref_cuda.py: non_local_search_*_forward -> anchor_self -> topk ->
    gather_add_*; backward kernels included), timed with CUDA events on this GPU, outside every product timing.
    C1 and C2 at full size; C3 on a 4-frame slice of one clip (its intermediates are 16 B per candidate)."""
    try:
        from oracle import ref_cuda
        if not ref_cuda.available():
            return {"unavailable": "oracle/_ref/stnls_cuda_ref*.so not present"}
        ref_cuda.mod()
    except Exception as exc:
        return {"unavailable": f"{type(exc).__name__}: {exc}"[:160]}
    res = {"what": "reference CUDA kernels (lib/csrc, unmodified but for .type()->.scalar_type()) driven as "
                   "lib/stnls/search/impl/non_local_search.py + lib/stnls/agg/gather_add.py do; CUDA events, median of 3"}
    for name in ("c1", "c2", "c3"):
        wl = dict(WORKLOADS[name])
        if name == "c3":
            wl.update(B=1, T=budget_frames)
        q, k, v, fl, gout = [t.to(dev) for t in make_inputs(wl, 0, 1, pin=False)]
        kw = dict(nheads=wl["HD"], stride0=wl["stride0"], dist_type=wl["dist_type"], itype=wl["itype"])
        flt = wl["itype"] == "float"
        sign = -1.0 if wl["dist_type"] == "l2" else 1.0
        bwd = wl.get("passes", "fwd+bwd") != "fwd"

        def once():
            d, i = ref_cuda.nls_forward(q, k, fl, wl["ws"], wl["wt"], wl["ps"], wl["K"], stride1=wl["stride1"],
                                        self_action=wl["self_action"], **kw)
            d.requires_grad_(bwd)
            w = th.softmax(sign * wl["scale"] * d, -1)
            out, counts = ref_cuda.gather_add_forward(v, w.detach(), i, wl["ps"], wl["stride0"], itype=wl["itype"])
            if bwd:
                gv, gw, gf = ref_cuda.gather_add_backward(gout, v, w.detach(), i, counts, wl["ps"], wl["stride0"],
                                                          itype=wl["itype"])
                w.backward(gw)
                ref_cuda.search_backward(q, k, fl, d.grad, gf if flt else None, i, wl["wt"], wl["ps"], flows_grad=flt,
                                         dists=d.detach(), **kw)
            return out
        try:
            once()
            th.cuda.synchronize()
            ts = []
            for _ in range(3):
                e0, e1 = th.cuda.Event(enable_timing=True), th.cuda.Event(enable_timing=True)
                e0.record()
                once()
                e1.record()
                e1.synchronize()
                ts.append(e0.elapsed_time(e1))
            ms = sorted(ts)[1]
            res[name] = {"ms_per_step": ms, "Gcandidates_per_s": candidates_per_step(wl) / (ms * 1e-3) / 1e9,
                         "passes": wl.get("passes", "fwd+bwd"),
                         "sample": "full size" if name != "c3" else f"{budget_frames}-frame slice of one clip (B=1)"}
        except Exception as exc:
            res[name] = {"failed": f"{type(exc).__name__}: {exc}"[:160]}
        del q, k, v, fl, gout
        th.cuda.empty_cache()
    return res

# ------------------------------------------------------------------------------------------------
#   CPU arm: the C restatement of the reference kernels (oracle/, OpenMP over the host cores)
# ------------------------------------------------------------------------------------------------

def cpu_pass(wl, crop, threads):
    """One fwd+bwd pass of the same pipeline on the host.  `crop` = spatial size of the bounded sample."""
    from oracle import ref
    os.environ["OMP_NUM_THREADS"] = str(threads)
    th.set_num_threads(threads)
    w = dict(wl, H=crop, W=crop)
    q, k, v, flows, gout = make_inputs(w, 0, 1, pin=False)
    kw = dict(nheads=w["HD"], stride0=w["stride0"], dist_type=w["dist_type"], itype=w["itype"])
    flt = w["itype"] == "float"
    t0 = time.perf_counter()
    d, i = ref.nls_forward(q, k, flows, w["ws"], w["wt"], w["ps"], w["K"], stride1=w["stride1"],
                           self_action=w["self_action"], **kw)
    d.requires_grad_(True)
    sign = -1.0 if w["dist_type"] == "l2" else 1.0
    wts = th.softmax(sign * w["scale"] * d, -1)
    out, counts = ref.gather_add_forward(v, wts.detach(), i, w["ps"], w["stride0"], itype=w["itype"])
    if wl.get("passes") == "fwd":
        return candidates_per_step(w), time.perf_counter() - t0
    gv, gw, gf = ref.gather_add_backward(gout, v, wts.detach(), i, counts, w["ps"], w["stride0"], itype=w["itype"])
    wts.backward(gw)
    gi = gf if flt else None
    ref.search_backward(q, k, flows, d.grad, gi, i, w["wt"], w["ps"], flows_grad=flt, **kw)
    dt = time.perf_counter() - t0
    return candidates_per_step(w), dt


def cpu_sample_size(wl):
    return min(wl["H"], wl["W"], 128)


def cpu_sample(wl, cores, budget_s=12.0, max_passes=6):
    """Bounded CPU sample: whole passes of the pipeline on the crop until ~budget_s of work is done."""
    crop = cpu_sample_size(wl)
    tot_c, tot_t, n = 0, 0.0, 0
    while n < max_passes and (n == 0 or tot_t < budget_s):
        c, t = cpu_pass(wl, crop, cores)
        tot_c += c; tot_t += t; n += 1
    return tot_c, tot_t, n, crop


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return None
    wl = WORKLOADS[args.workload]
    cores = os.cpu_count() or 1
    crop = cpu_sample_size(wl)
    cpu_pass(dict(wl, T=max(2 * wl["wt"] + 1, 2)), 16, cores)  # load + warm the library
    # one step = one pass of the pipeline on the crop; K and W are honoured unless they would push the run
    # past ~3 minutes of host time (then they are clamped and the line says so)
    _, t1 = cpu_pass(wl, crop, cores)
    warm = 1 + sum(1 for _ in range(max(0, min(args.warmup, int(30 / max(t1, 1e-3))) - 1)) if cpu_pass(wl, crop, cores))
    steps = max(1, min(args.steps, int(150 / max(t1, 1e-3))))
    tot_c, tot_t = 0, 0.0
    for _ in range(steps):
        c, t = cpu_pass(wl, crop, cores)
        tot_c += c
        tot_t += t
    val = tot_c / tot_t / 1e9
    sample = f"{steps} pass(es) of the same pipeline on a {crop}x{crop} crop (T={wl['T']}), OpenMP {cores} threads"
    line = {"impl": "reference", "metric": "search+agg fwd+bwd throughput", "value": val,
            "unit": "Gcandidates/s", "n_gpus": args.gpus, "steps": steps, "warmup": warm,
            "ms_per_step": tot_t / steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": describe(wl, args.workload, 1),
            "cpu_baseline": {"value": val, "unit": "Gcandidates/s", "cores": cores, "kind": "port",
                             "sample": sample},
            "e2e": {"value": val, "unit": "Gcandidates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "the reference has no CPU path (SURVEY.md 0.4); this is oracle/stnls_oracle.c, a C "
                    "restatement of its CUDA kernels, pinned against the reference kernels (oracle/_ref)"}
    print(json.dumps(line), flush=True)
    return line


# ------------------------------------------------------------------------------------------------
#   assemble the JSON line
# ------------------------------------------------------------------------------------------------

def kernel_table(timeline, steps):
    agg = {}
    for name, e0, e1 in timeline:
        ms = e0.elapsed_time(e1)
        a = agg.setdefault(name, [0.0, 0])
        a[0] += ms
        a[1] += 1
    return {n: dict(ms_total=t, launches=c, ms_per_launch=t / c, ms_per_step=t / steps) for n, (t, c) in agg.items()}


def run_extras(args, r):
    """After the main workload, on every rank: the 1080p block (always), the RefineSearch-4K block (N = 1 and 8), and --
    single GPU only -- the measured fp32 FMA peak and the reference's own CUDA kernels on this GPU."""
    if args.blocks == "none":
        return {}
    import stnls_b200  # noqa: F401
    from stnls_b200 import ops
    world, rank = r["world"], r["rank"]
    dev = th.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    out = {}
    groups = {}
    for name, when in (("c4_1080p", True), ("c5", world in (1, 8))):
        if not when or (args.blocks != "all" and name not in args.blocks.split(",")):
            continue
        try:
            out[name] = run_block(name, args, world, rank, dev, steps=2, warmup=1, groups=groups)
        except Exception as exc:   # a block must not take the headline line down with it
            out[name] = {"failed": f"{type(exc).__name__}: {exc}"[:300]}
            th.cuda.synchronize()
            th.cuda.empty_cache()
    if rank == 0:
        out["fp32_fma_peak"] = {"tflops": ops.ffma_peak_tflops(dev), "how": "stnls_ffma_peak: 8 CTAs x 256 threads per SM, 16 "
                                "independent FFMA chains x 4096 steps per thread, CUDA events, best of 5 (csrc/peak.cu)"}
        if world == 1 and not args.no_cpu_baseline:
            out["reference_cuda_on_b200"] = reference_cuda_on_b200(dev)
    return out


def main_block(args):
    """--workload c1 / c4_1080p / c5: one line from the block runner (device-resident inputs; no e2e leg)."""
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not th.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product arm has no CPU path)")
    th.cuda.set_device(local)
    dev = th.device("cuda", local)
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    try:
        if world > 1:
            dist.init_process_group("nccl", device_id=dev)
        with ClockSampler(local) as clk:
            blk = run_block(args.workload, args, world, rank, dev, args.steps, args.warmup)
    finally:
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        os.close(real_stdout)
    if rank == 0:
        wl = WORKLOADS[args.workload]
        line = {"metric": "search+agg throughput (" + blk["passes"] + ")", "value": blk["Gcandidates_per_s"],
                "unit": "Gcandidates/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": blk["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32", "data": blk["data"], "config": dict(describe(wl, args.workload, world), **blk["config"]),
                "frames_per_s": blk["frames_per_s"], "clocks": clk.summary(), "kernels": blk["kernels"],
                "halo": blk["halo"], "peak_device_GiB": blk["peak_device_GiB"]}
        if world == 1 and not args.no_cpu_baseline and not wl.get("refine") and wl["H"] <= 256:
            cores = os.cpu_count() or 1
            c, t, n, crop = cpu_sample(wl, cores)
            line["cpu_baseline"] = {"value": c / t / 1e9, "unit": "Gcandidates/s", "cores": cores, "kind": "port",
                                    "sample": f"{n} pass(es) on a {crop}x{crop} crop, oracle/stnls_oracle.c with OpenMP"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="product", choices=["product", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--halo", default="auto", choices=["auto", "peer", "nccl"])
    ap.add_argument("--blocks", default="all",
                    help="extra blocks after the main workload: all | none | comma list of c4_1080p,c5")
    ap.add_argument("--graph", default="auto", choices=["auto", "off"],
                    help="auto: `value` from CUDA-graph replays of the step when capture works")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "product" else args.warmup

    if args.impl == "reference":
        run_reference(args)
        return
    wl0 = WORKLOADS[args.workload]
    if wl0.get("passes") or wl0.get("Tmax"):
        return main_block(args)

    # stdout carries the one JSON line only: whatever libraries print on fd 1 meanwhile (NCCL's version line at N>1)
    # is sent to stderr, and fd 1 comes back for the final print
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    extra = {}
    try:
        r = run_product(args)
        extra = run_extras(args, r)
    finally:
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        os.close(real_stdout)
    if r["rank"] != 0:
        return
    wl, world, steps = r["wl"], r["world"], args.steps
    cands = candidates_per_step(wl, world)
    eager_ms = r["t_ms"] / steps
    graphed = r["graph_ms"] is not None
    ms_per_step = r["graph_ms"] / steps if graphed else eager_ms
    value = cands / (ms_per_step * 1e-3) / 1e9
    e2e_ms = r["e2e_ms"] / steps
    e2e_val = cands / (e2e_ms * 1e-3) / 1e9

    kt = kernel_table(r["timeline"], steps)
    plan = r["plan"]
    alg = algorithmic(wl, plan.fb - plan.fa)
    main_k = [n for n in kt if n in ("search_forward", "search_backward", "wpsum_forward", "wpsum_backward")]
    dom = max(main_k, key=lambda n: kt[n]["ms_total"])
    peak, peak_src = load_peaks()
    clocks = r["clocks"]
    mhz = clocks.get("sm_mhz") or clocks.get("sm_max_mhz") or 1965
    fma_meas = (extra.get("fp32_fma_peak") or {}).get("tflops")
    fma_peak = fma_meas or 148 * 128 * 2 * mhz * 1e6 / 1e12
    fma_src = ("measured in this run: stnls_ffma_peak microbenchmark (csrc/peak.cu), CUDA events" if fma_meas else
               f"formula 148 SM x 128 lanes x 2 x {mhz} MHz (microbenchmark not run)")
    dur_s = kt[dom]["ms_per_launch"] * 1e-3
    hbm_ach = alg[dom]["bytes"] / dur_s / 1e9
    traffic = load_traffic(args.workload, dom) if world == 1 else None
    roofline = {"kernel": dom, "traffic": traffic,
                "traffic_source": "profiles/traffic.json: dram__bytes_read.sum + dram__bytes_write.sum of one launch, from the "
                                  "committed `ncu --set full` capture of this kernel (not re-measured in this run)",
                "algorithmic_bytes_per_launch": alg[dom]["bytes"], "ms_per_launch": kt[dom]["ms_per_launch"],
                "share_of_step": kt[dom]["ms_per_step"] / ms_per_step,
                "hbm": {"achieved": hbm_ach, "peak": peak, "unit": "GB/s", "frac": hbm_ach / peak, "peak_source": peak_src}}
    if dom == "search_forward":
        # the search forward is bound by the fp32 FMA / issue pipes (arithmetic intensity >> machine balance), not by HBM
        exe = alg[dom]["flops_executed"] / dur_s / 1e12
        roofline.update({"bound": "fp32_fma", "achieved": exe, "peak": fma_peak, "unit": "TFLOP/s", "frac": exe / fma_peak,
                         "peak_source": fma_src,
                         "flops": "executed: 4 per interpolated lattice point + 2 (prod) / 3 (l2) per patch element"})
    else:
        roofline.update({"bound": "hbm", "achieved": hbm_ach, "peak": peak, "unit": "GB/s", "frac": hbm_ach / peak,
                         "peak_source": peak_src,
                         "note": "scatter kernel: limited by gather latency / the L1 pipe at ~21 % occupancy (vector REDs measured at 13-19 % of it, DESIGN.md 3.2); DRAM traffic == algorithmic"})
    fwd = kt.get("search_forward")
    roofline_fma = None
    if fwd:
        af = alg["search_forward"]["flops"] / (fwd["ms_per_launch"] * 1e-3) / 1e12
        ae = alg["search_forward"]["flops_executed"] / (fwd["ms_per_launch"] * 1e-3) / 1e12
        roofline_fma = {"kernel": "search_forward", "bound": "fp32 fma pipe", "achieved": af, "peak": fma_peak,
                        "unit": "TFLOP/s", "frac": af / fma_peak, "peak_source": fma_src,
                        "algorithmic_flops_per_launch": alg["search_forward"]["flops"],
                        "note": "SURVEY.md 8d flop count (8 flops of bilinear work per patch element of every candidate); "
                                "the kernel interpolates each lattice point once, so `executed` is the pipe's real load",
                        "executed": {"achieved": ae, "frac": ae / fma_peak,
                                     "flops_per_launch": alg["search_forward"]["flops_executed"]}}
    for n in kt:
        if n in alg and alg[n]["bytes"]:
            kt[n]["alg_GBs"] = alg[n]["bytes"] / (kt[n]["ms_per_launch"] * 1e-3) / 1e9

    line = {"metric": "search+agg fwd+bwd throughput", "value": value, "unit": "Gcandidates/s",
            "n_gpus": world, "steps": steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": dict(describe(wl, args.workload, world), halo=r["halo"], **({"halo_path": r["halo_path"]} if r.get("halo_path") else {})),
            "frames_per_s": wl["B"] * wl["T"] * world / (ms_per_step * 1e-3),
            "e2e": {"value": e2e_val, "unit": "Gcandidates/s", "h2d_bytes_per_step": r["h2d"],
                    "d2h_bytes_per_step": r["d2h"], "ms_per_step": e2e_ms,
                    "host_link_GBs_per_rank": {"h2d": r.get("h2d_gbs"), "d2h": r.get("d2h_gbs"),
                                               "how": "one timed pinned<->device copy of a step's tensors, all ranks at once"},
                    "latency_ms_per_step": r["lat_ms"] / steps,
                    "through": "stnls_b200.parallel.GraphedStep (two buffer slots, one CUDA graph of the step each)"
                               if r["e2e_from"] == "graph" else "eager Python front end",
                    "eager_ms_per_step": r["e2e_eager_ms"] / steps,
                    "what": "every step: pinned host q,k,v,flows,grad_out -> device, search -> softmax -> aggregation fwd+bwd, "
                            "out + grads of q,k,v,flows -> pinned host; wall clock, one synchronize after the K steps.  "
                            "value / ms_per_step: the K steps pipelined -- the next step's uploads run on a copy-in stream "
                            "while the current step computes, results leave on a copy-out stream.  eager_ms_per_step: the "
                            "same pipeline driven through the eager modules (host-bound when many ranks share the host).  "
                            "latency_ms_per_step: eager, with a barrier + synchronize on both sides of every step"},
            "gpu_launches": r["per_graph"] * steps if graphed else r["launches"],
            "timing": {"value_from": "cuda graph replay of the step (one launch per step; CUDA events around each replay)"
                                     if graphed else "eager enqueue through the Python front end",
                       "ms_per_step_eager": eager_ms, "kernel_launches_per_step": r["per_graph"] if graphed else None,
                       "graphs": r["unit"] if graphed else None,
                       "note": r["graph_note"] or "`kernels` and `roofline` come from the eager pass (per-launch CUDA events)"},
            "clocks": clocks, "roofline": roofline,
            "roofline_fma": roofline_fma, "kernels": kt}
    if world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        cpu_pass(dict(wl, T=max(2 * wl["wt"] + 1, 2)), 16, cores)
        c, t, n, crop = cpu_sample(wl, cores)
        line["cpu_baseline"] = {"value": c / t / 1e9, "unit": "Gcandidates/s", "cores": cores, "kind": "port",
                                "sample": f"{n} pass(es) of the same pipeline on a {crop}x{crop} crop (T={wl['T']}), "
                                          f"oracle/stnls_oracle.c with OpenMP, {t:.1f} s of CPU work"}
    line.update(extra)
    if world > 1:
        line["config"]["host_affinity"] = (f"rank threads bound to the {r['numa']} CPUs NVML lists for their GPU" if r["numa"]
                                           else "not bound (NVML affinity unavailable)")
    print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
