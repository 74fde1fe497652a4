#!/usr/bin/env python
"""Host <-> device copy bandwidth with pinned buffers, both directions at once, on every rank at the same time.
Explains the e2e numbers at N>1: usage  python -m torch.distributed.run --nproc-per-node N tools/pcie_bw.py [MiB]"""
import os, sys, time
import torch as th
import torch.distributed as dist

world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
th.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=th.device("cuda", local))
mib = int(sys.argv[1]) if len(sys.argv) > 1 else 24
n = mib << 20
hin, hout = th.empty(n, dtype=th.uint8).pin_memory(), th.empty(n, dtype=th.uint8).pin_memory()
din, dout = th.empty(n, dtype=th.uint8, device="cuda"), th.zeros(n, dtype=th.uint8, device="cuda")
up, down = th.cuda.Stream(), th.cuda.Stream()
for reps in (5, 40):
    if world > 1:
        dist.barrier()
    th.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        with th.cuda.stream(up):
            din.copy_(hin, non_blocking=True)
        with th.cuda.stream(down):
            hout.copy_(dout, non_blocking=True)
    th.cuda.synchronize()
    dt = time.perf_counter() - t0
t = th.tensor([dt], device="cuda", dtype=th.float64)
if world > 1:
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
if int(os.environ.get("RANK", "0")) == 0:
    print(f"ranks={world} {mib} MiB each way x {reps}: {n * reps / float(t[0]) / 1e9:.1f} GB/s per direction per GPU "
          f"(slowest rank), {world * n * reps / float(t[0]) / 1e9:.1f} GB/s per direction over all ranks")
if world > 1:
    dist.destroy_process_group()
