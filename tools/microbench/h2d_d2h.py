"""Pinned host<->device copy bandwidth with 1..N GPUs copying at the same time (the bound under the N>1 e2e number).

torchrun --nproc-per-node N tools/microbench/h2d_d2h.py [MB ...]   (default 100 and 805 MB: the 16384^2 RGB image)
Every rank copies its own pinned buffer to / from its own GPU; the ranks start together (barrier), time with CUDA events
and report per-rank GB/s plus the aggregate = total bytes / slowest rank's time, for H2D alone, D2H alone and both at once
(two streams).  One JSON line per size from rank 0.
"""
import json
import os
import sys

import torch
import torch.distributed as dist

local = int(os.environ.get("LOCAL_RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
sizes = [int(a) for a in sys.argv[1:]] or [100, 805]


def sync():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()


def allmax(x):
    t = torch.tensor([x], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


for mb in sizes:
    n = mb * 1000 * 1000
    h_a, h_b = torch.empty(n, dtype=torch.uint8).pin_memory(), torch.empty(n, dtype=torch.uint8).pin_memory()
    d_a, d_b = torch.empty(n, dtype=torch.uint8, device=dev), torch.empty(n, dtype=torch.uint8, device=dev)
    s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    reps = 5
    out = {"check": "pinned host<->device copies", "n_gpus": world, "mb_per_copy": mb, "reps": reps}

    def run(h2d, d2h):
        sync()
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        e0.record(torch.cuda.current_stream())
        s1.wait_event(e0); s2.wait_event(e0)
        for _ in range(reps):
            if h2d:
                with torch.cuda.stream(s1):
                    d_a.copy_(h_a, non_blocking=True)
            if d2h:
                with torch.cuda.stream(s2):
                    h_b.copy_(d_b, non_blocking=True)
        e1.record(s1); e2.record(s2)
        torch.cuda.synchronize()
        return max(e0.elapsed_time(e1), e0.elapsed_time(e2)) / reps      # ms per round of copies

    for name, (a, b) in {"h2d": (True, False), "d2h": (False, True), "both": (True, True)}.items():
        run(a, b)
        ms = allmax(run(a, b))
        per_dir = n / (ms * 1e-3) / 1e9
        out[name] = {"ms": round(ms, 3), "per_gpu_GBps_per_direction": round(per_dir, 2),
                     "aggregate_GBps_per_direction": round(per_dir * world, 2)}
    if local == 0 and int(os.environ.get("RANK", "0")) == 0:
        print(json.dumps(out), flush=True)
    del h_a, h_b, d_a, d_b
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
