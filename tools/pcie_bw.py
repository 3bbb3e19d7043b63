#!/usr/bin/env python
"""Plain pinned-memory copy bandwidth per rank and in aggregate, H2D and D2H running at the same
time (the traffic pattern of the *_host entry points), to put a number on the host-side ceiling
of the e2e figure.  One process per GPU:
    python tools/pcie_bw.py                       (1 GPU)
    torchrun --nproc-per-node N --master-addr 127.0.0.1 ... tools/pcie_bw.py
Rank 0 prints one JSON line."""
import json
import os
import time

import torch

rank = int(os.environ.get("RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist = None
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
nbytes = 1 << 30
h_in = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
h_out = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
h_in.fill_(1)
d_in = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
d_out = torch.ones(nbytes, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def run(do_h2d, do_d2h, reps=8):
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        if do_h2d:
            with torch.cuda.stream(s1):
                d_in.copy_(h_in, non_blocking=True)
        if do_d2h:
            with torch.cuda.stream(s2):
                h_out.copy_(d_out, non_blocking=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    if dist is not None:
        t = torch.tensor([dt], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
    return reps * nbytes / dt / 1e9


run(True, True, 2)
out = {"n_gpus": world, "bytes_per_copy": nbytes,
       "h2d_only_GBps_per_rank": run(True, False), "d2h_only_GBps_per_rank": run(False, True)}
both = run(True, True)
out["h2d_and_d2h_GBps_per_rank_each_direction"] = both
out["aggregate_both_directions_GBps"] = 2 * both * world
try:
    out["cpus"] = len(os.sched_getaffinity(0))
    import subprocess
    out["topo"] = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True,
                                 text=True).stdout.splitlines()[:12] if rank == 0 else None
except Exception:
    pass
if rank == 0:
    print(json.dumps(out), flush=True)
if dist is not None:
    dist.destroy_process_group()
