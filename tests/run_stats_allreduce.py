"""Multi-GPU check of sgx_stats_allreduce (run under torchrun, one rank per GPU; also valid
with a single rank).  Each rank solves its shard of the PTE workload, the statistics vectors
are summed through the C ABI's own NCCL communicator and compared with torch.distributed's sum.
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tests/run_stats_allreduce.py
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g  # noqa: E402


def main():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    import torch.distributed as dist

    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = g.load_package()
    L, W = pkg._lib, pkg.workloads
    dev = torch.device("cuda", local)
    n_global = 1 << 16
    lo, hi = pkg.dist.shard_range(n_global, rank, world)
    g3 = pkg.closure.set_eos_3mat_analytic()
    sie0 = []
    for m in range(3):
        o = torch.empty(1, dtype=torch.float64, device=dev)
        g3[m].InternalEnergyFromDensityTemperature(
            torch.tensor([W.PTE_RHO[m]], dtype=torch.float64, device=dev),
            torch.tensor([W.PTE_T], dtype=torch.float64, device=dev), o, 1)
        sie0.append(float(o.item()))
    st = W.pte_state_torch(hi - lo, sie0, dev, start=lo)
    stats = torch.zeros(L.PTE_NSTATS, dtype=torch.int64, device=dev)
    pkg.closure.PTESolverRhoT(g3, *st, params=pkg.closure.MixParams(), stats=stats)
    want = stats.clone()
    if world > 1:
        dist.all_reduce(want)
    comm = pkg.dist.StatsComm(rank, world)
    comm.allreduce(stats)
    torch.cuda.synchronize()
    comm.close()
    ok = bool((stats == want).all().item()) and int(stats[0].item()) == n_global
    print(f"rank {rank}/{world}: ncell {int(stats[0])} converged {int(stats[1])} "
          f"{'OK' if ok else 'MISMATCH'}", flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
