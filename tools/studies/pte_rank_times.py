"""Per-rank PTE RhoT step times, the bench's PTE step re-staged (same state generator, same
MixParams, same reset-copy + solve) so that each rank's own CUDA-event time is printed:
    python tools/studies/pte_rank_times.py                      (1 process)
    torchrun --nproc-per-node N ... tools/studies/pte_rank_times.py
START_OVERRIDE=k makes every rank solve the cells of rank k (separates "which cells" from
"how many processes")."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import __graft_entry__ as g

rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist = None
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=dev)
pkg = g.load_package(); L = pkg._lib; W = pkg.workloads
npte = 1 << int(os.environ.get("LOG2_CELLS", "27"))
g3 = pkg.closure.set_eos_3mat_analytic()
sie0 = []
for m in range(3):
    o = torch.empty(1, dtype=torch.float64, device=dev)
    g3[m].InternalEnergyFromDensityTemperature(
        torch.tensor([W.PTE_RHO[m]], dtype=torch.float64, device=dev),
        torch.tensor([W.PTE_T], dtype=torch.float64, device=dev), o, 1)
    sie0.append(float(o.item()))
srank = int(os.environ.get("START_OVERRIDE", rank))
state0 = W.pte_state_torch(npte, sie0, dev, start=srank * npte)
state = [a.clone() for a in state0]
status = torch.zeros(npte, dtype=torch.int64, device=dev)
prm = pkg.closure.test_pte_params()


def pte():
    for a, b in zip(state, state0):
        a.copy_(b)
    pkg.closure.PTESolverRhoT(g3, *state, params=prm, status=status)


for _ in range(2):
    pte()
torch.cuda.synchronize()
if dist is not None:
    dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(3):
    pte()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 3
niter = ((status >> 8) & 0xffff).double().mean().item()
print(f"rank {rank}/{world} cells-of-rank {srank}: {ms:.2f} ms/step = {npte / ms / 1e3:.2f} M solves/s; "
      f"converged {int((status & 1).sum())}/{npte}, mean status>>8&0xffff {niter:.3f}", flush=True)
if dist is not None:
    dist.barrier(); dist.destroy_process_group()
