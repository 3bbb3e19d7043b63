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
late = os.environ.get("ALLOC_BEFORE_INIT") == "1"  # state arrays first, communicator afterwards
if world > 1 and not late:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=dev)
    t_ = torch.ones(1, device=dev); dist.all_reduce(t_); torch.cuda.synchronize()
if os.environ.get("ENABLE_PEER") == "1":  # what a 2-rank NCCL communicator does to the context
    import ctypes
    rt = ctypes.CDLL("libcudart.so.12")
    torch.zeros(1, device=dev)
    peer = 1 - local
    rc = rt.cudaDeviceEnablePeerAccess(ctypes.c_int(peer), ctypes.c_uint(0))
    print(f"cudaDeviceEnablePeerAccess({peer}) -> {rc}", flush=True)
if os.environ.get("STACK_LIMIT"):  # cudaLimitStackSize = 0
    import ctypes
    rt = ctypes.CDLL("libcudart.so.12")
    torch.zeros(1, device=dev)
    rc = rt.cudaDeviceSetLimit(ctypes.c_int(0), ctypes.c_size_t(int(os.environ["STACK_LIMIT"])))
    print(f"cudaDeviceSetLimit(stack, {os.environ['STACK_LIMIT']}) -> {rc}", flush=True)
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
if world > 1 and late:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=dev)
    t_ = torch.ones(1, device=dev); dist.all_reduce(t_); torch.cuda.synchronize()
if True:
    import ctypes
    rt_ = ctypes.CDLL("libcudart.so.12"); v_ = ctypes.c_size_t(0)
    rt_.cudaDeviceGetLimit(ctypes.byref(v_), ctypes.c_int(0))
    print(f"rank {rank} cudaLimitStackSize now {v_.value}", flush=True)
print(f"rank {rank} state ptrs mod 2MiB:", [a.data_ptr() % (1 << 21) for a in state], "first ptr", hex(state[0].data_ptr()),
      "reserved MiB", torch.cuda.memory_reserved() >> 20, flush=True)


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
