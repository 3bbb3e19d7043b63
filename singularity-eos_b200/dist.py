"""Multi-GPU plumbing for the cell-parallel path.

Cells are independent: the global index range [0, N) is cut into `world` contiguous shards,
one per rank / GPU, tables are replicated per device at EOS-create time and the data path
needs no collective.  The only exchange is the optional all-reduce of the <= 64-entry
convergence-statistics vector (NCCL on GPUs, gloo in the CPU tests).
"""


def shard_range(n_global, rank, world):
    """Contiguous [lo, hi) of rank `rank`; sizes differ by at most one cell."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    base, rem = divmod(int(n_global), int(world))
    lo = rank * base + min(rank, rem)
    hi = lo + base + (1 if rank < rem else 0)
    return lo, hi


def env_rank_world():
    import os

    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def allreduce_stats(stats):
    """Sum the uint64[SGX_PTE_NSTATS] statistics vector (held as an int64 tensor) over ranks."""
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.SUM)
    return stats


def max_over_ranks(seconds, device):
    import torch
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(seconds)
    t = torch.tensor([seconds], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


class StatsComm:
    """The C ABI's own communicator for `sgx_stats_allreduce` (NCCL bound at run time inside
    the library).  Rank 0 makes the id; `exchange` ships 128 bytes to the other ranks (any
    transport: torch.distributed object broadcast here, MPI_Bcast in a host code)."""

    def __init__(self, rank, world, exchange=None):
        import ctypes as C

        from . import _lib

        self._lib = _lib
        lib = _lib.load()
        buf = C.create_string_buffer(128)
        if rank == 0:
            _lib.check(lib.sgx_comm_unique_id(buf))
        raw = bytes(buf.raw)
        if world > 1:
            if exchange is None:
                import torch.distributed as dist

                box = [raw]
                dist.broadcast_object_list(box, src=0)
                raw = box[0]
            else:
                raw = exchange(raw)
        self.h = C.c_void_p()
        _lib.check(lib.sgx_comm_create(C.create_string_buffer(raw, 128), world, rank,
                                       C.byref(self.h)))

    def allreduce(self, stats, stream=None):
        """In-place sum of a device int64/uint64 tensor over the ranks."""
        self._lib.check(self._lib.load().sgx_stats_allreduce(self.h, stats.data_ptr(),
                                                             stats.numel(), stream))
        return stats

    def close(self):
        if self.h:
            self._lib.load().sgx_comm_destroy(self.h)
            self.h = None
