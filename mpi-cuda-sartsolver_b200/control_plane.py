"""Process-level plumbing for one-process-per-GPU runs (what MPI_Init / mpirun give the reference,
/root/reference/source/main.cpp:63-68): rank discovery from the torchrun environment, a small host-side
collective set over torch.distributed (gloo) to ship the engine's 128-byte communicator id and to combine
timings, and the reference's row partition. No tensor of the data path goes through here -- the engines talk
to each other over their own NCCL communicator (sart_comm_init)."""
import datetime
import os


class ControlPlane:
    def __init__(self, rank=0, world=1, local_rank=0, dist=None):
        self.rank, self.world, self.local_rank, self._dist = rank, world, local_rank, dist

    @classmethod
    def from_env(cls, timeout_s=600):
        """RANK / WORLD_SIZE / LOCAL_RANK / MASTER_ADDR / MASTER_PORT as torchrun sets them; world 1 needs none."""
        world = int(os.environ.get("WORLD_SIZE", "1"))
        rank = int(os.environ.get("RANK", "0"))
        local_rank = int(os.environ.get("LOCAL_RANK", str(rank)))
        if world == 1:
            return cls(0, 1, local_rank if "LOCAL_RANK" in os.environ else 0)
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        if not dist.is_initialized():
            dist.init_process_group("gloo", rank=rank, world_size=world, timeout=datetime.timedelta(seconds=timeout_s))
        return cls(rank, world, local_rank, dist)

    def barrier(self):
        if self._dist:
            self._dist.barrier()

    def broadcast_bytes(self, payload, src=0):
        """Rank ``src`` passes bytes, the others pass None; every rank gets the bytes."""
        if not self._dist:
            return payload
        box = [payload if self.rank == src else None]
        self._dist.broadcast_object_list(box, src=src)
        return box[0]

    def _reduce(self, value, op):
        if not self._dist:
            return float(value)
        import torch
        t = torch.tensor([float(value)], dtype=torch.float64)
        self._dist.all_reduce(t, op=op)
        return float(t[0])

    def max(self, value):
        import torch.distributed as dist
        return self._reduce(value, dist.ReduceOp.MAX) if self._dist else float(value)

    def sum(self, value):
        import torch.distributed as dist
        return self._reduce(value, dist.ReduceOp.SUM) if self._dist else float(value)

    def gather(self, obj):
        """Every rank contributes a picklable object; every rank gets the list in rank order."""
        if not self._dist:
            return [obj]
        out = [None] * self.world
        self._dist.all_gather_object(out, obj)
        return out

    def shutdown(self):
        if self._dist and self._dist.is_initialized():
            self._dist.destroy_process_group()


def shard_range(npixel, nranks, rank):
    """offset_pixel / npixel_local of source/main.cpp:67-68 (host-side twin of sart_shard_range, for callers that
    have to slice frames before any library is loaded)."""
    base, rem = divmod(npixel, nranks)
    return rank * base + min(rank, rem), base + (1 if rank < rem else 0)
