"""Control-plane communicators for the Python front end.

`TorchComm` implements the small collective interface the C++ framework needs
(`tenzing::Communicator`, include/tenzing/communicator.hpp) on top of
``torch.distributed`` - NCCL when CUDA is available, gloo otherwise - and is
handed to C++ through ``_C.PyComm``.  The reference uses MPI_Bcast /
MPI_Allreduce(MAX) / MPI_Barrier for the same purpose.
"""
import os

import torch
import torch.distributed as dist

from .. import _C


class TorchComm:
    """rank/size/barrier/bcast_bytes/allreduce_max/allgather_bytes over a process group."""

    def __init__(self, group=None, device=None):
        if not dist.is_initialized():
            raise RuntimeError("torch.distributed is not initialised")
        self.group = group
        backend = dist.get_backend(group)
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" \
                else torch.device("cpu")
        self.device = device

    def rank(self):
        return dist.get_rank(self.group)

    def size(self):
        return dist.get_world_size(self.group)

    def barrier(self):
        # an all-reduce instead of dist.barrier(): no device_ids warnings, same effect
        t = torch.zeros(1, device=self.device)
        dist.all_reduce(t, group=self.group)
        if t.is_cuda:
            torch.cuda.current_stream().synchronize()

    def bcast_bytes(self, data, root):
        t = torch.frombuffer(bytearray(data), dtype=torch.uint8).to(self.device) if len(data) \
            else torch.zeros(0, dtype=torch.uint8, device=self.device)
        if len(data):
            dist.broadcast(t, src=dist.get_global_rank(self.group, root) if self.group else root,
                           group=self.group)
        return bytes(t.cpu().numpy().tobytes())

    def allreduce_max(self, vals):
        t = torch.tensor(vals, dtype=torch.float64, device=self.device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        return t.cpu().tolist()

    def allgather_bytes(self, data):
        n = len(data)
        mine = torch.frombuffer(bytearray(data), dtype=torch.uint8).to(self.device) if n \
            else torch.zeros(0, dtype=torch.uint8, device=self.device)
        out = [torch.empty(n, dtype=torch.uint8, device=self.device) for _ in range(self.size())]
        if n:
            dist.all_gather(out, mine, group=self.group)
        return [bytes(o.cpu().numpy().tobytes()) for o in out]


def init_process_group_from_env(backend=None):
    """Initialise torch.distributed from RANK/WORLD_SIZE/MASTER_* (torchrun) if needed.

    Returns (rank, world_size, local_rank)."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    if torch.cuda.is_available():
        torch.cuda.set_device(local % torch.cuda.device_count())
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        kwargs = {}
        if backend == "nccl":
            kwargs["device_id"] = torch.device("cuda", torch.cuda.current_device())
        dist.init_process_group(backend=backend, rank=rank, world_size=world, **kwargs)
    return rank, world, local


def make_comm(group=None):
    """The communicator for this process: torch.distributed if initialised, else world size 1."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        comm = _C.PyComm(TorchComm(group))
        _C.set_log_rank(comm.rank())
        return comm
    return _C.SelfComm()
