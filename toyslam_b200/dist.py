"""torch.distributed plumbing for the multi-GPU path: one process per GPU; the library's own NCCL
communicator is bootstrapped from a unique id that rank 0 creates and broadcasts."""
import torch
import torch.distributed as dist
from . import capi


def broadcast_unique_id(rank: int) -> bytes:
    t = torch.zeros(128, dtype=torch.uint8)
    if rank == 0:
        t = torch.frombuffer(bytearray(capi.nccl_unique_id()), dtype=torch.uint8).clone()
    if dist.get_backend() == "nccl":
        t = t.cuda()
    dist.broadcast(t, 0)
    return bytes(t.cpu().numpy().tobytes())
