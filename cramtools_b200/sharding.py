"""Multi-GPU layout of the hot path: containers are independent units and are dealt round-robin to the
ranks; the only exchange is an all-gather of per-container compressed sizes, from which every rank derives
the byte offset of each of its containers in the output file.

Reference behaviour being reproduced: the writer keeps a running byte `offset` and assigns
`container.offset = offset` before writing containers strictly in order
(CRAMContainerStreamWriter.java:229-230); its multi-threaded variants convert containers concurrently and
re-serialise them by ordinal (CRAMContainerAsynchWriter.java:128-189, paralell/OrderingConveyer.java:24-38).
"""
import torch
import torch.distributed as dist


def container_owner(ordinal, world):
    """rank that converts container `ordinal`"""
    return ordinal % world


def local_ordinals(n_total, rank, world):
    """global ordinals of the containers rank `rank` owns, in local order"""
    return list(range(rank, n_total, world))


def allgather_offsets(local_sizes, rank, world, base=0, group=None):
    """local_sizes: int64 tensor, compressed bytes of this rank's containers (local order; every rank holds the
    same count, pad with zeros otherwise).  Returns the int64 tensor of global byte offsets of this rank's
    containers: container with global ordinal g = j*world + rank starts at base + sum(sizes of ordinals < g).
    NCCL on GPUs, gloo on CPU (tests)."""
    if world == 1:
        return base + torch.cumsum(local_sizes, 0) - local_sizes
    n = local_sizes.numel()
    gathered = torch.empty(world * n, dtype=local_sizes.dtype, device=local_sizes.device)
    dist.all_gather_into_tensor(gathered, local_sizes.contiguous().view(-1), group=group)
    flat = gathered.view(world, n).t().reshape(-1)   # ordinal order: j*world + r
    excl = torch.cumsum(flat, 0) - flat
    return base + excl.view(-1, world)[:, rank].contiguous()
