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


# ---------------------------------------------------------------------------------------------- through the C ABI
import ctypes as _C

import numpy as _np

from . import _lib as _L


class ContainerSize(_C.Structure):
    _fields_ = [('compressed_bytes', _C.c_uint64), ('n_bases', _C.c_uint64), ('n_records', _C.c_uint32), ('reserved', _C.c_uint32)]


ALLGATHER_FN = _C.CFUNCTYPE(_C.c_int, _C.c_void_p, _C.c_void_p, _C.c_void_p, _C.c_uint64)
_L.SIGNATURES['cramgpu_allgather_sizes'] = (_C.c_int, [_C.c_void_p, _C.c_uint32, _C.POINTER(ContainerSize), _C.c_uint32, _C.c_uint32,
                                                       ALLGATHER_FN, _C.c_void_p, _C.c_uint64, _C.c_uint64, _L.u64p, _L.u64p,
                                                       _C.POINTER(ContainerSize)])


def allgather_sizes(compressed_bytes, n_records, n_bases, rank, world, base_offset=0, base_record=0, group=None):
    """cramgpu_allgather_sizes with torch.distributed as the exchange (NCCL needs device tensors: the callback stages the few
    bytes through the current CUDA device then; gloo takes them as they are).  Arguments: per local container.
    -> (offsets, record counters, totals dict)"""
    lib = _L.load()
    n = len(compressed_bytes)
    local = (ContainerSize * max(n, 1))()
    for i in range(n):
        local[i].compressed_bytes = int(compressed_bytes[i]); local[i].n_records = int(n_records[i]); local[i].n_bases = int(n_bases[i])

    def gather(user, send, recv, nbytes):
        try:
            nb = int(nbytes)
            src = _np.frombuffer((_C.c_uint8 * nb).from_address(send), _np.uint8)
            dev = 'cuda' if dist.get_backend(group) == 'nccl' else 'cpu'
            t_in = torch.from_numpy(src.copy()).to(dev)
            t_out = torch.empty(world * nb, dtype=torch.uint8, device=dev)
            dist.all_gather_into_tensor(t_out, t_in, group=group)
            _C.memmove(recv, t_out.cpu().numpy().ctypes.data, world * nb)
            return 0
        except Exception:      # noqa: BLE001
            return 1
    cb = ALLGATHER_FN(gather)
    off = _np.zeros(max(n, 1), _np.uint64); rec = _np.zeros(max(n, 1), _np.uint64)
    tot = ContainerSize()
    rc = lib.cramgpu_allgather_sizes(None, n, local, rank, world, cb, None, base_offset, base_record,
                                     off.ctypes.data_as(_L.u64p), rec.ctypes.data_as(_L.u64p), _C.byref(tot))
    if rc != _L.OK:
        raise _L.CramGpuError(rc, 'cramgpu_allgather_sizes failed')
    return off[:n], rec[:n], {'compressed_bytes': int(tot.compressed_bytes), 'n_records': int(tot.n_records), 'n_bases': int(tot.n_bases)}
