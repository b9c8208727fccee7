"""Device-resident entry points (cramgpu_*_dev) on torch tensors.

torch is used only for what it is good at here: owning HBM buffers and CUDA streams.  The kernels are
libcramgpu's.  These wrappers are what bench.py times for the HBM-resident `value`.
"""
import ctypes

import numpy as np
import torch

from . import _lib


def _round16(n):
    return (int(n) + 15) & ~15


def pack_blocks(blocks, device):
    """Concatenate host byte blocks into one device buffer, each starting at a multiple of 16.
    Returns (d_in u8 tensor, in_off np.uint64, in_len np.uint32)."""
    lens = np.array([len(b) for b in blocks], dtype=np.uint32)
    offs = np.zeros(len(blocks), dtype=np.uint64)
    total = 0
    for i, ln in enumerate(lens):
        offs[i] = total
        total += _round16(ln)
    host = np.zeros(max(total, 16), dtype=np.uint8)
    for i, b in enumerate(blocks):
        if lens[i]:
            host[int(offs[i]):int(offs[i]) + int(lens[i])] = np.frombuffer(bytes(b), dtype=np.uint8) \
                if not isinstance(b, np.ndarray) else b
    return torch.from_numpy(host).to(device), offs, lens


class DeviceBatchResult(object):
    def __init__(self, out, off, length, status):
        self.out, self.off, self.len, self.status = out, off, length, status

    def to_host(self):
        out = self.out.cpu().numpy()
        off = self.off.cpu().numpy().view(np.uint64)
        ln = self.len.cpu().numpy().view(np.uint32)
        st = self.status.cpu().numpy()
        return [out[int(off[i]):int(off[i]) + int(ln[i])].tobytes() for i in range(len(ln))], st


def _stream_ptr():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def rans_encode_dev(ctx, d_in, in_off, in_len, order, out=None, out_cap=None):
    """cramgpu_rans_encode_dev on torch's current stream.  Returns DeviceBatchResult (asynchronous)."""
    n = len(in_len)
    dev = d_in.device
    in_off = np.ascontiguousarray(in_off, dtype=np.uint64)
    in_len = np.ascontiguousarray(in_len, dtype=np.uint32)
    order = np.ascontiguousarray(order, dtype=np.uint8)
    if out is None:
        if out_cap is None:
            out_cap = int(sum(_round16(9 + 1040 + int(l) + int(l) // 4 + 96 + (263936 if o else 0))
                              for l, o in zip(in_len, order)))
        out = torch.empty(max(int(out_cap), 16), dtype=torch.uint8, device=dev)
    off = torch.empty(n + 1, dtype=torch.int64, device=dev)
    ln = torch.empty(max(n, 1), dtype=torch.int32, device=dev)
    st = torch.empty(max(n, 1), dtype=torch.int32, device=dev)
    rc = ctx.lib.cramgpu_rans_encode_dev(
        ctx.handle, n, ctypes.c_void_p(d_in.data_ptr()), in_off.ctypes.data_as(_lib.u64p),
        in_len.ctypes.data_as(_lib.u32p), order.ctypes.data_as(_lib.u8p), ctypes.c_void_p(out.data_ptr()),
        ctypes.c_uint64(out.numel()), ctypes.c_void_p(off.data_ptr()), ctypes.c_void_p(ln.data_ptr()),
        ctypes.c_void_p(st.data_ptr()), _stream_ptr())
    ctx.check(rc)
    return DeviceBatchResult(out, off, ln[:n], st[:n])


def rans_decode_dev(ctx, d_in, in_off, in_len, prefixes, out=None):
    """cramgpu_rans_decode_dev on torch's current stream.  `prefixes` is an (n, 9) uint8 host array."""
    n = len(in_len)
    dev = d_in.device
    in_off = np.ascontiguousarray(in_off, dtype=np.uint64)
    in_len = np.ascontiguousarray(in_len, dtype=np.uint32)
    prefixes = np.ascontiguousarray(prefixes, dtype=np.uint8).reshape(n, 9)
    if out is None:
        raw = prefixes[:, 5].astype(np.uint64) | (prefixes[:, 6].astype(np.uint64) << 8) | \
            (prefixes[:, 7].astype(np.uint64) << 16) | (prefixes[:, 8].astype(np.uint64) << 24)
        cap = int(sum(_round16(r) for r in raw)) + 16
        out = torch.empty(cap, dtype=torch.uint8, device=dev)
    off = torch.empty(n + 1, dtype=torch.int64, device=dev)
    ln = torch.empty(max(n, 1), dtype=torch.int32, device=dev)
    st = torch.empty(max(n, 1), dtype=torch.int32, device=dev)
    rc = ctx.lib.cramgpu_rans_decode_dev(
        ctx.handle, n, ctypes.c_void_p(d_in.data_ptr()), in_off.ctypes.data_as(_lib.u64p),
        in_len.ctypes.data_as(_lib.u32p), prefixes.ctypes.data_as(_lib.u8p), ctypes.c_void_p(out.data_ptr()),
        ctypes.c_uint64(out.numel()), ctypes.c_void_p(off.data_ptr()), ctypes.c_void_p(ln.data_ptr()),
        ctypes.c_void_p(st.data_ptr()), _stream_ptr())
    ctx.check(rc)
    return DeviceBatchResult(out, off, ln[:n], st[:n])
