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


# ------------------------------------------------------------------------------------------- record-feature work

def dev_to_numpy(ctx, ptr, n, dtype):
    """copy n elements of `dtype` from a raw device pointer (a cramgpu_features field) to a numpy array"""
    out = np.empty(int(n), dtype=dtype)
    if out.size:
        ctx.check(ctx.lib.cramgpu_copy_to_host(ctx.handle, ctypes.c_void_p(out.ctypes.data), ctypes.c_void_p(int(ptr)),
                                               ctypes.c_uint64(out.nbytes), _stream_ptr()))
    return out


class DeviceReads(object):
    """a batch of containers resident in HBM (torch tensors) + the small host tables the C ABI wants"""

    def __init__(self, rd, ref, cont_off, device, ref_off=None, ref_len=None):
        from . import records as rec
        self.n = len(rd['pos'])
        self.cont_off = np.ascontiguousarray(cont_off, np.uint64)
        self.nc = self.cont_off.size - 1
        ref = np.ascontiguousarray(ref, np.uint8)
        self.ref_off = np.ascontiguousarray(np.zeros(self.nc) if ref_off is None else ref_off, np.uint64)
        self.ref_len = np.ascontiguousarray(np.full(self.nc, ref.size) if ref_len is None else ref_len, np.uint64)
        t = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a, dt).view(np.uint8) if np.dtype(dt).itemsize > 1 else np.ascontiguousarray(a, dt)).to(device)
        self.t = dict(ref=t(ref, np.uint8), bases=t(rd['bases'], np.uint8), quals=t(rd['quals'], np.uint8),
                      read_off=t(rd['read_off'], np.uint64), cigar=t(rd['cigar'], np.uint32), cigar_off=t(rd['cigar_off'], np.uint64),
                      pos=t(rd['pos'], np.int32), flag=t(rd['flag'], np.uint16), mapq=t(rd['mapq'], np.uint8))
        R = rec.Reads()
        R.n_reads = self.n; R.n_containers = self.nc
        R.cont_off = self.cont_off.ctypes.data_as(_lib.u64p); R.ref_off = self.ref_off.ctypes.data_as(_lib.u64p)
        R.ref_len = self.ref_len.ctypes.data_as(_lib.u64p); R.ref_total = ref.size
        for k, v in self.t.items():
            setattr(R, k, ctypes.c_void_p(v.data_ptr()))
        R.rflags = None
        self.struct = R


def features_encode_dev(ctx, dreads, policies=None, n_policies=-1):
    """cramgpu_features_encode_dev on torch's current stream.  Returns (cramgpu_features struct with DEVICE
    pointers owned by ctx, stream_off, stream_len) -- the two tables as (n_containers, 12) numpy arrays."""
    from . import records as rec
    F = rec.Features()
    soff = np.zeros((dreads.nc, 12), np.uint64)
    slen = np.zeros((dreads.nc, 12), np.uint64)
    rc = ctx.lib.cramgpu_features_encode_dev(ctx.handle, ctypes.byref(dreads.struct), policies, n_policies, ctypes.byref(F),
                                             soff.ctypes.data_as(_lib.u64p), slen.ctypes.data_as(_lib.u64p), _stream_ptr())
    ctx.check(rc)
    return F, soff, slen
