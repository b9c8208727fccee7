"""Host-side mirror of the reference's block-compressor plug-in point for rANS.

Mirrors htsjdk.samtools.cram.io.ExternalCompression [ExternalCompression.java:89-113] and
htsjdk.samtools.cram.encoding.ExternalCompressor.createRANS [ExternalCompressor.java:76-80] -- same
names, argument meaning and error behaviour -- on top of the C ABI (include/cramgpu.h).  The JVM is not
available in this image, so this Python layer plays the role the Java shadow class of INTEGRATION.md
plays in production; both bind exactly the same exported symbols.

Only rANS goes to the GPU.  gzip/bzip2/xz blocks stay on the Java host by design.
"""
import ctypes

import numpy as np

from . import _lib
from ._lib import CramGpuError, u8p

__all__ = ['ExternalCompression', 'ExternalCompressor', 'ORDER_ZERO', 'ORDER_ONE', 'rans_bound']

ORDER_ZERO = 0   # RANS.ORDER.ZERO
ORDER_ONE = 1    # RANS.ORDER.ONE

_MESSAGES = {   # wording of the reference's RuntimeExceptions [RANS.java:37,44,55]
    _lib.ERR_INPUT_LENGTH: 'Incorrect input length.',
    _lib.ERR_UNKNOWN_ORDER: 'Unknown rANS order: %s',
    _lib.ERR_OUT_TOO_SMALL: 'Output buffer too small to fit %s bytes.',
    _lib.ERR_CORRUPT: 'Corrupt rANS block',
}


def rans_bound(n):
    """RANS.allocateIfNeeded [RANS.java:79-85]"""
    return int(_lib.load().cramgpu_rans_bound(int(n)))


def _as_u8(b):
    if isinstance(b, np.ndarray):
        return np.ascontiguousarray(b, dtype=np.uint8)
    return np.frombuffer(bytes(b), dtype=np.uint8)


def _ptr_array(arrs):
    n = len(arrs)
    pa = (u8p * n)()
    for i, a in enumerate(arrs):
        pa[i] = a.ctypes.data_as(u8p) if a.size else ctypes.cast(None, u8p)
    return pa


class ExternalCompression(object):
    """Static helpers with the reference's names; the *_batch forms carry all rANS blocks of >= 1 slices."""

    @staticmethod
    def rans_batch(blocks, orders, ctx=None):
        ctx = ctx or _lib.default_context()
        ins = [_as_u8(b) for b in blocks]
        n = len(ins)
        if n == 0:
            return []
        in_len = np.array([a.size for a in ins], dtype=np.uint32)
        order = np.array([int(o) & 0xFF for o in orders], dtype=np.uint8)
        for o in orders:
            if int(o) not in (0, 1):
                raise CramGpuError(_lib.ERR_UNKNOWN_ORDER, _MESSAGES[_lib.ERR_UNKNOWN_ORDER] % int(o))
        caps = np.array([rans_bound(a.size) for a in ins], dtype=np.uint32)
        outs = [np.empty(int(c), dtype=np.uint8) for c in caps]
        out_len = np.zeros(n, dtype=np.uint32)
        status = np.zeros(n, dtype=np.int32)
        rc = ctx.lib.cramgpu_rans_encode_batch(
            ctx.handle, n, _ptr_array(ins), in_len.ctypes.data_as(_lib.u32p), order.ctypes.data_as(u8p),
            _ptr_array(outs), caps.ctypes.data_as(_lib.u32p), out_len.ctypes.data_as(_lib.u32p),
            status.ctypes.data_as(_lib.i32p))
        if rc != _lib.OK:
            raise CramGpuError(rc, ctx.last_error())
        return [outs[i][:out_len[i]].tobytes() for i in range(n)]

    @staticmethod
    def unrans_batch(blobs, ctx=None):
        ctx = ctx or _lib.default_context()
        ins = [_as_u8(b) for b in blobs]
        n = len(ins)
        if n == 0:
            return []
        in_len = np.array([a.size for a in ins], dtype=np.uint32)
        raw = np.zeros(n, dtype=np.uint32)
        status = np.zeros(n, dtype=np.int32)
        pin = _ptr_array(ins)
        rc = ctx.lib.cramgpu_rans_raw_sizes(n, pin, in_len.ctypes.data_as(_lib.u32p), raw.ctypes.data_as(_lib.u32p),
                                            status.ctypes.data_as(_lib.i32p))
        if rc != _lib.OK:
            i = int(np.nonzero(status)[0][0])
            msg = _MESSAGES.get(rc, 'error')
            if rc == _lib.ERR_UNKNOWN_ORDER:
                msg = msg % int(ins[i][0])
            raise CramGpuError(rc, msg)
        outs = [np.empty(max(int(r), 1), dtype=np.uint8) for r in raw]
        out_len = np.zeros(n, dtype=np.uint32)
        rc = ctx.lib.cramgpu_rans_decode_batch(
            ctx.handle, n, pin, in_len.ctypes.data_as(_lib.u32p), _ptr_array(outs), raw.ctypes.data_as(_lib.u32p),
            out_len.ctypes.data_as(_lib.u32p), status.ctypes.data_as(_lib.i32p))
        if rc != _lib.OK:
            raise CramGpuError(rc, ctx.last_error())
        return [outs[i][:out_len[i]].tobytes() for i in range(n)]

    @staticmethod
    def rans(data, order, ctx=None):
        """byte[] rans(byte[] data, int order) / rans(byte[], RANS.ORDER) [ExternalCompression.java:89-109]"""
        return ExternalCompression.rans_batch([data], [order], ctx)[0]

    @staticmethod
    def unrans(data, ctx=None):
        """byte[] unrans(byte[] data) [ExternalCompression.java:112-113]"""
        return ExternalCompression.unrans_batch([data], ctx)[0]


class ExternalCompressor(object):
    """ExternalCompressor.createRANS(ORDER) [ExternalCompressor.java:76-80]: an object with compress(byte[])."""

    def __init__(self, order, ctx=None):
        self.order = int(order)
        self.ctx = ctx

    @staticmethod
    def createRANS(order, ctx=None):
        return ExternalCompressor(order, ctx)

    def compress(self, data):
        return ExternalCompression.rans(data, self.order, self.ctx)
