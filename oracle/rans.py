"""ctypes front-end of the C oracle (oracle/rans_oracle.c).  TEST INFRASTRUCTURE ONLY."""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, '_build', 'liboracle.so')


def build(force=False):
    srcs = [os.path.join(_HERE, f) for f in ('rans_oracle.c', 'records_oracle.c', 'pipeline_oracle.c', 'Makefile')]
    if force or not os.path.exists(_SO) or any(
            os.path.exists(s) and os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs):
        subprocess.check_call(['make', '-s', '-C', _HERE] + (['-B'] if force else []))
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
        u8p = ctypes.POINTER(ctypes.c_uint8)
        u32p = ctypes.POINTER(ctypes.c_uint32)
        _lib.rans_oracle_bound.restype = ctypes.c_uint32
        _lib.rans_oracle_bound.argtypes = [ctypes.c_uint32]
        _lib.rans_oracle_compress.restype = ctypes.c_int
        _lib.rans_oracle_compress.argtypes = [u8p, ctypes.c_uint32, ctypes.c_int, u8p, ctypes.c_uint32, u32p]
        _lib.rans_oracle_uncompress.restype = ctypes.c_int
        _lib.rans_oracle_uncompress.argtypes = [u8p, ctypes.c_uint32, u8p, ctypes.c_uint32, u32p]
    return _lib


ERRORS = {
    -1: 'Incorrect input length.',
    -2: 'Output buffer too small',
    -3: 'Unknown rANS order',
    -4: 'corrupt rANS block',
}


def _ptr(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_uint8))


def bound(n):
    return lib().rans_oracle_bound(n)


def compress(data, order):
    """RANS.compress restated; returns bytes."""
    a = np.frombuffer(bytes(data), dtype=np.uint8) if not isinstance(data, np.ndarray) else np.ascontiguousarray(data)
    n = a.size
    cap = bound(n)
    out = np.empty(cap, dtype=np.uint8)
    olen = ctypes.c_uint32(0)
    src = a if n else np.zeros(1, np.uint8)
    rc = lib().rans_oracle_compress(_ptr(src), n, int(order), _ptr(out), cap, ctypes.byref(olen))
    if rc:
        raise RuntimeError(ERRORS.get(rc, str(rc)))
    return out[:olen.value].tobytes()


def uncompress(blob):
    a = np.frombuffer(bytes(blob), dtype=np.uint8) if not isinstance(blob, np.ndarray) else np.ascontiguousarray(blob)
    if a.size == 0:
        return b''
    olen = ctypes.c_uint32(0)
    dummy = np.zeros(1, np.uint8)
    rc = lib().rans_oracle_uncompress(_ptr(a), a.size, _ptr(dummy), 0, ctypes.byref(olen))
    if rc not in (0, -2):
        raise RuntimeError(ERRORS.get(rc, str(rc)))
    out = np.empty(max(olen.value, 1), dtype=np.uint8)
    rc = lib().rans_oracle_uncompress(_ptr(a), a.size, _ptr(out), olen.value, ctypes.byref(olen))
    if rc:
        raise RuntimeError(ERRORS.get(rc, str(rc)))
    return out[:olen.value].tobytes()
