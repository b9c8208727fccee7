"""ctypes binding of libcramgpu.so (include/cramgpu.h).

There is no CPU fallback anywhere in this package: if the shared library has not been built
(`python -c "import __graft_entry__ as g; g.build()"` or `make -C cramtools_b200/csrc`) or no CUDA
device is usable, calls fail loudly.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, 'libcramgpu.so')

OK = 0
ERR_INPUT_LENGTH = -1
ERR_OUT_TOO_SMALL = -2
ERR_UNKNOWN_ORDER = -3
ERR_CORRUPT = -4
ERR_CUDA = -5
ERR_ARG = -6
ERR_UNSUPPORTED = -7

u8p = ctypes.POINTER(ctypes.c_uint8)
u8pp = ctypes.POINTER(u8p)
u32p = ctypes.POINTER(ctypes.c_uint32)
i32p = ctypes.POINTER(ctypes.c_int32)
u64p = ctypes.POINTER(ctypes.c_uint64)
vp = ctypes.c_void_p

# name -> (restype, argtypes); every symbol include/cramgpu.h declares
SIGNATURES = {
    'cramgpu_init': (ctypes.c_int, [ctypes.c_int, ctypes.POINTER(vp)]),
    'cramgpu_destroy': (None, [vp]),
    'cramgpu_last_error': (ctypes.c_char_p, [vp]),
    'cramgpu_sync': (ctypes.c_int, [vp]),
    'cramgpu_launch_count': (ctypes.c_uint64, [vp]),
    'cramgpu_set_option': (ctypes.c_int, [vp, ctypes.c_char_p, ctypes.c_uint64]),
    'cramgpu_copy_to_host': (ctypes.c_int, [vp, vp, vp, ctypes.c_uint64, vp]),
    'cramgpu_profile_enable': (ctypes.c_int, [vp, ctypes.c_int]),
    'cramgpu_profile_read': (ctypes.c_int, [vp, ctypes.c_uint32, ctypes.POINTER(ctypes.c_char_p),
                                            ctypes.POINTER(ctypes.c_float)]),
    'cramgpu_rans_bound': (ctypes.c_uint32, [ctypes.c_uint32]),
    'cramgpu_rans_encode_batch': (ctypes.c_int, [vp, ctypes.c_uint32, u8pp, u32p, u8p, u8pp, u32p, u32p, i32p]),
    'cramgpu_rans_raw_sizes': (ctypes.c_int, [ctypes.c_uint32, u8pp, u32p, u32p, i32p]),
    'cramgpu_rans_decode_batch': (ctypes.c_int, [vp, ctypes.c_uint32, u8pp, u32p, u8pp, u32p, u32p, i32p]),
    'cramgpu_rans_encode_dev': (ctypes.c_int, [vp, ctypes.c_uint32, vp, u64p, u32p, u8p, vp, ctypes.c_uint64,
                                               vp, vp, vp, vp]),
    'cramgpu_rans0_encode_static': (ctypes.c_int, [vp, ctypes.c_uint32, vp, ctypes.c_uint32, vp, vp, vp]),
    'cramgpu_rans_decode_dev': (ctypes.c_int, [vp, ctypes.c_uint32, vp, u64p, u32p, u8p, vp, ctypes.c_uint64,
                                               vp, vp, vp, vp]),
}

_lib = None


def load():
    """dlopen libcramgpu.so and type every export.  Raises if the library is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError('%s not built: run __graft_entry__.build() (nvcc, sm_100a). '
                               'cramtools_b200 has no CPU fallback.' % LIB_PATH)
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


class CramGpuError(RuntimeError):
    """Raised where the reference throws RuntimeException; .status holds the cramgpu_status."""

    def __init__(self, status, message):
        RuntimeError.__init__(self, message)
        self.status = status


class Context:
    """One cramgpu_ctx (CUDA stream + scratch + pinned staging).  One per calling thread."""

    def __init__(self, device=0):
        self.lib = load()
        h = vp()
        rc = self.lib.cramgpu_init(int(device), ctypes.byref(h))
        if rc != OK:
            raise CramGpuError(rc, self.lib.cramgpu_last_error(None).decode())
        self.handle = h
        self.device = int(device)

    def last_error(self):
        return self.lib.cramgpu_last_error(self.handle).decode()

    def check(self, rc):
        if rc != OK:
            raise CramGpuError(rc, self.last_error())

    def sync(self):
        self.check(self.lib.cramgpu_sync(self.handle))

    def set_option(self, name, value):
        self.check(self.lib.cramgpu_set_option(self.handle, name.encode(), int(value)))

    def launch_count(self):
        return int(self.lib.cramgpu_launch_count(self.handle))

    def profile_enable(self, on=True):
        self.check(self.lib.cramgpu_profile_enable(self.handle, 1 if on else 0))

    def profile_read(self, cap=65536):
        """[(kernel name, milliseconds)] since the last read; synchronise the stream first"""
        names = (ctypes.c_char_p * cap)()
        ms = (ctypes.c_float * cap)()
        n = self.lib.cramgpu_profile_read(self.handle, cap, names, ms)
        n = min(max(n, 0), cap)
        return [(names[i].decode(), float(ms[i])) for i in range(n)]

    def close(self):
        if self.handle:
            self.lib.cramgpu_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_default_ctx = {}


def default_context(device=0):
    ctx = _default_ctx.get(device)
    if ctx is None:
        ctx = _default_ctx[device] = Context(device)
    return ctx
