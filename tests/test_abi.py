"""CPU: the C-ABI library loads and exports every symbol include/cramgpu.h declares (no compute calls)."""
import ctypes
import os
import re

import pytest

from cramtools_b200 import _lib
from cramtools_b200 import records, sharding  # noqa: F401  (they register their signatures)

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    names = set()
    for fn in os.listdir(os.path.join(ROOT, 'include')):
        if fn.endswith('.h'):
            text = open(os.path.join(ROOT, 'include', fn)).read()
            text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
            names.update(re.findall(r'\b(cramgpu_[a-z0-9_]+)\s*\(', text))
    return sorted(names)


def test_library_built_and_loads():
    assert os.path.exists(_lib.LIB_PATH), 'run __graft_entry__.build()'
    _lib.load()


def test_every_declared_symbol_is_exported_and_bound():
    lib = ctypes.CDLL(_lib.LIB_PATH)
    syms = declared_symbols()
    assert len(syms) >= 10
    for s in syms:
        assert hasattr(lib, s), 'missing export ' + s
        assert s in _lib.SIGNATURES, 'python binding missing for ' + s
    for s in _lib.SIGNATURES:
        assert s in syms, 'binding for undeclared symbol ' + s


def test_bound_matches_reference_formula():
    lib = _lib.load()
    for n in (0, 1, 4, 1000, 1 << 20, 8 << 20):
        assert lib.cramgpu_rans_bound(n) == int(1.05 * n + 257 * 257 * 3 + 4)   # [RANS.java:79]


def test_raw_sizes_is_host_only(rans_golden):
    import numpy as np
    vec, err = rans_golden
    lib = _lib.load()
    blobs = [v[3] for v in vec]
    arrs = [np.frombuffer(b, dtype=np.uint8) for b in blobs]
    n = len(arrs)
    pa = (_lib.u8p * n)()
    for i, a in enumerate(arrs):
        pa[i] = a.ctypes.data_as(_lib.u8p) if a.size else ctypes.cast(None, _lib.u8p)
    ln = np.array([a.size for a in arrs], dtype=np.uint32)
    raw = np.zeros(n, dtype=np.uint32)
    st = np.zeros(n, dtype=np.int32)
    rc = lib.cramgpu_rans_raw_sizes(n, pa, ln.ctypes.data_as(_lib.u32p), raw.ctypes.data_as(_lib.u32p),
                                    st.ctypes.data_as(_lib.i32p))
    assert rc == 0
    assert [int(r) for r in raw] == [len(v[2]) for v in vec]
    for name, blob, throws in err:
        a = np.frombuffer(blob, dtype=np.uint8)
        pa1 = (_lib.u8p * 1)(a.ctypes.data_as(_lib.u8p))
        ln1 = np.array([a.size], dtype=np.uint32)
        rc = lib.cramgpu_rans_raw_sizes(1, pa1, ln1.ctypes.data_as(_lib.u32p), raw.ctypes.data_as(_lib.u32p),
                                        st.ctypes.data_as(_lib.i32p))
        if 'Incorrect input length' in throws:
            assert rc == _lib.ERR_INPUT_LENGTH, name
        else:
            assert rc == _lib.ERR_UNKNOWN_ORDER, name


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip('a GPU is present')
    with pytest.raises(_lib.CramGpuError):
        _lib.Context(0)
