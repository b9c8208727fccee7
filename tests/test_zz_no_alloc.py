"""GPU test of cramgpu_rans0_encode_static, the entry point for hosts that may not allocate device memory while they
run (all of HBM already belongs to the host's own memory manager, or a cudaMalloc's device-wide synchronisation is
not acceptable).  It keeps its scratch in a fixed device array inside the library, so the things to check are: the
result equals what the pooled entry point produces, a second call is not disturbed by what the first one left in
the shared array, the size limit is enforced with a readable message, and a call takes no memory from the device.

Collected last on purpose (file name), so that a run stopping at the first failure has already run everything else.
"""
import ctypes

import numpy as np
import pytest

from cramtools_b200 import synth, _lib

pytestmark = pytest.mark.gpu

LIMIT = 1 << 20


def _static_encode(lib, torch, data):
    """one call through the C ABI; returns (bytes, status)"""
    n = int(data.size)
    d_in = torch.from_numpy(np.concatenate([data, np.zeros(16, np.uint8)])).cuda()      # readable padding after the block
    cap = int(lib.cramgpu_rans_bound(n))
    d_out = torch.empty(cap, dtype=torch.uint8, device='cuda')
    d_len = torch.full((1,), 123456, dtype=torch.int32, device='cuda')
    d_st = torch.full((1,), -99, dtype=torch.int32, device='cuda')
    torch.cuda.synchronize()
    rc = lib.cramgpu_rans0_encode_static(ctypes.c_void_p(d_in.data_ptr()), n, ctypes.c_void_p(d_out.data_ptr()), cap,
                                         ctypes.c_void_p(d_len.data_ptr()), ctypes.c_void_p(d_st.data_ptr()),
                                         ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, lib.cramgpu_last_error(None)
    torch.cuda.synchronize()
    ln = int(d_len.item())
    return d_out[:ln].cpu().numpy().tobytes(), int(d_st.item())


def test_static_scratch_entry_point_matches_pooled_entry_point(gpu_ctx):
    import torch
    from cramtools_b200 import ExternalCompression as EC
    lib = _lib.load()
    # two calls in a row: a large block first, then a much smaller one with a different alphabet, so that anything the
    # first call left behind in the shared array (histogram, table, payload) would corrupt the second result
    first = synth.qs_stream(900000, 41)
    second = synth.rn_stream(70001, 42)
    want = EC.rans_batch([first, second], [0, 0], gpu_ctx)          # existing entry point (context-owned scratch)
    got1, st1 = _static_encode(lib, torch, first)
    free_before = torch.cuda.mem_get_info()[0]
    got2, st2 = _static_encode(lib, torch, second)
    torch.cuda.synchronize()
    # torch's caching allocator serves the test's own small tensors from memory it already holds; the entry point itself
    # must not have taken anything from the device
    assert torch.cuda.mem_get_info()[0] >= free_before - (64 << 20)
    assert st1 == 0 and st2 == 0
    assert got1 == want[0]            # integer codec: byte-identical, no tolerance
    assert got2 == want[1]
    # and once more the first input: the array now holds the second call's leftovers
    got3, st3 = _static_encode(lib, torch, first)
    assert st3 == 0 and got3 == want[0]


def test_static_scratch_entry_point_against_the_jar_and_the_oracle(rans_golden):
    """the bytes themselves, not another GPU path: every order-0 vector the reference jar's bytecode produced
    (tests/golden/rans_jvm.json, RANS.compress -> E04.compress) and the oracle on synthetic series"""
    import torch
    from oracle import rans as orans
    lib = _lib.load()
    vec, _ = rans_golden
    seen = 0
    for name, order, data, want in vec:
        if int(order) != 0:
            continue
        got, st = _static_encode(lib, torch, np.frombuffer(data, np.uint8))
        assert st == 0 and got == want, name
        seen += 1
    assert seen >= 20
    for data in (synth.qs_stream(300001, 7), synth.ba_stream(65536, 8), synth.itf8_stream(12345, 9)):
        got, st = _static_encode(lib, torch, data)
        assert st == 0 and got == orans.compress(data, 0)


def test_static_scratch_entry_point_edges(gpu_ctx):
    import torch
    from cramtools_b200 import ExternalCompression as EC
    lib = _lib.load()
    for data in (np.zeros(0, np.uint8), np.frombuffer(b'A', np.uint8), np.frombuffer(b'ACG', np.uint8),
                 synth.ba_stream(4097, 5), synth.itf8_stream(LIMIT, 6)):
        got, st = _static_encode(lib, torch, data)
        assert st == 0
        assert got == EC.rans(data, 0, gpu_ctx), data.size
    # one byte above the limit is refused, with a message that says why
    big = torch.zeros(LIMIT + 32, dtype=torch.uint8, device='cuda')
    out = torch.empty(int(lib.cramgpu_rans_bound(LIMIT + 1)), dtype=torch.uint8, device='cuda')
    ln = torch.zeros(1, dtype=torch.int32, device='cuda'); st = torch.zeros(1, dtype=torch.int32, device='cuda')
    rc = lib.cramgpu_rans0_encode_static(ctypes.c_void_p(big.data_ptr()), LIMIT + 1, ctypes.c_void_p(out.data_ptr()), out.numel(),
                                         ctypes.c_void_p(ln.data_ptr()), ctypes.c_void_p(st.data_ptr()), None)
    assert rc == _lib.ERR_ARG
    msg = lib.cramgpu_last_error(None).decode()
    assert 'limit' in msg and str(LIMIT) in msg
