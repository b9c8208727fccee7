"""GPU parity for the CRAM v3 block layer (block part): cramgpu_blocks_frame_dev wraps device-resident compressed
blocks as Block.write(3, ...) does -- header, data, CRC32.  Pinned to the blocks the jar itself wrote in the mini-JVM
(tests/golden/streams_jvm.json, 'blocks_v3') and, at sizes up to megabytes, to the oracle's restatement (zlib.crc32)."""
import base64
import ctypes as C

import numpy as np
import pytest

from tests import golden_streams as gs

pytestmark = pytest.mark.gpu


def _frame(ctx, d_comp, off, ln, raw_len, cid, method, ctype):
    """-> list of framed blocks (bytes)"""
    import torch
    from cramtools_b200 import records as R, device as dv       # noqa: F401  (records registers the signature)
    n = len(ln)
    t_off = torch.from_numpy(np.ascontiguousarray(off, np.uint64).view(np.uint8).copy()).cuda()
    t_len = torch.from_numpy(np.ascontiguousarray(ln, np.uint32).view(np.uint8).copy()).cuda()
    t_raw = torch.from_numpy(np.ascontiguousarray(raw_len, np.uint32).view(np.uint8).copy()).cuda()
    cid = np.ascontiguousarray(cid, np.int32); method = np.ascontiguousarray(method, np.uint8); ctype = np.ascontiguousarray(ctype, np.uint8)
    out, olen, ooff = C.c_void_p(), C.c_uint64(), C.c_void_p()
    torch.cuda.synchronize()
    ctx.check(ctx.lib.cramgpu_blocks_frame_dev(ctx.handle, n, C.c_void_p(d_comp.data_ptr()), C.c_void_p(t_off.data_ptr()), C.c_void_p(t_len.data_ptr()),
                                               C.c_void_p(t_raw.data_ptr()), cid.ctypes.data_as(C.POINTER(C.c_int32)),
                                               method.ctypes.data_as(C.POINTER(C.c_uint8)), ctype.ctypes.data_as(C.POINTER(C.c_uint8)),
                                               C.byref(out), C.byref(olen), C.byref(ooff), None))
    ctx.sync()
    buf = dv.dev_to_numpy(ctx, out.value, olen.value, np.uint8)
    o = dv.dev_to_numpy(ctx, ooff.value, n + 1, np.uint64)
    hdr = 2 + np.array([len(_itf8(int(c))) + len(_itf8(int(l))) + len(_itf8(int(r))) for c, l, r in zip(cid, ln, raw_len)])
    assert (o[:-1] % 16 == 0).all()
    return [buf[int(o[i]):int(o[i]) + int(hdr[i]) + int(ln[i]) + 4].tobytes() for i in range(n)]


def _itf8(v):
    from oracle import records as orec
    return orec.itf8(v)


def test_framed_rans_blocks_match_jar(gpu_ctx):
    """series -> rANS on the device -> framed blocks, byte for byte what the jar's ExternalCompression.rans + Block.write wrote"""
    from cramtools_b200 import device as dv
    g, ref, conts = gs.load()
    for c, rd, series in conts:
        cids = sorted(int(k) for k in c['blocks_v3'])
        raws = [series[k] for k in cids]
        orders = np.array([c['blocks_v3'][str(k)]['order'] for k in cids], np.uint8)
        d_in, off, ln = dv.pack_blocks(raws, 'cuda:0')
        r = dv.rans_encode_dev(gpu_ctx, d_in, off, ln, orders)
        gpu_ctx.sync()
        import torch
        torch.cuda.synchronize()
        c_off = r.off.cpu().numpy().view(np.uint64)[:len(cids)]; c_len = r.len.cpu().numpy().view(np.uint32)
        got = _frame(gpu_ctx, r.out, c_off, c_len, ln, cids, [4] * len(cids), [4] * len(cids))
        for k, b in zip(cids, got):
            assert b == base64.b64decode(c['blocks_v3'][str(k)]['block']), (c['tag'], k)


def test_framing_sizes_and_crc_against_oracle(gpu_ctx):
    """one byte to megabytes, chunk and piece boundaries, every header width: the CRC combine is exact"""
    from cramtools_b200 import device as dv
    from oracle import records as orec
    rng = np.random.default_rng(3)
    sizes = [1, 2, 55, 56, 57, 58, 59, 60, 63, 64, 65, 127, 128, 129, 4096, 16370, 16375, 16376, 16377, 16378, 16384, 16385, 32760, 32768,
             100000, 1 << 20, 1500000, 3 * (1 << 20) + 7]
    blocks = [rng.integers(0, 256, size=s, dtype=np.uint8) for s in sizes]
    cid = [0, 1, 127, 128, 16383, 16384, 2097151, 2097152, 1 << 27, (1 << 28) - 1, 1 << 28, 0x7FFFFFFF] + list(range(len(sizes) - 12))
    raw = [int(s) * 3 + 1 for s in sizes]
    method = [i % 5 for i in range(len(sizes))]; ctype = [4 + (i & 1) for i in range(len(sizes))]
    d_in, off, ln = dv.pack_blocks(blocks, 'cuda:0')
    got = _frame(gpu_ctx, d_in, off, ln, raw, cid, method, ctype)
    for i, b in enumerate(blocks):
        want = orec.cram_block_v3(method[i], ctype[i], cid[i], raw[i], b.tobytes())
        assert got[i] == want, (i, sizes[i])
