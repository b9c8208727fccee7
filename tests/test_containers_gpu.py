"""GPU parity for the CRAM v3 container layer (container part): cramgpu_containers_assemble_dev lays out whole
containers -- container header, compression header block, slice header block, core and external blocks -- as
ContainerIO.writeContainer does.  Pinned to the containers the jar itself wrote in the mini-JVM
(tests/golden/streams_jvm.json, 'container_v3') and, with rANS-coded blocks of realistic size, to the oracle's restatement."""
import base64

import numpy as np
import pytest

from tests import golden_streams as gs

pytestmark = pytest.mark.gpu


def _assemble(ctx, parts, tags=True):
    """parts: list of (hdr, grc, comp header bytes, ids, core (method, raw_len, bytes), ext [(cid, method, raw_len, bytes)])
    -> list of container bytes"""
    import torch
    from cramtools_b200 import records as R, device as dv
    blocks, cid, method, ctype, raw, bf = [], [], [], [], [], [0]
    for hdr, grc, chb, ids, core, ext in parts:
        blocks.append(np.frombuffer(core[2], np.uint8)); cid.append(0); method.append(core[0]); ctype.append(5); raw.append(core[1])
        for c, m, rl, data in ext:
            blocks.append(np.frombuffer(data, np.uint8)); cid.append(c); method.append(m); ctype.append(4); raw.append(rl)
        bf.append(len(blocks))
    d_in, off, ln = dv.pack_blocks(blocks, 'cuda:0')
    t_off = torch.from_numpy(np.ascontiguousarray(off, np.uint64).view(np.uint8).copy()).cuda()
    t_len = torch.from_numpy(np.ascontiguousarray(ln, np.uint32).view(np.uint8).copy()).cuda()
    t_raw = torch.from_numpy(np.ascontiguousarray(raw, np.uint32).view(np.uint8).copy()).cuda()
    t_sl = torch.from_numpy(R.slice_header_array([p[0] for p in parts])).cuda()
    torch.cuda.synchronize()
    out, olen, ooff = R.containers_assemble(t_sl.data_ptr(), [p[1] for p in parts], [p[2] for p in parts], bf, d_in.data_ptr(),
                                            t_off.data_ptr(), t_len.data_ptr(), t_raw.data_ptr(), cid, method, ctype, tags=tags, ctx=ctx)
    ctx.sync()
    buf = dv.dev_to_numpy(ctx, out, olen, np.uint8)
    o = dv.dev_to_numpy(ctx, ooff, len(parts) + 1, np.uint64)
    assert int(o[-1]) == olen and int(o[0]) == 0
    return [buf[int(o[i]):int(o[i + 1])].tobytes() for i in range(len(parts))]


def test_containers_match_jar(gpu_ctx):
    """the three golden containers in one call: byte for byte what ContainerIO.writeContainer wrote"""
    g, ref, conts = gs.load()
    parts, want = [], []
    for c, rd, series in conts:
        hdr, grc, chb, ids, tags, core, ext = gs.container_parts(c, series)
        parts.append((hdr, grc, chb, ids, (0, len(core), core), ext))
        want.append(base64.b64decode(c['container_v3']['bytes']))
    got = _assemble(gpu_ctx, parts)
    for c, a, b in zip(conts, got, want):
        assert a == b, c[0]['tag']


def test_containers_of_coded_blocks_against_oracle(gpu_ctx):
    """many containers with blocks from bytes to megabytes, every ITF8 / LTF8 width in the headers, an unplaced slice, an empty
    compression header, with and without the digest tags"""
    from oracle import records as orec
    rng = np.random.default_rng(11)
    parts = []
    grcs = [0, 127, 128, (1 << 14) - 1, 1 << 14, 1 << 21, 1 << 28, 1 << 35, 1 << 42, 1 << 49, (1 << 56) - 1, 1 << 56, (1 << 63) + 5]
    sizes = [0, 1, 57, 4000, 16376, 16377, 70000, 1 << 20, 1500000]
    for k, grc in enumerate(grcs):
        seq = [-1, -2, 0, 5, 300][k % 5]
        hdr = dict(sequence_id=seq, alignment_start=[-1, 0, 1, 130, 20000, 3000000, 270000000][k % 7], alignment_span=[0, 100, 20000, 1 << 22][k % 4],
                   n_records=[0, 1, 200, 10000, 1 << 22][k % 5], bases=[0, 150, 1500000, 1 << 40][k % 4], ref_md5=bytes(rng.integers(0, 256, 16, dtype=np.uint8)),
                   bd=int(rng.integers(0, 1 << 32)), sd=int(rng.integers(0, 1 << 32)))
        nex = [0, 1, 5, 28][k % 4]
        ids = sorted(int(v) for v in rng.choice(200000, size=nex, replace=False))
        core = bytes(rng.integers(0, 256, sizes[k % len(sizes)], dtype=np.uint8))
        ext = []
        for j, cid in enumerate(ids):
            data = bytes(rng.integers(0, 256, sizes[(k + j + 1) % len(sizes)], dtype=np.uint8))
            ext.append((cid, int(rng.integers(0, 5)), len(data) * 3 + 2, data))
        chb = bytes(rng.integers(0, 256, [0, 170, 5000][k % 3], dtype=np.uint8))
        parts.append((hdr, grc, chb, ids, (0, len(core), core), ext))
    s1, s5 = orec.SHA1_EMPTY, orec.SHA512_EMPTY
    for with_tags in (True, False):
        got = _assemble(gpu_ctx, parts, tags=with_tags)
        for k, (hdr, grc, chb, ids, core, ext) in enumerate(parts):
            tags = [('B1', s1), ('S1', s1), ('B5', s5), ('S5', s5), ('BD', hdr['bd'].to_bytes(4, 'little')), ('SD', hdr['sd'].to_bytes(4, 'little'))]
            want = orec.container_v3(hdr, grc, chb, ids, hdr['ref_md5'], tags if with_tags else None, core, ext)
            assert got[k] == want, (k, with_tags)


def test_container_argument_errors(gpu_ctx):
    from cramtools_b200 import records as R, _lib
    import torch
    t = torch.zeros(64, dtype=torch.uint8, device='cuda:0')
    with pytest.raises(_lib.CramGpuError):
        R.containers_assemble(t.data_ptr(), [0], [b''], [0, 0], t.data_ptr(), t.data_ptr(), t.data_ptr(), t.data_ptr(), [], [], [], ctx=gpu_ctx)
