"""GPU parity for SURVEY 8f-2 (CIGAR rebuild + BAM record serialisation + BAM-native record input), through the C ABI with
device-resident buffers.  Pinned to the reference jar: tests/golden/streams_jvm.json holds, for every golden record, the CIGAR
Cram2SamRecordFactory.getCigar2 rebuilt from the features the jar's reader read back, and the BAM record cramtools'
BAMRecordView laid out (both produced in the mini-JVM, oracle/make_golden_streams.py)."""
import base64
import ctypes as C

import numpy as np
import pytest

from tests import golden_streams as gs

pytestmark = pytest.mark.gpu
OPS = 'MIDNSHP=X'


@pytest.fixture(scope='module')
def R():
    from cramtools_b200 import records
    return records


@pytest.fixture(scope='module')
def G():
    return gs.load()


def _dev(a, dt):
    import torch
    a = np.ascontiguousarray(a, dt)
    if a.size == 0:
        a = np.zeros(1, dt)
    return torch.from_numpy(a.view(np.uint8).copy()).cuda()


def _host(ptr, n, dt):
    """n items of dt at device address ptr -> numpy"""
    import cramtools_b200
    from cramtools_b200 import device as dv
    import torch
    torch.cuda.synchronize()
    return dv.dev_to_numpy(cramtools_b200.default_context(0), int(ptr), int(n), dt) if int(n) else np.zeros(0, dt)


def _vp(t):
    return C.c_void_p(t.data_ptr())


def _cigar_rebuild(R, ctx, flag, read_off, feat_off, f_op, f_pos, f_val, f_aux):
    import torch
    t = [_dev(flag, np.uint16), _dev(read_off, np.uint64), _dev(feat_off, np.uint64), _dev(f_op, np.uint8), _dev(f_pos, np.int32),
         _dev(f_val, np.int32), _dev(f_aux, np.int32)]
    cig, off, total = C.c_void_p(), C.c_void_p(), C.c_uint64()
    ctx.check(ctx.lib.cramgpu_cigar_rebuild_dev(ctx.handle, len(flag), *[_vp(x) for x in t], C.byref(cig), C.byref(off), C.byref(total), None))
    ctx.sync()
    torch.cuda.synchronize()
    n = len(flag)
    return _host(cig.value, total.value, np.uint32), _host(off.value, n + 1, np.uint64)


def test_cigar_rebuild_matches_jar(R, G, gpu_ctx):
    g, ref, conts = G
    for c, rd, series in conts:
        enc = R.features_encode(rd, ref, None, ctx=gpu_ctx)
        words, off = _cigar_rebuild(R, gpu_ctx, rd['flag'], rd['read_off'], enc.feat_off, enc.f_op, enc.f_pos, enc.f_val, enc.f_aux)
        for i, rb in enumerate(c['read_back']):
            got = ''.join('%d%s' % (w >> 4, OPS[w & 15]) for w in words[int(off[i]):int(off[i + 1])])
            want = '' if rd['flag'][i] & 4 else rb['cigar2']
            assert got == want, (c['tag'], i, got, want)


def _records_struct(R, rd, sc, cig_words, cig_off, tags, keep):
    n = len(rd['pos'])
    paired = (rd['flag'] & 1) != 0
    names = [x.encode('latin-1') if isinstance(x, str) else bytes(x) for x in rd['names']]
    tag_off = np.concatenate([[0], np.cumsum([len(t) for t in tags])]).astype(np.uint64)
    arrs = dict(ref_id=(rd['ref_id'], np.int32), pos=(rd['pos'], np.int32),
                mate_ref_id=(np.where(paired, [s['mate_seq_id'] for s in sc], -1), np.int32),
                mate_pos=(np.where(paired, [s['mate_start'] for s in sc], 0), np.int32), tlen=([s['tlen'] for s in sc], np.int32),
                flag=(rd['flag'], np.uint16), mapq=(rd['mapq'], np.uint8), rflags=(rd['rflags'], np.uint8),
                names=(np.frombuffer(b''.join(names), np.uint8), np.uint8),
                name_off=(np.concatenate([[0], np.cumsum([len(x) for x in names])]), np.uint64), cigar=(cig_words, np.uint32),
                cigar_off=(cig_off, np.uint64), bases=(rd['bases'], np.uint8), quals=(rd['quals'], np.uint8), read_off=(rd['read_off'], np.uint64),
                tags=(np.frombuffer(b''.join(tags), np.uint8), np.uint8), tag_off=(tag_off, np.uint64))
    B = R.BamRecords()
    B.n_reads = n
    for k, (v, dt) in arrs.items():
        t = _dev(v, dt)
        keep.append(t)
        setattr(B, k, t.data_ptr())
    return B


def test_bam_pack_matches_jar_and_unpack_inverts_it(R, G, gpu_ctx):
    """features -> CIGAR -> BAM records: byte for byte the records BAMRecordView laid out in the mini-JVM; then the stripe goes
    back through cramgpu_bam_unpack_dev and every field returns (the rebuilt CIGAR, 1-based positions, upper-case bases)"""
    import torch
    g, ref, conts = G
    ctx = gpu_ctx
    for c, rd, series in conts:
        n = len(rd['pos'])
        enc = R.features_encode(rd, ref, None, ctx=ctx)
        words, off = _cigar_rebuild(R, ctx, rd['flag'], rd['read_off'], enc.feat_off, enc.f_op, enc.f_pos, enc.f_val, enc.f_aux)
        tags = [base64.b64decode(rb['bam_tags']) for rb in c['read_back']]
        keep = []
        B = _records_struct(R, rd, c['scalars'], words, off, tags, keep)
        st = torch.full((n,), -9, dtype=torch.int32, device='cuda')
        bam, blen, roff = C.c_void_p(), C.c_uint64(), C.c_void_p()
        ctx.check(ctx.lib.cramgpu_bam_pack_dev(ctx.handle, C.byref(B), C.byref(bam), C.byref(blen), C.byref(roff), _vp(st), None))
        ctx.sync()
        stripe = _host(bam.value, blen.value, np.uint8).tobytes()
        rec_off = _host(roff.value, n + 1, np.uint64)
        assert (st.cpu().numpy() == 0).all()
        want = [base64.b64decode(rb['bam']) for rb in c['read_back']]
        for i in range(n):
            assert stripe[int(rec_off[i]):int(rec_off[i + 1])] == want[i], (c['tag'], i)
        assert stripe == b''.join(want)
        # ---- and back
        d_stripe = _dev(np.frombuffer(stripe, np.uint8), np.uint8); d_off = _dev(rec_off, np.uint64)
        U = R.BamRecords()
        st2 = torch.full((n,), -9, dtype=torch.int32, device='cuda')
        ctx.check(ctx.lib.cramgpu_bam_unpack_dev(ctx.handle, _vp(d_stripe), _vp(d_off), n, C.byref(U), _vp(st2), None))
        ctx.sync()
        assert (st2.cpu().numpy() == 0).all()
        assert (_host(U.pos, n, np.int32) == rd['pos']).all() and (_host(U.flag, n, np.uint16) == rd['flag']).all()
        assert (_host(U.ref_id, n, np.int32) == rd['ref_id']).all() and (_host(U.mapq, n, np.uint8) == rd['mapq']).all()
        assert (_host(U.bin, n, np.uint16) == [rb['bam_bin'] for rb in c['read_back']]).all()
        assert (_host(U.tlen, n, np.int32) == [s['tlen'] for s in c['scalars']]).all()
        assert (_host(U.read_off, n + 1, np.uint64) == rd['read_off']).all() and (_host(U.cigar_off, n + 1, np.uint64) == off).all()
        assert (_host(U.cigar, U.cigar_len, np.uint32) == words).all()
        assert _host(U.bases, U.bases_len, np.uint8).tobytes() == rd['bases'].tobytes().upper().replace(b'.', b'N')
        q = _host(U.quals, U.bases_len, np.uint8)
        rfl = _host(U.rflags, n, np.uint8)
        for i in range(n):
            a, b = int(rd['read_off'][i]), int(rd['read_off'][i + 1])
            if rd['rflags'][i] & 1:
                assert (rfl[i] & 1) and (q[a:b] == 0xFF).all()
            else:
                assert not (rfl[i] & 1) and (q[a:b] == rd['quals'][a:b]).all()
        assert _host(U.names, U.names_len, np.uint8).tobytes() == b''.join(x.encode('latin-1') for x in rd['names'])
        assert _host(U.tags, U.tags_len, np.uint8).tobytes() == b''.join(tags)


def test_bam_round_trip_parity_with_oracle_large(R, gpu_ctx):
    """20 000 simulated records with indels, clips and ragged lengths: the stripe equals the oracle's records one by one; a
    corrupt record (block_size off) and a base outside the BAM alphabet are reported per record"""
    import torch
    from oracle import records as orec
    from cramtools_b200 import synth
    from tests.test_records_gpu import _concat_reads
    ctx = gpu_ctx
    ref = synth.reference(50000, 8)
    rd = _concat_reads([synth.reads(ref, 12000, 9, read_len=150, indel_frac=0.3, clip_frac=0.2, sub_rate=0.01),
                        synth.reads(ref, 5000, 10, read_len=101, indel_frac=0.2), synth.reads(ref, 3000, 11, read_len=36, clip_frac=0.3)])
    n = len(rd['pos'])
    rd['ref_id'] = np.zeros(n, np.int32); rd['rflags'] = np.zeros(n, np.uint8); rd['rflags'][::53] = 1
    enc = R.features_encode(rd, ref, None, ctx=ctx)
    words, off = _cigar_rebuild(R, ctx, rd['flag'], rd['read_off'], enc.feat_off, enc.f_op, enc.f_pos, enc.f_val, enc.f_aux)
    # the rebuilt CIGAR against the oracle's getCigar2 and (M/I/D/S reads) against the original CIGAR
    ex = orec.extract(rd, ref)
    for i in range(0, n, 7):
        feats = []
        for f in range(int(ex.feat_off[i]), int(ex.feat_off[i + 1])):
            op = chr(ex.f_op[f])
            feats.append((op, int(ex.f_pos[f]), int(ex.f_aux[f]) if op in 'SI' else int(ex.f_val[f])))
        want = orec.cigar_from_features(feats, int(rd['read_off'][i + 1] - rd['read_off'][i]))
        got = [(int(w >> 4), OPS[w & 15]) for w in words[int(off[i]):int(off[i + 1])]]
        assert got == want, i
        orig = [(int(w >> 4), OPS[w & 15]) for w in rd['cigar'][int(rd['cigar_off'][i]):int(rd['cigar_off'][i + 1])]]
        assert got == orig, (i, got, orig)
    sc = [dict(mate_seq_id=-1, mate_start=0, tlen=0)] * n
    tags = [b'RGZgrp%d\0' % (i % 3) if i % 4 == 0 else b'' for i in range(n)]
    keep = []
    B = _records_struct(R, dict(rd, names=rd['names']), sc, words, off, tags, keep)
    st = torch.full((n,), -9, dtype=torch.int32, device='cuda')
    bam, blen, roff = C.c_void_p(), C.c_uint64(), C.c_void_p()
    ctx.check(ctx.lib.cramgpu_bam_pack_dev(ctx.handle, C.byref(B), C.byref(bam), C.byref(blen), C.byref(roff), _vp(st), None))
    ctx.sync()
    stripe = bytearray(_host(bam.value, blen.value, np.uint8).tobytes())
    rec_off = _host(roff.value, n + 1, np.uint64)
    assert (st.cpu().numpy() == 0).all()
    for i in range(0, n, 11):
        a, b = int(rd['read_off'][i]), int(rd['read_off'][i + 1])
        cg = [(int(w >> 4), OPS[w & 15]) for w in words[int(off[i]):int(off[i + 1])]]
        want = orec.bam_record(0, int(rd['pos'][i]), bytes(rd['names'][i]) if not isinstance(rd['names'][i], str) else rd['names'][i].encode(),
                               int(rd['mapq'][i]), int(rd['flag'][i]), cg, -1, 0, 0, rd['bases'][a:b].tobytes(),
                               b'' if rd['rflags'][i] & 1 else rd['quals'][a:b].tobytes(), tags[i])
        assert bytes(stripe[int(rec_off[i]):int(rec_off[i + 1])]) == want, i
    # ---- unpack, with record 5 corrupted (block_size one too large)
    o5 = int(rec_off[5])
    stripe[o5] = (stripe[o5] + 1) & 0xFF
    d_stripe = _dev(np.frombuffer(bytes(stripe), np.uint8), np.uint8); d_off = _dev(rec_off, np.uint64)
    U = R.BamRecords()
    st2 = torch.full((n,), -9, dtype=torch.int32, device='cuda')
    ctx.check(ctx.lib.cramgpu_bam_unpack_dev(ctx.handle, _vp(d_stripe), _vp(d_off), n, C.byref(U), _vp(st2), None))
    ctx.sync()
    s2 = st2.cpu().numpy()
    assert s2[5] == -4 and (np.delete(s2, 5) == 0).all()
    uro = _host(U.read_off, n + 1, np.uint64)
    assert uro[6] == uro[5] and int(uro[n]) == int(rd['read_off'][n]) - int(rd['read_off'][6] - rd['read_off'][5])
    ub = _host(U.bases, U.bases_len, np.uint8)
    assert ub[:int(uro[5])].tobytes() == rd['bases'][:int(rd['read_off'][5])].tobytes()
    assert ub[int(uro[6]):].tobytes() == rd['bases'][int(rd['read_off'][6]):].tobytes()
    assert (np.delete(_host(U.pos, n, np.int32), 5) == np.delete(rd['pos'], 5)).all()
    # ---- a base BAM cannot hold
    bad = rd['bases'].copy(); bad[int(rd['read_off'][9]) + 3] = ord('J')
    keep2 = []
    B2 = _records_struct(R, dict(rd, bases=bad), sc, words, off, tags, keep2)
    ctx.check(ctx.lib.cramgpu_bam_pack_dev(ctx.handle, C.byref(B2), C.byref(bam), C.byref(blen), C.byref(roff), _vp(st), None))
    ctx.sync()
    s3 = st.cpu().numpy()
    assert s3[9] == -6 and (np.delete(s3, 9) == 0).all()
    with pytest.raises(ValueError):
        orec.bam_record(0, 1, b'x', 0, 0, [(4, 'M')], -1, 0, 0, b'ACJT', b'####')
