"""GPU parity tests for the record-feature work (rows F1-F5, Q, W, N1, N2), through the C ABI."""
import base64

import numpy as np
import pytest

from cramtools_b200 import synth, _lib
from tests import golden_records as gr

pytestmark = pytest.mark.gpu

SPECS = ['*8', 'm5', 'M30', '*40', '*0', 'N40-D8', 'R8-N40', 'I40', 'N8', 'D40-I8', 'm20_8-N40', 'U']


@pytest.fixture(scope='module')
def R():
    from cramtools_b200 import records
    return records


@pytest.fixture(scope='module')
def G():
    return gr.load()


def _drop_code(rows):
    return [[f[:4] if f[0] == 'X' else f for f in (row or [])] for row in rows]


def test_golden_features_and_matrix(R, G, gpu_ctx):
    g, ref, rd = G
    n = len(rd['pos'])
    enc = R.features_encode(rd, ref, None, ctx=gpu_ctx)
    got = gr.features_as_lists(enc.f_op, enc.f_pos, enc.f_val, enc.f_aux, enc.feat_off, enc.payload, n)
    for i in range(n):
        assert got[i] == (g['features_coded'][i] or []), (i, rd['names'][i])
    assert enc.matrix[0, :5].tobytes() == base64.b64decode(g['subst_matrix'])
    # lossless: force iff the record has quality scores (CramSerilization.java:160-161)
    for i in range(n):
        assert (enc.force_qs[i] & 1) == (0 if rd['rflags'][i] & 1 else 1)


@pytest.mark.parametrize('spec', SPECS)
def test_golden_lossy_models(R, G, gpu_ctx, spec):
    g, ref, rd = G
    n = len(rd['pos'])
    enc = R.features_encode(rd, ref, spec, ctx=gpu_ctx)
    got = _drop_code(gr.features_as_lists(enc.f_op, enc.f_pos, enc.f_val, enc.f_aux, enc.feat_off, enc.payload, n))
    rows = g['lossy'][spec]
    for i, row in enumerate(rows):
        if 'throws' in row:
            continue
        assert (enc.force_qs[i] & 1) == row['force'], (spec, i, rd['names'][i])
        assert got[i] == _drop_code([row['features']])[0], (spec, i, rd['names'][i])
        if row['force'] and row['scores'] not in (None, '*'):
            o = int(enc.read_qs_off[i]); ln = int(rd['read_off'][i + 1] - rd['read_off'][i])
            assert enc.streams[o:o + ln].tobytes() == base64.b64decode(row['scores']), (spec, i)


def test_lossy_spec_parser(R):
    # src/test/java/htsjdk/samtools/cram/lossy/TestQualityScorePreservation.java:46-114 (host-only, but lives with its users)
    p = R.QualityScorePreservation.lossyFromSpec('m999_8').getPreservationPolicies()
    assert len(p) == 1 and p[0].read_cat == 3 and p[0].read_param == 999 and p[0].treatment == 1 and p[0].treat_param == 8
    p = R.QualityScorePreservation.lossyFromSpec('R8-N40').getPreservationPolicies()
    assert [x.treatment for x in p] == [1, 0] and p[0].base_cat[0] == 1 and p[1].base_cat[0] == 2
    with pytest.raises(_lib.CramGpuError):
        R.QualityScorePreservation.lossyFromSpec('Z9')
    p = R.QualityScorePreservation.lossyFromSpec('X10-P2_8').getPreservationPolicies()       # coverage / pile-up selectors (tests/test_tracks_gpu.py)
    assert [x.treatment for x in p] == [1, 0] and p[0].base_cat[0] == 6 | (2 << 8) and p[1].base_cat[0] == 5 | (10 << 8)


def _sim(n_reads, n_cont, seed, unmapped_tail=0):
    ref = synth.reference(60000, seed)
    rd = synth.reads(ref, n_reads, seed + 1, sub_rate=0.004, indel_frac=0.1, clip_frac=0.05, unmapped_tail=unmapped_tail)
    n = len(rd['pos'])
    cont_off = np.linspace(0, n, n_cont + 1).astype(np.uint64)
    return ref, rd, cont_off


@pytest.mark.parametrize('spec', [None, '*8', 'm5', 'N40-D8', 'R8-N40', 'D8-I40-m30'])
def test_parity_with_oracle_multi_container(R, gpu_ctx, spec):
    from oracle import records as orec
    ref, rd, cont_off = _sim(6000, 5, 11, unmapped_tail=150)
    enc = R.features_encode(rd, ref, spec, cont_off=cont_off, ctx=gpu_ctx)
    n = len(rd['pos'])
    foff = 0
    for c in range(5):
        a, b = int(cont_off[c]), int(cont_off[c + 1])
        sub = dict(bases=rd['bases'][rd['read_off'][a]:rd['read_off'][b]], quals=rd['quals'][rd['read_off'][a]:rd['read_off'][b]],
                   read_off=rd['read_off'][a:b + 1] - rd['read_off'][a], cigar=rd['cigar'][rd['cigar_off'][a]:rd['cigar_off'][b]],
                   cigar_off=rd['cigar_off'][a:b + 1] - rd['cigar_off'][a], pos=rd['pos'][a:b], flag=rd['flag'][a:b], mapq=rd['mapq'][a:b])
        ex = orec.extract(sub, ref)
        force, scores = orec.quality(sub, ref, spec)
        want = orec.write_streams(sub, ex, force, scores)
        assert enc.matrix[c, :5].tobytes() == ex.matrix, c
        assert (enc.force_qs[a:b] == force).all(), c
        for name in R.SERIES:
            assert enc.series(c, name) == want[name], (spec, c, name)
        # feature arrays minus the Q features equal the oracle's extraction
        f0, f1 = int(enc.feat_off[a]), int(enc.feat_off[b])
        keep = enc.f_op[f0:f1] != ord('Q')
        assert (enc.f_op[f0:f1][keep] == ex.f_op).all()
        assert (enc.f_pos[f0:f1][keep] == ex.f_pos).all()
        isS = ex.f_op == ord('S')
        assert (enc.f_val[f0:f1][keep][~isS] == ex.f_val[~isS]).all()
        assert (enc.f_aux[f0:f1][keep] == ex.f_aux).all()
        foff = f1
    assert foff == enc.n_feat == int(enc.feat_off[n])


def test_golden_restore_bases(R, G, gpu_ctx):
    g, ref, rd = G
    n = len(rd['pos'])
    enc = R.features_encode(rd, ref, None, ctx=gpu_ctx)
    matrix = np.frombuffer(base64.b64decode(g['subst_matrix']) + b'\0\0\0', np.uint8)
    rfl = np.zeros(n, np.uint8)
    rfl[(rd['flag'] & 4) != 0] |= 2
    for name, case in g['restore_bases'].items():
        ws, we = case['window_start0'], case['window_end0']
        bases, quals, star, status = R.normalize_decode(rd['pos'], rd['read_off'], rfl, enc, ref[ws:we], ws, matrix, ctx=gpu_ctx,
                                                        check=False)
        for i, want in enumerate(case['bases']):
            if want is None:
                continue
            lo, hi = int(rd['read_off'][i]), int(rd['read_off'][i + 1])
            if isinstance(want, dict):
                assert status[i] == _lib.ERR_CORRUPT, (name, i)
            else:
                assert status[i] == 0, (name, i, rd['names'][i])
                assert bases[lo:hi].tobytes() == base64.b64decode(want), (name, i, rd['names'][i])


@pytest.mark.parametrize('spec', ['lossless'] + SPECS)
def test_golden_restore_quals(R, G, gpu_ctx, spec):
    g, ref, rd = G
    n = len(rd['pos'])
    enc = R.features_encode(rd, ref, None if spec == 'lossless' else spec, ctx=gpu_ctx)
    rfl = (enc.force_qs & 1).astype(np.uint8)
    rfl[(rd['flag'] & 4) != 0] |= 2
    bases, quals, star, status = R.normalize_decode(rd['pos'], rd['read_off'], rfl, enc, ref, 0, enc.matrix[0], qs=enc.streams,
                                                    read_qs_off=enc.read_qs_off, ba=enc.streams, read_ba_off=enc.read_ba_off,
                                                    ctx=gpu_ctx, check=False)
    for i, want in enumerate(g['restore_quals'][spec]):
        if isinstance(want, dict):
            continue
        lo, hi = int(rd['read_off'][i]), int(rd['read_off'][i + 1])
        if want == '*':
            assert star[i] == 1, (spec, i, rd['names'][i])
        else:
            assert star[i] == 0, (spec, i, rd['names'][i])
            assert quals[lo:hi].tobytes() == base64.b64decode(want), (spec, i, rd['names'][i])


def test_round_trip_lossless_multi_container(R, gpu_ctx):
    """decode(encode(reads)) gives back bases and qualities bit for bit (upper-case ACGT reads, lossless)."""
    ref, rd, cont_off = _sim(20000, 7, 23, unmapped_tail=300)
    n = len(rd['pos'])
    enc = R.features_encode(rd, ref, None, cont_off=cont_off, ctx=gpu_ctx)
    rfl = (enc.force_qs & 1).astype(np.uint8)
    rfl[(rd['flag'] & 4) != 0] |= 2
    # per container region = [min start, max end] as restoreBases builds it (:212-225)
    nc = 7
    win_off = np.zeros(nc, np.uint64); win_len = np.zeros(nc, np.uint64); ws0 = np.zeros(nc, np.int64)
    for c in range(nc):
        a, b = int(cont_off[c]), int(cont_off[c + 1])
        mp = [i for i in range(a, b) if rd['pos'][i] > 0]
        s = min(int(rd['pos'][i]) for i in mp) - 1
        e = max(int(rd['pos'][i]) + 170 for i in mp)
        win_off[c] = s; win_len[c] = min(e, ref.size) - s; ws0[c] = s
    bases, quals, star, status = R.normalize_decode(rd['pos'], rd['read_off'], rfl, enc, ref, ws0, enc.matrix, cont_off=cont_off,
                                                    win_off=win_off, win_len=win_len, qs=enc.streams, read_qs_off=enc.read_qs_off,
                                                    ba=enc.streams, read_ba_off=enc.read_ba_off, ctx=gpu_ctx)
    assert (status == 0).all() and (star == 0).all()
    assert (bases == rd['bases']).all()
    assert (quals == rd['quals']).all()


def test_empty_batch(R, gpu_ctx):
    rd = dict(bases=np.zeros(0, np.uint8), quals=np.zeros(0, np.uint8), read_off=np.zeros(1, np.uint64), cigar=np.zeros(0, np.uint32),
              cigar_off=np.zeros(1, np.uint64), pos=np.zeros(0, np.int32), flag=np.zeros(0, np.uint16), mapq=np.zeros(0, np.uint8))
    enc = R.features_encode(rd, b'ACGT', None, cont_off=[0, 0], ctx=gpu_ctx)
    assert enc.n_feat == 0


def test_host_chain_features_then_block_codec(R, gpu_ctx):
    """the call sequence of the Java host: convert() -> compress() of the QS/BA series -> uncompress() -> normalize();
    results of the first call must survive the later ones (they live in context-owned pinned memory)"""
    from cramtools_b200 import ExternalCompression as EC
    from oracle import rans as orans
    ref, rd, cont_off = _sim(4000, 2, 31)
    enc = R.features_encode(rd, ref, None, cont_off=cont_off, ctx=gpu_ctx)
    series = [enc.series(c, nm) for c in range(2) for nm in ('QS', 'BA')]
    blobs = EC.rans_batch(series, [1] * len(series), gpu_ctx)
    for s_, b in zip(series, blobs):
        assert b == orans.compress(s_, 1)
    back = EC.unrans_batch(blobs, gpu_ctx)
    assert back == series
    # the first call's results are still intact after the other calls reused the staging buffers
    enc2 = R.features_encode(rd, ref, None, cont_off=cont_off, ctx=gpu_ctx)
    assert enc2.streams.tobytes() == enc.streams.tobytes()


def test_resident_series_and_device_pointers(R, gpu_ctx):
    """The host entry points with intermediates left on the device: encode with QS/BA resident, code them from device
    pointers, decode QS into device memory, restore from it -- same bytes as the all-host path at every stage."""
    import ctypes as C
    import torch
    from cramtools_b200 import ExternalCompression, rans_bound
    from cramtools_b200._lib import u8p, u32p, i32p, u8pp
    ref, rd, cont_off = _sim(12000, 5, 41, unmapped_tail=200)
    n, nc = len(rd['pos']), 5
    host = R.features_encode(rd, ref, None, cont_off=cont_off, ctx=gpu_ctx)
    enc, dev = R.features_encode_resident(rd, ref, None, cont_off=cont_off, ctx=gpu_ctx)
    KQS, KBA = R.SERIES.index('QS'), R.SERIES.index('BA')
    assert (enc.stream_off == host.stream_off).all() and (enc.stream_len == host.stream_len).all()
    for c in range(nc):
        for name in R.SERIES:
            if name not in ('QS', 'BA'):
                assert enc.series(c, name) == host.series(c, name), (c, name)
    for name in ('f_op', 'f_pos', 'f_val', 'f_aux', 'feat_off', 'payload', 'force_qs', 'read_qs_off', 'read_ba_off'):
        assert (getattr(enc, name) == getattr(host, name)).all(), name
    # rANS of QS and BA straight from the device-resident series, plus one ordinary host block, in one call
    extra = np.frombuffer(b'host block ' * 700, np.uint8).copy()
    blocks = [(int(dev.streams) + int(enc.stream_off[c, k]), int(enc.stream_len[c, k])) for c in range(nc) for k in (KQS, KBA)]
    blocks.append((extra.ctypes.data, extra.size))
    nb = len(blocks)
    in_ptr = np.array([b[0] for b in blocks], np.uint64); in_len = np.array([b[1] for b in blocks], np.uint32)
    order = np.ones(nb, np.uint8)
    caps = np.array([rans_bound(int(x)) for x in in_len], np.uint32)
    outs = [np.zeros(int(cp), np.uint8) for cp in caps]
    out_ptr = np.array([o.ctypes.data for o in outs], np.uint64)
    out_len = np.zeros(nb, np.uint32); status = np.zeros(nb, np.int32)
    gpu_ctx.check(gpu_ctx.lib.cramgpu_rans_encode_batch(gpu_ctx.handle, nb, C.cast(in_ptr.ctypes.data, u8pp), in_len.ctypes.data_as(u32p),
                                                        order.ctypes.data_as(u8p), C.cast(out_ptr.ctypes.data, u8pp),
                                                        caps.ctypes.data_as(u32p), out_len.ctypes.data_as(u32p), status.ctypes.data_as(i32p)))
    want = [host.series(c, name) for c in range(nc) for name in ('QS', 'BA')] + [extra.tobytes()]
    for i in range(nb):
        assert outs[i][:out_len[i]].tobytes() == ExternalCompression.rans(want[i], 1, ctx=gpu_ctx), i
    # decode: QS and BA land in device memory (one buffer laid out like `streams`), the host block in host memory
    d_streams = torch.zeros(int(host.streams.size) + 64, dtype=torch.uint8, device='cuda')
    back_host = np.zeros(extra.size + 16, np.uint8)
    dst = [int(d_streams.data_ptr()) + int(enc.stream_off[c, k]) for c in range(nc) for k in (KQS, KBA)] + [back_host.ctypes.data]
    dst_ptr = np.array(dst, np.uint64)
    dcap = np.array([int(x) + 16 for x in in_len], np.uint32)
    back_len = np.zeros(nb, np.uint32)
    gpu_ctx.check(gpu_ctx.lib.cramgpu_rans_decode_batch(gpu_ctx.handle, nb, C.cast(out_ptr.ctypes.data, u8pp), out_len.ctypes.data_as(u32p),
                                                        C.cast(dst_ptr.ctypes.data, u8pp), dcap.ctypes.data_as(u32p),
                                                        back_len.ctypes.data_as(u32p), status.ctypes.data_as(i32p)))
    assert (back_len == in_len).all()
    assert back_host[:extra.size].tobytes() == extra.tobytes()
    got = d_streams.cpu().numpy()
    for c in range(nc):
        for k in (KQS, KBA):
            o, ln = int(enc.stream_off[c, k]), int(enc.stream_len[c, k])
            assert got[o:o + ln].tobytes() == host.streams[o:o + ln].tobytes(), (c, k)
    # restore with qs / ba as device pointers
    rfl = (enc.force_qs & 1).astype(np.uint8)
    rfl[(rd['flag'] & 4) != 0] |= 2
    keep = []
    I = R.NormIn()
    co = np.ascontiguousarray(cont_off, np.uint64); wo = np.zeros(nc, np.uint64); wl = np.full(nc, ref.size, np.uint64); w0 = np.zeros(nc, np.int64)
    arrs = dict(pos=np.ascontiguousarray(rd['pos'], np.int32), read_off=np.ascontiguousarray(rd['read_off'], np.uint64), rflags=rfl,
                feat_off=enc.feat_off, f_op=enc.f_op if enc.f_op.size else np.zeros(1, np.uint8),
                f_pos=enc.f_pos if enc.f_pos.size else np.zeros(1, np.int32), f_val=enc.f_val if enc.f_val.size else np.zeros(1, np.int32),
                f_aux=enc.f_aux if enc.f_aux.size else np.zeros(1, np.int32), payload=enc.payload if enc.payload.size else np.zeros(1, np.uint8),
                read_qs_off=enc.read_qs_off, read_ba_off=enc.read_ba_off, matrix=np.ascontiguousarray(enc.matrix, np.uint8))
    refa = np.ascontiguousarray(ref, np.uint8)
    keep.extend([co, wo, wl, w0, refa] + list(arrs.values()))
    I.n_reads = n; I.n_containers = nc
    from cramtools_b200._lib import u64p
    i64p = C.POINTER(C.c_int64)
    I.cont_off = co.ctypes.data_as(u64p); I.win_off = wo.ctypes.data_as(u64p); I.win_len = wl.ctypes.data_as(u64p)
    I.win_start0 = w0.ctypes.data_as(i64p); I.ref = refa.ctypes.data; I.ref_total = refa.size
    for k, v in arrs.items():
        setattr(I, k, v.ctypes.data)
    I.payload_len = int(enc.payload.size)
    I.qs = d_streams.data_ptr(); I.ba = d_streams.data_ptr()
    total = int(rd['read_off'][n])
    bases = np.zeros(total, np.uint8); quals = np.zeros(total, np.uint8); star = np.zeros(n, np.uint8); st = np.zeros(n, np.int32)
    gpu_ctx.check(gpu_ctx.lib.cramgpu_normalize_decode(gpu_ctx.handle, C.byref(I), bases.ctypes.data_as(u8p), quals.ctypes.data_as(u8p),
                                                       star.ctypes.data_as(u8p), st.ctypes.data_as(i32p)))
    assert (st == 0).all() and (star == 0).all()
    assert (bases == rd['bases']).all() and (quals == rd['quals']).all()
    # ---- the decode pipeline that never sees a feature on the host: series (the small ones in host memory, QS / BA where the
    # block codec left them on the device) -> cramgpu_features_unpack_resident -> cramgpu_normalize_decode with device pointers
    S = R.Series()
    sa = np.zeros((nc, 12), np.uint64)
    sa[:] = np.uint64(host.streams.ctypes.data) + host.stream_off
    sa[:, KQS] = np.uint64(d_streams.data_ptr()) + enc.stream_off[:, KQS]; sa[:, KBA] = np.uint64(d_streams.data_ptr()) + enc.stream_off[:, KBA]
    sl = np.ascontiguousarray(host.stream_len, np.uint64)
    S.n_reads = n; S.n_containers = nc; S.cont_off = co.ctypes.data_as(u64p); S.streams = None; S.streams_len = 0
    S.stream_off = sa.ctypes.data_as(u64p); S.stream_len = sl.ctypes.data_as(u64p)
    S.read_off = arrs['read_off'].ctypes.data; S.rflags = rfl.ctypes.data
    U = R.Features(); ust = np.zeros(nc, np.int32); umq = np.zeros(n, np.uint8)
    gpu_ctx.check(gpu_ctx.lib.cramgpu_features_unpack_resident(gpu_ctx.handle, C.byref(S), C.byref(U), umq.ctypes.data_as(u8p), ust.ctypes.data_as(i32p)))
    assert (ust == 0).all() and int(U.n_feat) == int(host.n_feat)
    mapped = (rd['flag'] & 4) == 0
    assert (umq[mapped] == rd['mapq'][mapped]).all()
    I.feat_off = U.feat_off; I.f_op = U.f_op; I.f_pos = U.f_pos; I.f_val = U.f_val; I.f_aux = U.f_aux; I.payload = U.payload; I.payload_len = U.payload_len
    I.qs = None; I.ba = None; I.read_qs_off = U.read_qs_off; I.read_ba_off = U.read_ba_off
    bases[:] = 0; quals[:] = 0
    gpu_ctx.check(gpu_ctx.lib.cramgpu_normalize_decode(gpu_ctx.handle, C.byref(I), bases.ctypes.data_as(u8p), quals.ctypes.data_as(u8p),
                                                       star.ctypes.data_as(u8p), st.ctypes.data_as(i32p)))
    assert (st == 0).all() and (star == 0).all()
    assert (bases == rd['bases']).all() and (quals == rd['quals']).all()
    # ---- and the encode side that keeps the feature arrays on the device
    lean, dev2 = R.features_encode_resident(rd, ref, None, cont_off=cont_off, ctx=gpu_ctx, features_on_device_only=True)
    assert lean.f_op.size == 0 and lean.read_qs_off.size == 0 and (lean.force_qs == host.force_qs).all() and (lean.matrix == host.matrix).all()
    assert (lean.stream_off == host.stream_off).all() and (lean.stream_len == host.stream_len).all()
    for c in range(nc):
        for name in R.SERIES:
            if name not in ('QS', 'BA'):
                assert lean.series(c, name) == host.series(c, name), (c, name)
    from cramtools_b200 import device as dv
    assert (dv.dev_to_numpy(gpu_ctx, int(dev2.f_op), int(host.n_feat), np.uint8) == host.f_op).all()


def _nf_flags(enc, rd):
    rfl = (enc.force_qs & 1).astype(np.uint8)
    rfl[(rd['flag'] & 4) != 0] |= 2
    if rd.get('rflags') is not None:
        rfl[(np.asarray(rd['rflags']) & 2) != 0] |= 4          # NO_BASES -> unknown bases
    return rfl


@pytest.mark.parametrize('spec', [None, '*8', 'N40-D8', 'D8-I40-m30'])
def test_unpack_parity_with_oracle(R, gpu_ctx, spec):
    """row P: series -> features on the GPU against the Python restatement of CramRecordReader.read, per container"""
    from oracle import records as orec
    ref, rd, cont_off = _sim(5000, 4, 19, unmapped_tail=120)
    n, nc = len(rd['pos']), 4
    enc = R.features_encode(rd, ref, spec, cont_off=cont_off, ctx=gpu_ctx)
    rfl = _nf_flags(enc, rd)
    up = R.features_unpack(enc.streams, enc.stream_off, enc.stream_len, rd['read_off'], rfl, cont_off=cont_off, ctx=gpu_ctx)
    assert (up.status == 0).all()
    rl = np.diff(np.asarray(rd['read_off'], np.int64))
    for c in range(nc):
        a, b = int(cont_off[c]), int(cont_off[c + 1])
        st = {name: enc.series(c, name) for name in R.SERIES}
        rows, mapq, qs_off, ba_off = orec.read_streams(st, rl[a:b], rfl[a:b])
        for i in range(a, b):
            f0, f1 = int(up.feat_off[i]), int(up.feat_off[i + 1])
            got = []
            for f in range(f0, f1):
                op = chr(up.f_op[f])
                if op == 'S':
                    o, ln = int(up.f_val[f]), int(up.f_aux[f])
                    got.append([op, int(up.f_pos[f]), up.payload[o:o + ln].tobytes()])
                else:
                    got.append([op, int(up.f_pos[f]), int(up.f_val[f])])
            assert got == rows[i - a], (spec, i)
            assert up.mapq[i] == mapq[i - a]
            qo, bo = qs_off[i - a], ba_off[i - a]
            assert int(up.read_qs_off[i]) == (0xFFFFFFFFFFFFFFFF if qo is None else int(enc.stream_off[c, R.SERIES.index('QS')]) + qo)
            assert int(up.read_ba_off[i]) == (0xFFFFFFFFFFFFFFFF if bo is None else int(enc.stream_off[c, R.SERIES.index('BA')]) + bo)
    assert int(up.feat_off[n]) == up.n_feat == enc.n_feat


def test_unpack_then_normalize_round_trip(R, gpu_ctx):
    """decode side without the encoder's feature arrays: series -> features (row P) -> bases and qualities (rows N1, N2)"""
    ref, rd, cont_off = _sim(20000, 6, 29, unmapped_tail=250)
    n = len(rd['pos'])
    enc = R.features_encode(rd, ref, None, cont_off=cont_off, ctx=gpu_ctx)
    rfl = _nf_flags(enc, rd)
    up = R.features_unpack(enc.streams, enc.stream_off, enc.stream_len, rd['read_off'], rfl, cont_off=cont_off, ctx=gpu_ctx)
    assert (up.mapq[(rd['flag'] & 4) == 0] == rd['mapq'][(rd['flag'] & 4) == 0]).all()
    nc = 6
    ws0 = np.zeros(nc, np.int64)
    bases, quals, star, status = R.normalize_decode(rd['pos'], rd['read_off'], rfl, up, ref, ws0, enc.matrix, cont_off=cont_off,
                                                    qs=enc.streams, read_qs_off=up.read_qs_off, ba=enc.streams,
                                                    read_ba_off=up.read_ba_off, ctx=gpu_ctx)
    assert (status == 0).all() and (star == 0).all()
    assert (bases == rd['bases']).all() and (quals == rd['quals']).all()


def test_unpack_corrupt_and_foreign_series(R, gpu_ctx):
    """a series that ends early -> CRAMGPU_ERR_CORRUPT for that container only; an operator the library does not write ->
    CRAMGPU_ERR_UNSUPPORTED"""
    ref, rd, cont_off = _sim(3000, 3, 31)
    enc = R.features_encode(rd, ref, None, cont_off=cont_off, ctx=gpu_ctx)
    rfl = _nf_flags(enc, rd)
    sl = enc.stream_len.copy()
    sl[1, R.SERIES.index('FP')] -= 1                     # container 1 loses the last byte of its FP series
    up = R.features_unpack(enc.streams, enc.stream_off, sl, rd['read_off'], rfl, cont_off=cont_off, ctx=gpu_ctx, check=False)
    assert list(up.status) == [0, _lib.ERR_CORRUPT, 0]
    st = enc.streams.copy()
    o = int(enc.stream_off[2, R.SERIES.index('FC')])
    assert enc.stream_len[2, R.SERIES.index('FC')] > 0
    st[o] = ord('B')                                     # ReadBase: needs the BA+QS pair this library never writes
    up = R.features_unpack(st, enc.stream_off, enc.stream_len, rd['read_off'], rfl, cont_off=cont_off, ctx=gpu_ctx, check=False)
    assert list(up.status) == [0, 0, _lib.ERR_UNSUPPORTED]


def _concat_reads(parts):
    """several synth.reads results over one reference -> one SoA batch"""
    out = {}
    for k in ('bases', 'quals', 'cigar', 'pos', 'flag', 'mapq'):
        out[k] = np.concatenate([p[k] for p in parts])
    ro, co = [np.zeros(1, np.int64)], [np.zeros(1, np.int64)]
    for p in parts:
        ro.append(p['read_off'][1:] + ro[-1][-1]); co.append(p['cigar_off'][1:] + co[-1][-1])
    out['read_off'] = np.concatenate(ro); out['cigar_off'] = np.concatenate(co)
    out['names'] = sum([list(p.get('names', [b''] * len(p['pos']))) for p in parts], [])
    return out


def test_mixed_read_lengths_and_region_edges(R, gpu_ctx):
    """the warp-wide copies only apply to 32 records of one length well inside the reference: records of other lengths, runs
    that change length inside a warp, records at the very start / end of the reference and short (< 16 bases) records take
    the per-record paths -- same results (oracle parity on encode, bit-exact round trip on decode)"""
    from oracle import records as orec
    ref = synth.reference(30000, 77)
    parts = [synth.reads(ref, 700, 78, read_len=150, sub_rate=0.01), synth.reads(ref, 333, 79, read_len=100, sub_rate=0.01),
             synth.reads(ref, 150, 80, read_len=40, indel_frac=0.2, clip_frac=0.1), synth.reads(ref, 97, 81, read_len=33),
             synth.reads(ref, 64, 82, read_len=150, indel_frac=0.0, clip_frac=0.0, contig_start=0, contig_end=260),           # at the start
             synth.reads(ref, 64, 83, read_len=150, indel_frac=0.0, clip_frac=0.0, contig_start=ref.size - 400, contig_end=ref.size)]
    rd = _concat_reads(parts)
    n = len(rd['pos'])
    cont_off = np.array([0, 500, 1100, n], np.uint64)
    enc = R.features_encode(rd, ref, None, cont_off=cont_off, ctx=gpu_ctx)
    for c in range(3):
        a, b = int(cont_off[c]), int(cont_off[c + 1])
        sub = dict(bases=rd['bases'][rd['read_off'][a]:rd['read_off'][b]], quals=rd['quals'][rd['read_off'][a]:rd['read_off'][b]],
                   read_off=rd['read_off'][a:b + 1] - rd['read_off'][a], cigar=rd['cigar'][rd['cigar_off'][a]:rd['cigar_off'][b]],
                   cigar_off=rd['cigar_off'][a:b + 1] - rd['cigar_off'][a], pos=rd['pos'][a:b], flag=rd['flag'][a:b], mapq=rd['mapq'][a:b])
        ex = orec.extract(sub, ref)
        force, scores = orec.quality(sub, ref, None)
        want = orec.write_streams(sub, ex, force, scores)
        for name in R.SERIES:
            assert enc.series(c, name) == want[name], (c, name)
    rfl = (enc.force_qs & 1).astype(np.uint8)
    # decode with the tightest regions: [min start, max end] of each container
    win_off = np.zeros(3, np.uint64); win_len = np.zeros(3, np.uint64); ws0 = np.zeros(3, np.int64)
    rl = np.diff(rd['read_off'])
    for c in range(3):
        a, b = int(cont_off[c]), int(cont_off[c + 1])
        s = int(rd['pos'][a:b].min()) - 1
        e = min(int((rd['pos'][a:b] + rl[a:b] + 8).max()), ref.size)
        win_off[c] = s; win_len[c] = e - s; ws0[c] = s
    up = R.features_unpack(enc.streams, enc.stream_off, enc.stream_len, rd['read_off'], rfl, cont_off=cont_off, ctx=gpu_ctx)
    bases, quals, star, status = R.normalize_decode(rd['pos'], rd['read_off'], rfl, up, ref, ws0, enc.matrix, cont_off=cont_off,
                                                    win_off=win_off, win_len=win_len, qs=enc.streams, read_qs_off=up.read_qs_off,
                                                    ba=enc.streams, read_ba_off=up.read_ba_off, ctx=gpu_ctx)
    assert (status == 0).all() and (star == 0).all()
    assert (bases == rd['bases']).all() and (quals == rd['quals']).all()
    # and under a read-level lossy model: binned arrays of mixed lengths
    encb = R.features_encode(rd, ref, '*8', cont_off=cont_off, ctx=gpu_ctx)
    sub = dict(rd); forceb, scoresb = orec.quality(rd, ref, '*8')
    assert ((encb.force_qs & 1) == (np.asarray(forceb) & 1)).all()
    for i in range(n):
        o, ln = int(encb.read_qs_off[i]), int(rl[i])
        assert encb.streams[o:o + ln].tobytes() == np.asarray(scoresb[rd['read_off'][i]:rd['read_off'][i + 1]], np.uint8).tobytes(), i


@pytest.mark.parametrize('spec', ['*8', 'N40-D8'])
def test_full_size_containers_lossy_pipeline(R, gpu_ctx, spec):
    """BASELINE config 4 at the reference's container size (10 000 reads, 1.5 MB of scores per slice): features under a lossy
    model -> rANS order 1 of the QS series (byte-identical to the oracle) -> decode -> series unpack -> restoration.  Size-
    independent properties: bases come back bit for bit; under *8 every score comes back as its Illumina bin; under N40-D8
    a record either keeps all scores that the model keeps (array) or exactly its Q features, and every kept score is
    the original one or its bin."""
    from oracle import rans as orans
    from cramtools_b200 import ExternalCompression as EC
    ref = synth.reference(140000, 91)
    rd = synth.reads(ref, 40000, 92, with_names=False)
    n, nc = len(rd['pos']), 4
    cont_off = np.arange(nc + 1, dtype=np.uint64) * 10000
    enc = R.features_encode(rd, ref, spec, cont_off=cont_off, ctx=gpu_ctx)
    KQS = R.SERIES.index('QS')
    qs_blocks = [enc.series(c, 'QS') for c in range(nc)]
    comp = EC.rans_batch(qs_blocks, [1] * nc, gpu_ctx)
    for c in range(nc):
        assert comp[c] == orans.compress(np.frombuffer(qs_blocks[c], np.uint8), 1), c
    back = EC.unrans_batch(comp, gpu_ctx)
    assert back == qs_blocks
    rfl = (enc.force_qs & 1).astype(np.uint8)
    up = R.features_unpack(enc.streams, enc.stream_off, enc.stream_len, rd['read_off'], rfl, cont_off=cont_off, ctx=gpu_ctx)
    assert (up.status == 0).all() and up.n_feat == enc.n_feat
    bases, quals, star, status = R.normalize_decode(rd['pos'], rd['read_off'], rfl, up, ref, np.zeros(nc, np.int64), enc.matrix,
                                                    cont_off=cont_off, qs=enc.streams, read_qs_off=up.read_qs_off, ba=enc.streams,
                                                    read_ba_off=up.read_ba_off, ctx=gpu_ctx)
    assert (status == 0).all()
    assert (bases == rd['bases']).all()
    binned = np.array([q if q < 2 else 6 if q < 10 else 15 if q < 20 else 22 if q < 25 else 27 if q < 30 else 33 if q < 35 else 37 if q < 40 else 40
                       for q in range(256)], np.uint8)            # Binning.Illumina_binning_matrix [Binning.java:63-71]
    q_in = np.asarray(rd['quals'], np.uint8)
    if spec == '*8':
        assert (rfl == 1).all() and (star == 0).all()
        assert (quals == binned[q_in]).all()
    else:
        dflt = ord('?') - ord('!')
        ok = (quals == q_in) | (quals == binned[q_in]) | (quals == dflt)
        assert ok.all()
        # a record without a stored array has exactly its Q features' positions set, everything else the default score
        fo = up.feat_off.astype(np.int64)
        isq = up.f_op == ord('Q')
        nq = np.add.reduceat(np.concatenate([isq, [False]]).astype(np.int64), np.minimum(fo[:-1], isq.size)) * (fo[1:] > fo[:-1])
        ro = np.asarray(rd['read_off'], np.int64)
        notd = np.add.reduceat((quals != dflt).astype(np.int64), ro[:-1])
        unforced = rfl == 0
        assert (notd[unforced] <= nq[unforced]).all()


def test_normalize_rejects_out_of_range_features(R, gpu_ctx):
    """A corrupt series can hand the normaliser any position, length or payload offset (the reference throws
    ArrayIndexOutOfBoundsException there): the record must come back CRAMGPU_ERR_CORRUPT and every other record's bases and
    scores must be what they are without the corruption -- nothing may be written outside the record."""
    import types
    ref, rd, cont_off = _sim(3000, 2, 41)
    n = len(rd['pos'])
    enc = R.features_encode(rd, ref, 'N40-I40', cont_off=cont_off, ctx=gpu_ctx)          # sparse Q features, no stored arrays
    rfl = _nf_flags(enc, rd)
    win = dict(window=ref, window_start0=0, matrix=enc.matrix, cont_off=cont_off, win_off=np.zeros(2), win_len=np.full(2, ref.size))
    args = (rd['pos'], rd['read_off'], rfl)
    kw = dict(qs=enc.streams, read_qs_off=enc.read_qs_off, ba=enc.streams, read_ba_off=enc.read_ba_off, ctx=gpu_ctx, check=False, **win)
    base_b, base_q, _, base_st = R.normalize_decode(*args, enc, **kw)
    assert (base_st == 0).all()
    fo = np.asarray(enc.feat_off, np.int64)
    op = np.asarray(enc.f_op)

    def first_with(ch, skip=0):
        idx = np.nonzero(op == ord(ch))[0]
        return int(idx[skip])
    bad = types.SimpleNamespace(f_op=op.copy(), f_pos=np.asarray(enc.f_pos).copy(), f_val=np.asarray(enc.f_val).copy(),
                                f_aux=np.asarray(enc.f_aux).copy(), feat_off=enc.feat_off, payload=enc.payload)
    hit = []
    f = first_with('Q'); bad.f_pos[f] = 0; hit.append(f)                       # score written in front of the record
    f = first_with('Q', 40); bad.f_pos[f] = 1 << 30; hit.append(f)            # ... a gigabyte behind it
    f = first_with('Q', 90); bad.f_pos[f] = -7; hit.append(f)
    f = first_with('X', 5); bad.f_pos[f] = 100000; hit.append(f)
    f = first_with('S'); bad.f_aux[f] = -5; hit.append(f)                      # negative soft-clip length
    f = first_with('S', 3); bad.f_val[f] = int(enc.payload.size) + 1000; hit.append(f)   # payload offset outside the payload
    f = first_with('S', 6); bad.f_aux[f] = 1 << 20; hit.append(f)
    f = first_with('i', 2); bad.f_op[f] = ord('B'); bad.f_pos[f] = 5000; hit.append(f)   # ReadBase far outside
    recs = sorted({int(np.searchsorted(fo, f, side='right') - 1) for f in hit})
    assert len(recs) == len(hit)
    b, q, _, st = R.normalize_decode(*args, bad, **kw)
    ro = np.asarray(rd['read_off'], np.int64)
    for r in range(n):
        if r in recs:
            assert st[r] == _lib.ERR_CORRUPT, r
        else:
            assert st[r] == 0, r
            assert (b[ro[r]:ro[r + 1]] == base_b[ro[r]:ro[r + 1]]).all() and (q[ro[r]:ro[r + 1]] == base_q[ro[r]:ro[r + 1]]).all(), r


@pytest.mark.parametrize('spec', ['*8', 'N40-D8', 'R8-N40', 'm20_8-N40', '*8-*40', 'm5'])
def test_binning_table_overrun_is_reported_per_read(R, gpu_ctx, spec):
    """Binning.Illumina_binning_matrix has 110 entries and a signed byte index: a BIN treatment that meets a score of 110..255
    makes the reference throw ArrayIndexOutOfBoundsException (lossy/Binning.java:63-71, QualityScorePreservation.java:258-266).
    The call fails with CRAMGPU_ERR_UNSUPPORTED, bit 7 of force_qs marks exactly the records the oracle marks, and the
    other records' series are what they are without the fault."""
    from oracle import records as orec
    ref, rd, cont_off = _sim(3000, 3, 23, unmapped_tail=60)
    rng = np.random.default_rng(4)
    quals = rd['quals'].copy()
    for r, q in zip(rng.choice(len(rd['pos']), 40, replace=False), rng.choice([110, 111, 127, 128, 200, 255], 40)):
        lo, hi = int(rd['read_off'][r]), int(rd['read_off'][r + 1])
        quals[lo + rng.integers(0, hi - lo)] = q
    rd = dict(rd, quals=quals)
    want_force = []
    for c in range(3):
        a, b = int(cont_off[c]), int(cont_off[c + 1])
        sub = dict(bases=rd['bases'][rd['read_off'][a]:rd['read_off'][b]], quals=rd['quals'][rd['read_off'][a]:rd['read_off'][b]],
                   read_off=rd['read_off'][a:b + 1] - rd['read_off'][a], cigar=rd['cigar'][rd['cigar_off'][a]:rd['cigar_off'][b]],
                   cigar_off=rd['cigar_off'][a:b + 1] - rd['cigar_off'][a], pos=rd['pos'][a:b], flag=rd['flag'][a:b], mapq=rd['mapq'][a:b])
        want_force.append(orec.quality(sub, ref, spec)[0])
    want_force = np.concatenate(want_force)
    faults = int((want_force & 0x80 != 0).sum())
    if faults == 0:                                   # a model without BIN ('m5'), or none of the scores was binned
        enc = R.features_encode(rd, ref, spec, cont_off=cont_off, ctx=gpu_ctx)
        assert spec == 'm5'
    else:
        with pytest.raises(_lib.CramGpuError) as e:
            R.features_encode(rd, ref, spec, cont_off=cont_off, ctx=gpu_ctx)
        assert e.value.status == _lib.ERR_UNSUPPORTED and 'ArrayIndexOutOfBounds' in str(e.value)
        enc = R.features_encode(rd, ref, spec, cont_off=cont_off, ctx=gpu_ctx, allow_bin_fault=True)
    assert (enc.force_qs == want_force).all(), (spec, faults, int((enc.force_qs & 0x80 != 0).sum()))
    # the context is usable afterwards and a clean call succeeds
    ok = R.features_encode(dict(rd, quals=np.minimum(rd['quals'], 60)), ref, spec, cont_off=cont_off, ctx=gpu_ctx)
    assert not (ok.force_qs & 0x80).any()
