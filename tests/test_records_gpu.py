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
    with pytest.raises(_lib.CramGpuError):
        R.QualityScorePreservation.lossyFromSpec('X10')


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
