"""GPU parity for the lossy selectors X (coverage below n) and P (more than n mismatches piled up), SURVEY 8f-4:
cramgpu_tracks_update_dev + cramgpu_features_encode against the oracle's ReferenceTracks restatement, which
tests/test_oracle_tracks.py pins to the jar run container after container in the mini-JVM (tests/golden/tracks_jvm.json).
The containers go in as ONE batch (container c must see the counts of containers 0..c only), as two calls (the tracks carry
over), with a window that covers the reference, one that moves and clamps at the reference end, and one smaller than a
container's span (refused where the reference would move the window between records)."""
import numpy as np
import pytest

from cramtools_b200 import _lib
from tests import golden_tracks as gt

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def G():
    return gt.load()


def _concat(conts):
    out = {}
    for k in ('bases', 'quals', 'cigar', 'pos', 'flag', 'mapq', 'rflags'):
        out[k] = np.concatenate([c[k] for c in conts])
    for k, base in (('read_off', 'bases'), ('cigar_off', 'cigar')):
        offs, tot = [np.zeros(1, np.int64)], 0
        for c in conts:
            offs.append(c[k][1:] + tot); tot += int(c[k][-1])
        out[k] = np.concatenate(offs)
    cont_off = np.concatenate([[0], np.cumsum([len(c['pos']) for c in conts])]).astype(np.uint64)
    return out, cont_off


def _check(R, gpu_ctx, ref, conts, spec, tracks, otracks, tag):
    """one call for `conts` on the device tracks and on the oracle's; every series and force flag must agree"""
    from oracle import records as orec
    rd, cont_off = _concat(conts)
    st = tracks.update(rd, spec, cont_off)
    assert st == [0] * len(conts), tag
    enc = R.features_encode(rd, ref, spec, cont_off=cont_off, ctx=gpu_ctx)
    for c, sub in enumerate(conts):
        a, b = int(cont_off[c]), int(cont_off[c + 1])
        otracks.container(sub)
        force, scores = orec.quality(sub, ref, spec, tracks=otracks)
        assert otracks.thrown == 0
        ex = orec.extract(sub, ref)
        want = orec.write_streams(sub, ex, force, scores)
        assert (enc.force_qs[a:b] == force).all(), tag + (c, np.flatnonzero(enc.force_qs[a:b] != force)[:5].tolist())
        for name in R.SERIES:
            assert enc.series(c, name) == want[name], tag + (c, name)
    assert tracks.position == otracks.position, tag


@pytest.mark.parametrize('window', [1000000, 1500])
def test_selectors_match_oracle_one_batch_and_two_calls(G, gpu_ctx, window):
    from cramtools_b200 import records as R
    from oracle import records as orec
    g, ref, conts = G
    for spec in sorted({s['spec'] for s in g['scenarios']}):
        t = R.ReferenceTracks(ref, window, ctx=gpu_ctx); o = orec.Tracks(ref, window)
        _check(R, gpu_ctx, ref, conts, spec, t, o, (window, spec, 'one batch'))
        t.close()
        t = R.ReferenceTracks(ref, window, ctx=gpu_ctx); o = orec.Tracks(ref, window)
        _check(R, gpu_ctx, ref, conts[:5], spec, t, o, (window, spec, 'first call'))
        _check(R, gpu_ctx, ref, conts[5:], spec, t, o, (window, spec, 'second call'))
        t.close()


def test_window_smaller_than_a_container_is_refused_where_the_reference_moves_it(G, gpu_ctx):
    from cramtools_b200 import records as R
    from oracle import records as orec
    g, ref, conts = G
    sc = [s for s in g['scenarios'] if s['window'] == 700 and s['spec'] == 'X3'][0]
    jar_moves = [c for c, row in enumerate(sc['containers']) if row['position_at_end'] != row['position_after_ensure']]
    assert jar_moves, 'the golden scenario must hold a window that moves between records'
    rd, cont_off = _concat(conts)
    t = R.ReferenceTracks(ref, 700, ctx=gpu_ctx)
    with pytest.raises(_lib.CramGpuError) as e:
        t.update(rd, 'X3', cont_off)
    assert e.value.status == _lib.ERR_UNSUPPORTED and 'window' in str(e.value)
    first = next(c for c, s in enumerate(e.value.container_status) if s)
    assert first <= jar_moves[0] and all(s == _lib.ERR_UNSUPPORTED for s in e.value.container_status[first:])
    with pytest.raises(_lib.CramGpuError):          # the tracks stay unusable, as after the reference's exception
        t.update(rd, 'X3', cont_off)
    t.close()
    if first:                                       # what comes before the refused container is exact
        t = R.ReferenceTracks(ref, 700, ctx=gpu_ctx); o = orec.Tracks(ref, 700)
        _check(R, gpu_ctx, ref, conts[:first], 'X3', t, o, (700, 'X3', 'prefix'))
        t.close()


def test_selectors_need_tracks_and_backwards_window_is_an_error(G, gpu_ctx):
    from cramtools_b200 import records as R
    g, ref, conts = G
    rd, cont_off = _concat(conts[:2])
    with pytest.raises(_lib.CramGpuError) as e:
        R.features_encode(rd, ref, 'X3', cont_off=cont_off, ctx=gpu_ctx)
    assert 'requires reference tracks' in str(e.value)
    t = R.ReferenceTracks(ref, 1500, ctx=gpu_ctx)
    rd2, co2 = _concat(conts[6:8])
    t.update(rd2, 'P1_8', co2)
    with pytest.raises(_lib.CramGpuError) as e:     # a different model than the tracks call prepared
        R.features_encode(rd2, ref, 'X3', cont_off=co2, ctx=gpu_ctx)
    assert 'requires reference tracks' in str(e.value)
    with pytest.raises(_lib.CramGpuError) as e:     # containers that lie before the window
        t.update(rd, 'P1_8', cont_off)
    assert e.value.status == _lib.ERR_ARG and 'backwards' in str(e.value)
    t.close()
    # a model without X / P does not touch the tracks
    t = R.ReferenceTracks(ref, 1500, ctx=gpu_ctx)
    assert t.update(rd2, 'N40-D8', co2) == [0, 0] and t.position == 1
    t.close()


def test_larger_batch_against_oracle(gpu_ctx):
    """simulated reads at depth on a 60 kbp reference, 12 containers in one call and a 20 kbp window that moves three times"""
    from cramtools_b200 import records as R, synth
    from oracle import records as orec
    ref = synth.reference(60000, 5)
    rd = synth.reads(ref, 4800, 6, sub_rate=0.01, indel_frac=0.15, clip_frac=0.1)
    n = len(rd['pos'])
    cont_off = np.linspace(0, n, 13).astype(np.uint64)
    conts = []
    for c in range(12):
        a, b = int(cont_off[c]), int(cont_off[c + 1])
        conts.append(dict(bases=rd['bases'][rd['read_off'][a]:rd['read_off'][b]], quals=rd['quals'][rd['read_off'][a]:rd['read_off'][b]],
                          read_off=rd['read_off'][a:b + 1] - rd['read_off'][a], cigar=rd['cigar'][rd['cigar_off'][a]:rd['cigar_off'][b]],
                          cigar_off=rd['cigar_off'][a:b + 1] - rd['cigar_off'][a], pos=rd['pos'][a:b], flag=rd['flag'][a:b], mapq=rd['mapq'][a:b],
                          rflags=np.zeros(b - a, np.uint8)))
    for spec in ('X9_8-P1_40', 'P0_8-N40-X12'):
        t = R.ReferenceTracks(ref, 20000, ctx=gpu_ctx); o = orec.Tracks(ref, 20000)
        _check(R, gpu_ctx, ref, conts, spec, t, o, (20000, spec))
        assert t.position > 1
        t.close()


def _unplaced(n, seed):
    rng = np.random.default_rng(seed)
    ln = rng.integers(20, 60, n)
    return dict(bases=rng.choice(np.frombuffer(b'ACGTN', np.uint8), int(ln.sum())).astype(np.uint8), quals=rng.integers(2, 41, int(ln.sum())).astype(np.uint8),
                read_off=np.concatenate([[0], np.cumsum(ln)]).astype(np.int64), cigar=np.zeros(0, np.uint32), cigar_off=np.zeros(n + 1, np.int64),
                pos=np.zeros(n, np.int32), flag=np.full(n, 4, np.uint16), mapq=np.zeros(n, np.uint8), ref_id=np.full(n, -1, np.int32),
                rflags=np.zeros(n, np.uint8))


def test_unplaced_records_and_containers_leave_the_tracks_alone(G, gpu_ctx):
    """a container of unplaced reads between placed ones, and unplaced reads at the end of a placed container: not in the spans,
    not counted, no per-base masks (addQS returns at NO_ALIGNMENT_START)"""
    from cramtools_b200 import records as R
    from oracle import records as orec
    g, ref, conts = G
    tail, _ = _concat([conts[2], _unplaced(3, 1)])
    tail['ref_id'] = np.concatenate([conts[2]['ref_id'], np.full(3, -1, np.int32)])
    mixed = [conts[0], conts[1], _unplaced(5, 2), tail, conts[3]]
    for spec in ('X3', 'P0_40-X4_8-D8'):
        t = R.ReferenceTracks(ref, 1500, ctx=gpu_ctx); o = orec.Tracks(ref, 1500)
        _check(R, gpu_ctx, ref, mixed, spec, t, o, (1500, spec, 'unplaced'))
        t.close()


def test_record_without_bases_is_refused(G, gpu_ctx):
    from cramtools_b200 import records as R
    g, ref, conts = G
    c = {k: v.copy() for k, v in conts[1].items()}
    # the third record loses its bases (SAMRecord.NULL_SEQUENCE) but keeps its CIGAR: updateTracks indexes an empty array
    lo, hi = int(c['read_off'][2]), int(c['read_off'][3])
    c['bases'] = np.concatenate([c['bases'][:lo], c['bases'][hi:]]); c['quals'] = np.concatenate([c['quals'][:lo], c['quals'][hi:]])
    c['read_off'] = np.concatenate([c['read_off'][:3], c['read_off'][3:] - (hi - lo)])
    rd, cont_off = _concat([conts[0], c])
    t = R.ReferenceTracks(ref, 1500, ctx=gpu_ctx)
    with pytest.raises(_lib.CramGpuError) as e:
        t.update(rd, 'X3', cont_off)
    assert e.value.status == _lib.ERR_UNSUPPORTED and e.value.container_status == [0, _lib.ERR_UNSUPPORTED]
    t.close()
