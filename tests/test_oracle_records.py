"""CPU: the C oracle of the record path (oracle/records_oracle.c) against golden vectors that came out of the
reference jar's own bytecode (tests/golden/records_jvm.json, made by oracle/make_golden_records.py)."""
import base64

import numpy as np
import pytest

from oracle import records as orec
from tests import golden_records as gr


@pytest.fixture(scope='module')
def G():
    return gr.load()


def test_itf8_matches_jar(G):
    g, _, _ = G
    for v, b in g['itf8']:
        assert orec.itf8(v) == base64.b64decode(b), v


def test_normalize_base():
    L = orec.lib()
    for b in range(256):
        exp = {'a': 'A', 'A': 'A', 'c': 'C', 'C': 'C', 'g': 'G', 'G': 'G', 't': 'T', 'T': 'T'}.get(chr(b), 'N')
        assert chr(L.rec_oracle_normalize_base(b)) == exp


def test_feature_extraction_matches_jar(G):
    g, ref, rd = G
    ex = orec.extract(rd, ref)
    n = len(rd['pos'])
    got = gr.features_as_lists(ex.f_op, ex.f_pos, ex.f_val, ex.f_aux, ex.feat_off, ex.payload, n)
    for i in range(n):
        want = g['features_coded'][i] or []
        assert got[i] == want, (i, rd['names'][i])


def test_substitution_matrix_matches_jar(G):
    g, ref, rd = G
    ex = orec.extract(rd, ref)
    assert ex.matrix == base64.b64decode(g['subst_matrix'])
    assert ex.subst_counts == g['subst_counts']
    for e in g['subst_extra']:
        m, _ = orec.subst_matrix(e['counts'])
        assert m == base64.b64decode(e['matrix']), e['counts']


def test_lossy_spec_parser_matches_reference_unit_tests():
    # src/test/java/htsjdk/samtools/cram/lossy/TestQualityScorePreservation.java:46-114
    arr, n = orec.parse_policies('m999_8')
    assert n == 1 and arr[0].read_cat == 3 and arr[0].read_param == 999 and arr[0].treatment == 1 and arr[0].treat_param == 8
    arr, n = orec.parse_policies('R8-N40')
    assert n == 2
    assert arr[0].treatment == 1 and arr[0].base_cat[0] == 1          # BIN first: bases MATCH
    assert arr[1].treatment == 0 and arr[1].base_cat[0] == 2          # then PRESERVE: bases MISMATCH
    with pytest.raises(ValueError):
        orec.parse_policies('Z3')


def _q_features(force, scores, rlen):
    if force & 1 or not (force & 2):
        return []
    return [['Q', i + 1, int(scores[i])] for i in range(rlen) if scores[i] < 128]


def test_quality_preservation_matches_jar(G):
    g, ref, rd = G
    n = len(rd['pos'])
    ex = orec.extract(rd, ref)
    base_feats = gr.features_as_lists(ex.f_op, ex.f_pos, ex.f_val, ex.f_aux, ex.feat_off, ex.payload, n, coded=False)
    for spec, rows in g['lossy'].items():
        force, scores = orec.quality(rd, ref, spec)
        for i, row in enumerate(rows):
            if 'throws' in row:
                continue                      # the reference itself fails on this read (e.g. no bases)
            lo, hi = int(rd['read_off'][i]), int(rd['read_off'][i + 1])
            assert (force[i] & 1) == row['force'], (spec, i, rd['names'][i])
            if row['scores'] == '*':
                assert not (force[i] & 2)
            elif row['scores'] is not None:
                assert scores[lo:hi].tobytes() == base64.b64decode(row['scores']), (spec, i, rd['names'][i])
            # feature list after addQualityScores: originals + Q features, stably sorted by position
            merged = sorted(base_feats[i] + _q_features(int(force[i]), scores[lo:hi], hi - lo), key=lambda f: f[1])
            assert merged == (row['features'] or []), (spec, i, rd['names'][i])


def test_restore_bases_matches_jar(G):
    g, ref, rd = G
    ex = orec.extract(rd, ref)
    matrix = base64.b64decode(g['subst_matrix'])
    for name, case in g['restore_bases'].items():
        ws, we = case['window_start0'], case['window_end0']
        for i, want in enumerate(case['bases']):
            if want is None:
                continue
            rlen = int(rd['read_off'][i + 1] - rd['read_off'][i])
            got = orec.restore_bases(int(rd['pos'][i]), rlen, ex, i, ref[ws:we], ws, matrix)
            if isinstance(want, dict):
                assert got is None, (name, i)
            else:
                assert got == base64.b64decode(want), (name, i, rd['names'][i])


def test_restore_quals_matches_jar(G):
    g, ref, rd = G
    for spec, rows in g['restore_quals'].items():
        force, scores = orec.quality(rd, ref, spec)
        for i, want in enumerate(rows):
            if isinstance(want, dict):
                continue
            lo, hi = int(rd['read_off'][i]), int(rd['read_off'][i + 1])
            sc = scores[lo:hi]
            if force[i] & 1:
                got = orec.restore_quals(True, hi - lo, sc, [], [])
            else:
                q = _q_features(int(force[i]), sc, hi - lo)
                got = orec.restore_quals(False, hi - lo, None, [f[1] for f in q], [f[2] for f in q])
            if want == '*':
                assert got == '*', (spec, i)
            else:
                assert got == base64.b64decode(want), (spec, i, rd['names'][i])


def test_streams_consistent_with_features(G):
    """rows W and P are restated from the disassembly only: check them against each other and against the jar -- the
    series write_streams makes from the jar's feature lists, read back by read_streams the way CramRecordReader would,
    must reproduce those lists (FN counts, FP deltas, operator values), every mapping quality, and use up every byte."""
    g, ref, rd = G
    n = len(rd['pos'])
    ex = orec.extract(rd, ref)
    rl = np.diff(np.asarray(rd['read_off'], np.int64))
    for spec in ('lossless', 'N40-D8', '*8'):
        force, scores = orec.quality(rd, ref, spec)
        st = orec.write_streams(rd, ex, force, scores)
        rfl = (np.asarray(force) & 1).astype(np.uint8)
        rfl[(np.asarray(rd['flag']) & 4) != 0] |= 2
        rows, mapq, qs_off, ba_off = orec.read_streams(st, rl, rfl)
        want = g['features_coded'] if spec == 'lossless' else None
        for i in range(n):
            if rd['flag'][i] & 4:
                assert rows[i] == [] and ba_off[i] is not None
                continue
            assert mapq[i] == rd['mapq'][i]
            row = [[f[0], f[1], base64.b64encode(f[2]).decode()] if f[0] == 'S' else f for f in rows[i]]
            if want is not None:
                exp = [[f[0], f[1], f[4]] if f[0] == 'X' else f for f in (want[i] or [])]
                assert row == exp, (i, rd['names'][i])
            elif spec in g['lossy']:
                exp = [[f[0], f[1]] for f in (g['lossy'][spec][i].get('features') or [])]
                assert [[f[0], f[1]] for f in row] == exp, (spec, i)
            assert (qs_off[i] is not None) == bool(force[i] & 1)
