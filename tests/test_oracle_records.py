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
    """row W is restated from the disassembly only: check its structure -- FN counts, FP deltas and payloads must
    reproduce the jar's feature lists when read back the way CramRecordReader would."""
    g, ref, rd = G
    n = len(rd['pos'])
    ex = orec.extract(rd, ref)
    for spec in ('lossless', 'N40-D8', '*8'):
        force, scores = orec.quality(rd, ref, spec)
        st = orec.write_streams(rd, ex, force, scores)
        want = g['features_coded'] if spec == 'lossless' else None
        # walk the streams
        cur = {k: 0 for k in st}

        def itf8(name):
            b = st[name]; p = cur[name]; v = b[p]
            if v < 0x80:
                cur[name] = p + 1; return v
            if v < 0xC0:
                cur[name] = p + 2; return ((v & 0x7F) << 8) | b[p + 1]
            if v < 0xE0:
                cur[name] = p + 3; return ((v & 0x3F) << 16) | (b[p + 1] << 8) | b[p + 2]
            raise AssertionError('long ITF8 not expected here')

        def byte(name):
            v = st[name][cur[name]]; cur[name] += 1; return v
        for i in range(n):
            if rd['flag'][i] & 4:
                rl = int(rd['read_off'][i + 1] - rd['read_off'][i])
                cur['BA'] += rl                      # unmapped: bases go to BA one by one [Writer.java:254-258]
                if force[i] & 1:
                    cur['QS'] += rl
                continue
            nf = itf8('FN')
            prev = 0
            row = []
            for _ in range(nf):
                op = chr(byte('FC')); prev += itf8('FP')
                if op == 'X':
                    row.append(['X', prev, byte('BS')])
                elif op == 'D':
                    row.append(['D', prev, itf8('DL')])
                elif op == 'N':
                    row.append(['N', prev, itf8('RS')])
                elif op == 'H':
                    row.append(['H', prev, itf8('HC')])
                elif op == 'P':
                    row.append(['P', prev, itf8('PD')])
                elif op == 'i':
                    row.append(['i', prev, byte('BA')])
                elif op == 'Q':
                    row.append(['Q', prev, byte('QS')])
                elif op == 'S':
                    e = st['SC'].index(b'\t', cur['SC'])
                    row.append(['S', prev, base64.b64encode(st['SC'][cur['SC']:e]).decode()]); cur['SC'] = e + 1
            assert itf8('MQ') == rd['mapq'][i]
            if force[i] & 1:
                rl = int(rd['read_off'][i + 1] - rd['read_off'][i])
                cur['QS'] += rl
            if want is not None:
                exp = [[f[0], f[1], f[4]] if f[0] == 'X' else f for f in (want[i] or [])]
                assert row == exp, (i, rd['names'][i])
            elif spec in g['lossy']:
                exp = [[f[0], f[1]] for f in (g['lossy'][spec][i].get('features') or [])]
                assert [[f[0], f[1]] for f in row] == exp, (spec, i)
        for k in st:
            assert cur[k] == len(st[k]), (spec, k)      # every byte of every stream accounted for
