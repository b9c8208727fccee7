"""CPU: the oracle's ReferenceTracks / updateTracks / X and P selectors against what the reference jar itself
computed, container after container on one tracks object, in the mini-JVM (tests/golden/tracks_jvm.json, written by
oracle/make_golden_tracks.py): window position, the coverage and mismatch tracks after every container, the score array and
the force flag of every record -- for a window that covers the reference, one that moves, and one smaller than a
container's span (it moves while the records are looked at)."""
import base64

import numpy as np
import pytest

from oracle import records as orec
from tests import golden_tracks as gt


@pytest.fixture(scope='module')
def G():
    return gt.load()


def test_policy_parser_accepts_coverage_and_pileup_selectors():
    arr, n = orec.parse_policies('P0_40-X4_8-D8')
    got = [(arr[k].treatment, [arr[k].base_cat[b] for b in range(arr[k].n_base_cats)]) for k in range(n)]
    # sorted DROP, BIN, PRESERVE: P0 reads its (absent) treatment as 0 = drop, then "_40" makes it preserve
    assert got == [(1, [5 | (4 << 8)]), (1, [3]), (0, [6 | (0 << 8)])]
    arr, n = orec.parse_policies('X6P1_40')
    assert n == 1 and [arr[0].base_cat[b] for b in range(2)] == [5 | (6 << 8), 6 | (1 << 8)] and arr[0].treatment == 0


def test_tracks_and_selectors_match_jar(G):
    g, ref, conts = G
    seen_move = seen_mid = seen_scores = 0
    for sc in g['scenarios']:
        t = orec.Tracks(ref, sc['window'])
        for c, (rd, row) in enumerate(zip(conts, sc['containers'])):
            tag = (sc['window'], sc['spec'], c)
            span = t.container(rd)
            if 'throws' in row:
                assert orec.Tracks.THROWN[t.thrown] == row['throws'], tag
                break
            assert t.thrown == 0, tag
            assert list(span) == row['span'], tag
            if 'coverage' in row:
                assert t.coverage.tolist() == row['coverage'] and t.mismatches.tolist() == row['mismatches'], tag
            positions = []
            force, scores = orec.quality(rd, ref, sc['spec'], tracks=t, positions=positions)
            assert t.thrown == 0, tag
            for r, rec in enumerate(row['records']):
                lo, hi = int(rd['read_off'][r]), int(rd['read_off'][r + 1])
                assert 'throws' not in rec, tag
                assert positions[r] == rec['position'], tag + (r,)
                assert (force[r] & 1) == rec['force'], tag + (r,)
                if rec['scores'] is None or rec['scores'] == '*':
                    assert not (force[r] & 2) or rec['scores'] is None, tag + (r,)
                else:
                    assert scores[lo:hi].tobytes() == base64.b64decode(rec['scores']), tag + (r,)
                    seen_scores += 1
            seen_move += row['position_after_ensure'] > 1
            seen_mid += row['position_at_end'] != row['position_after_ensure']
            assert t.position == row['position_at_end'], tag
    assert seen_move > 20 and seen_mid > 5 and seen_scores > 2000
