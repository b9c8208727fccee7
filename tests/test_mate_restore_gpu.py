"""GPU parity for the decoder's pairing restoration (the inverse of mate linking): cramgpu_mate_restore_dev against what
CramNormalizer.normalize left in the records the jar's reader read back (mini-JVM, tests/golden/streams_jvm.json
'normalized_mates'), against the oracle's restatement on adversarial link structures (chains of three and more, several
records pointing to one, links into detached records, links out of the container), and as a round trip with
cramgpu_mate_link at scale."""
import numpy as np
import pytest

from cramtools_b200 import _lib
from tests import golden_streams as gs

pytestmark = pytest.mark.gpu


def _restore(R, ctx, a, cont_off=None):
    return R.mate_restore(a['flag'], a['cf'], a['mate_flags'], a['nf'], a['ref_id'], a['pos'], a['mate_ref_id'], a['mate_pos'], a['tlen'],
                          a['read_off'], a['f_op'], a['f_val'], a['f_aux'], a['feat_off'], cont_off=cont_off, ctx=ctx)


def test_pairing_restoration_matches_jar(gpu_ctx):
    from cramtools_b200 import records as R
    g, ref, conts = gs.load()
    for c, rd, series in conts:
        a = gs.read_back_arrays(c)
        got = _restore(R, gpu_ctx, a)
        want = c['normalized_mates']
        for k, key in (('mate_flags', 'mate_flags'), ('mate_ref_id', 'mate_seq_id'), ('mate_pos', 'mate_start'), ('tlen', 'tlen'), ('nf', 'nf'),
                       ('pos', 'alignment_start')):
            assert got[k].tolist() == [w[key] for w in want], (c['tag'], k)
        assert got['status'].tolist() == [0]


def _random_records(rng, n, per_cont):
    a = dict(flag=(rng.integers(0, 2, n) | (rng.integers(0, 2, n) << 2) | (rng.integers(0, 2, n) << 4) | (rng.integers(0, 4, n) << 6)).astype(np.uint16),
             cf=(rng.choice([1, 3, 5, 5, 5, 7], n)).astype(np.uint8), mate_flags=rng.integers(0, 4, n).astype(np.uint8),
             nf=rng.choice([0, 0, 1, 2, 5], n).astype(np.int32), ref_id=rng.choice([0, 0, 0, 1, -1], n).astype(np.int32),
             pos=rng.integers(1, 5000, n).astype(np.int32), mate_ref_id=rng.integers(-1, 3, n).astype(np.int32),
             mate_pos=rng.integers(0, 5000, n).astype(np.int32), tlen=rng.integers(-500, 500, n).astype(np.int32))
    a['flag'] |= 1                                        # mostly multi-fragment, a few not
    a['flag'][rng.random(n) < 0.1] &= 0xFFFE
    rl = rng.integers(30, 151, n)
    a['read_off'] = np.concatenate([[0], np.cumsum(rl)]).astype(np.uint64)
    nfe = rng.integers(0, 3, n)
    a['feat_off'] = np.concatenate([[0], np.cumsum(nfe)]).astype(np.uint64)
    m = int(nfe.sum())
    a['f_op'] = rng.choice(np.frombuffer(b'XiSDNQ', np.uint8), max(m, 1)).astype(np.uint8)
    a['f_val'] = rng.integers(1, 9, max(m, 1)).astype(np.int32); a['f_aux'] = rng.integers(1, 9, max(m, 1)).astype(np.int32)
    a['n'] = n
    cont_off = np.arange(0, n + 1, per_cont).astype(np.uint64)
    if cont_off[-1] != n:
        cont_off = np.append(cont_off, np.uint64(n))
    return a, cont_off


def test_adversarial_links_against_oracle(gpu_ctx):
    from cramtools_b200 import records as R
    from oracle import records as orec
    rng = np.random.default_rng(5)
    a, cont_off = _random_records(rng, 6000, 500)
    got = _restore(R, gpu_ctx, a, cont_off)
    chains3 = shared = 0
    for c in range(cont_off.size - 1):
        lo, hi = int(cont_off[c]), int(cont_off[c + 1])
        sub = {k: a[k][lo:hi] for k in ('flag', 'cf', 'nf', 'ref_id', 'pos', 'mate_flags', 'mate_ref_id', 'mate_pos', 'tlen')}
        sa = dict(sub, n=hi - lo, read_off=a['read_off'][lo:hi + 1] - a['read_off'][lo], feat_off=a['feat_off'][lo:hi + 1] - a['feat_off'][lo],
                  f_op=a['f_op'][int(a['feat_off'][lo]):], f_val=a['f_val'][int(a['feat_off'][lo]):], f_aux=a['f_aux'][int(a['feat_off'][lo]):])
        try:
            want = orec.restore_mates(sub['flag'], sub['cf'], sub['nf'], sub['ref_id'], sub['pos'], sub['mate_flags'], sub['mate_ref_id'],
                                      sub['mate_pos'], sub['tlen'], gs.read_back_ends(sa))
        except IndexError:
            assert got['status'][c] == _lib.ERR_CORRUPT, c
            continue
        assert got['status'][c] == 0, c
        for k, key in (('mate_flags', 'mate_flags'), ('mate_ref_id', 'mate_seq_id'), ('mate_pos', 'mate_pos'), ('tlen', 'tlen'), ('nf', 'nf'), ('pos', 'pos')):
            assert got[k][lo:hi].tolist() == [int(x) for x in want[key]], (c, k)
        nx, pv = want['next'], want['previous']
        chains3 += sum(1 for i in range(hi - lo) if nx[i] >= 0 and nx[nx[i]] >= 0)
        tgt = [x for x in nx if x >= 0]
        shared += len(tgt) - len(set(tgt))
    assert chains3 > 20 and shared > 20 and (got['status'] != 0).any() and (got['status'] == 0).any()


def test_link_then_restore_round_trip(gpu_ctx):
    """proper pairs at scale: what cramgpu_mate_link links, cramgpu_mate_restore_dev gives back"""
    from cramtools_b200 import records as R
    rng = np.random.default_rng(9)
    npair, per = 20000, 10000
    p1 = np.sort(rng.integers(1, 3000000, npair)).astype(np.int64)
    ins = rng.integers(160, 600, npair)
    pos = np.concatenate([p1, p1 + ins - 100]); mate = np.concatenate([p1 + ins - 100, p1])
    flag = np.concatenate([np.full(npair, 0x63), np.full(npair, 0x93)])           # first forward / last reverse
    tlen = np.concatenate([ins, -ins]); names = np.concatenate([np.arange(npair), np.arange(npair)])
    order = np.argsort(pos, kind='stable')
    pos, mate, flag, tlen, names = pos[order], mate[order], flag[order], tlen[order], names[order]
    n = 2 * npair
    read_off = (np.arange(n + 1) * 100).astype(np.uint64)
    feat_off = np.zeros(n + 1, np.uint64)
    z8, z32 = np.zeros(1, np.uint8), np.zeros(1, np.int32)
    cont_off = np.arange(0, n + 1, per).astype(np.uint64)
    cf, nf = R.mate_link(flag.astype(np.uint16), [b'r%d' % k for k in names], pos.astype(np.int32), np.zeros(n, np.int32), mate.astype(np.int32),
                         tlen.astype(np.int32), read_off, z8, z32, z32, feat_off, ref_id=np.zeros(n, np.int32), cont_off=cont_off, ctx=gpu_ctx)
    linked = (cf & 2) == 0
    assert linked.sum() > npair                    # most pairs fall into one container and are linked
    # what the reader hands over: CF bit 0 set (scores as array), detached records with their mate fields, linked ones without
    a = dict(flag=flag.astype(np.uint16), cf=(cf | 1).astype(np.uint8), nf=nf.astype(np.int32), ref_id=np.zeros(n, np.int32), pos=pos.astype(np.int32),
             mate_flags=np.where(linked, 0, np.where(flag & 0x20, 1, 0)).astype(np.uint8), mate_ref_id=np.where(linked, -1, 0).astype(np.int32),
             mate_pos=np.where(linked, 0, mate).astype(np.int32), tlen=np.where(linked, 0, tlen).astype(np.int32),
             read_off=read_off, feat_off=feat_off, f_op=z8, f_val=z32, f_aux=z32)
    got = _restore(R, gpu_ctx, a, cont_off)
    assert (got['status'] == 0).all()
    assert (got['mate_pos'] == mate).all() and (got['mate_ref_id'] == 0).all() and (got['tlen'] == tlen).all()
    assert (got['mate_flags'] == np.where(flag & 0x20, 1, 0)).all()


def test_backward_and_out_of_container_links_are_corrupt(gpu_ctx):
    """a link that points at or before its own record (the reference's list walk would never end) or beyond the container
    (records.get() throws): the container is reported, its neighbours are restored"""
    from cramtools_b200 import records as R
    n = 12
    z32 = np.zeros(n, np.int32)
    a = dict(flag=np.full(n, 0x41, np.uint16), cf=np.full(n, 3, np.uint8), mate_flags=np.zeros(n, np.uint8), nf=np.full(n, -1, np.int32),
             ref_id=z32.copy(), pos=(np.arange(n) * 10 + 1).astype(np.int32), mate_ref_id=np.full(n, -1, np.int32), mate_pos=z32.copy(), tlen=z32.copy(),
             read_off=(np.arange(n + 1) * 50).astype(np.uint64), feat_off=np.zeros(n + 1, np.uint64), f_op=np.zeros(1, np.uint8),
             f_val=np.zeros(1, np.int32), f_aux=np.zeros(1, np.int32))
    a['cf'][0] = 5; a['nf'][0] = 1; a['cf'][2] = 1                     # container 0: record 0 -> record 2, fine
    a['cf'][4] = 5; a['nf'][4] = -1                                    # container 1: points at itself
    a['cf'][9] = 5; a['nf'][9] = 5                                     # container 2: points beyond the container
    got = _restore(R, gpu_ctx, a, np.array([0, 4, 8, 12], np.uint64))
    assert got['status'].tolist() == [0, _lib.ERR_CORRUPT, _lib.ERR_CORRUPT]
    assert got['mate_pos'][0] == 21 and got['mate_pos'][2] == 1 and got['tlen'][0] == 21 and got['tlen'][2] == -21
