"""GPU parity for rows W and P against what the reference jar itself wrote and read (tests/golden/streams_jvm.json):
the series cramgpu_features_encode packs must equal, byte for byte, the external blocks' raw content that
ContainerFactory.buildSlice / Writer.write produced in the mini-JVM; cramgpu_features_unpack over the jar's own series
must return the records the jar's CramRecordReader read back."""
import base64

import numpy as np
import pytest

from tests import golden_streams as gs

pytestmark = pytest.mark.gpu

CID = {'BF': 0, 'AP': 1, 'FP': 2, 'FC': 3, 'QS': 4, 'DL': 5, 'BA': 6, 'TN': 7, 'NF': 8, 'RL': 9, 'RG': 10, 'MQ': 11, 'RN': 12,
       'NP': 13, 'TS': 14, 'FN': 15, 'BS': 16, 'IN': 17, 'TC': 18, 'MF': 19, 'NS': 20, 'CF': 21, 'TL': 24, 'RI': 25, 'RS': 26,
       'SC': 27, 'HC': 28, 'PD': 29}


@pytest.fixture(scope='module')
def R():
    from cramtools_b200 import records
    return records


@pytest.fixture(scope='module')
def G():
    return gs.load()


def test_feature_series_match_jar(R, G, gpu_ctx):
    g, ref, conts = G
    for c, rd, series in conts:
        enc = R.features_encode(rd, ref, None, ctx=gpu_ctx)
        assert enc.matrix[0, :5].tobytes() == base64.b64decode(c['subst_matrix']), c['tag']
        for nm in R.SERIES:
            assert enc.series(0, nm) == series.get(CID[nm], b''), (c['tag'], nm)


def test_all_containers_in_one_call_match_jar(R, G, gpu_ctx):
    """the three slices as one batch of three containers"""
    g, ref, conts = G
    cat = {}
    for k in ('bases', 'quals', 'cigar'):
        cat[k] = np.concatenate([rd[k] for _, rd, _ in conts])
    for k in ('pos', 'flag', 'mapq', 'rflags'):
        cat[k] = np.concatenate([rd[k] for _, rd, _ in conts])
    ro, co, cont_off = [np.zeros(1, np.int64)], [np.zeros(1, np.int64)], [0]
    for _, rd, _ in conts:
        ro.append(rd['read_off'][1:] + ro[-1][-1]); co.append(rd['cigar_off'][1:] + co[-1][-1]); cont_off.append(cont_off[-1] + len(rd['pos']))
    cat['read_off'] = np.concatenate(ro); cat['cigar_off'] = np.concatenate(co)
    enc = R.features_encode(cat, ref, None, cont_off=np.array(cont_off, np.uint64), ctx=gpu_ctx)
    for ci, (c, rd, series) in enumerate(conts):
        assert enc.matrix[ci, :5].tobytes() == base64.b64decode(c['subst_matrix']), c['tag']
        for nm in R.SERIES:
            assert enc.series(ci, nm) == series.get(CID[nm], b''), (c['tag'], nm)


def test_unpack_matches_jar_reader(R, G, gpu_ctx):
    g, ref, conts = G
    for c, rd, series in conts:
        back = c['read_back']
        n = len(back)
        # the jar's series laid out the way the decoder leaves them: every series at a multiple of 16
        buf, off, ln = bytearray(), np.zeros((1, len(R.SERIES)), np.uint64), np.zeros((1, len(R.SERIES)), np.uint64)
        for i, nm in enumerate(R.SERIES):
            b = series.get(CID[nm], b'')
            off[0, i] = len(buf); ln[0, i] = len(b)
            buf += b + b'\0' * (-len(b) % 16)
        streams = np.frombuffer(bytes(buf) + b'\0' * 16, np.uint8)
        read_off = np.concatenate([[0], np.cumsum([b['read_length'] for b in back])]).astype(np.uint64)
        rfl = np.array([(b['cf'] & 1) | (2 if b['flags'] & 4 else 0) | (4 if b['cf'] & 8 else 0) for b in back], np.uint8)
        up = R.features_unpack(streams, off, ln, read_off, rfl, ctx=gpu_ctx)
        assert (up.status == 0).all()
        qs0, ba0 = int(off[0, R.SERIES.index('QS')]), int(off[0, R.SERIES.index('BA')])
        for i, b in enumerate(back):
            f0, f1 = int(up.feat_off[i]), int(up.feat_off[i + 1])
            got = []
            for f in range(f0, f1):
                op = chr(up.f_op[f])
                if op == 'S':
                    o, l = int(up.f_val[f]), int(up.f_aux[f])
                    got.append([op, int(up.f_pos[f]), base64.b64encode(up.payload[o:o + l].tobytes()).decode()])
                else:
                    got.append([op, int(up.f_pos[f]), int(up.f_val[f])])
            want = [[f[0], f[1], f[4]] if f[0] == 'X' else f for f in (b['features'] or [])]
            assert got == want, (c['tag'], i, b['name'])
            if not b['flags'] & 4:
                assert int(up.mapq[i]) == b['mapq'], (c['tag'], i)
            rl = b['read_length']
            if b['cf'] & 1 and b['quals'] is not None:
                o = int(up.read_qs_off[i])
                assert streams[o:o + rl].tobytes() == base64.b64decode(b['quals']), (c['tag'], i)
            if b['flags'] & 4 and b['bases'] is not None:
                o = int(up.read_ba_off[i])
                assert streams[o:o + rl].tobytes() == base64.b64decode(b['bases']), (c['tag'], i)


def _host_fields(c, rd):
    sc = c['scalars']
    return dict(flag=rd['flag'], pos=rd['pos'], read_off=rd['read_off'], slice_start=[c['slice']['alignment_start']],
                multi_ref=[1 if c['slice']['sequence_id'] == -2 else 0], ap_delta=bool(c['ap_delta']),
                names_included=bool(c['read_names_included']), ref_id=[s['seq_id'] for s in sc],
                read_group=[s['read_group'] for s in sc], names=[s['name'] for s in sc], mate_ref_id=[s['mate_seq_id'] for s in sc],
                mate_pos=[s['mate_start'] for s in sc], tlen=[s['tlen'] for s in sc], nf=[s['nf'] for s in sc])


def test_host_series_match_jar(R, G, gpu_ctx):
    """BF CF RI RL AP RG RN MF NS NP TS NF TL from the records' scalars (mate linking done by the jar's setMateInfo): BF / MF
    derived from the SAM flags on the device, CF once handed over whole and once composed from its parts"""
    g, ref, conts = G
    for c, rd, series in conts:
        sc = c['scalars']
        kw = _host_fields(c, rd)
        full = R.host_series_pack(cf=[s['cf'] for s in sc], ctx=gpu_ctx, **kw)
        enc = R.features_encode(rd, ref, None, ctx=gpu_ctx)
        parts = R.host_series_pack(mate_cf=[s['cf'] & 6 for s in sc], force_qs=enc.force_qs, rflags=rd['rflags'], ctx=gpu_ctx, **kw)
        for nm in R.HOST_SERIES:
            want = series.get(CID[nm], b'')
            assert full.series(0, nm) == want, (c['tag'], nm)
            assert parts.series(0, nm) == want, (c['tag'], nm, 'cf composed')


def test_host_series_multi_container_and_options(R, G, gpu_ctx):
    """the three slices as one batch (per-container slice start / multi-reference flag), against the oracle's restatement too"""
    from oracle import records as orec
    g, ref, conts = G
    same = [x for x in conts if x[0]['ap_delta'] == conts[0][0]['ap_delta'] and x[0]['read_names_included'] == conts[0][0]['read_names_included']]
    assert len(same) >= 2
    cat = {k: [] for k in ('flag', 'pos', 'ref_id', 'read_group', 'names', 'mate_ref_id', 'mate_pos', 'tlen', 'nf', 'cf')}
    ro, cont_off, ss, mr = [np.zeros(1, np.int64)], [0], [], []
    for c, rd, _ in same:
        kw = _host_fields(c, rd)
        for k in ('flag', 'pos', 'ref_id', 'read_group', 'names', 'mate_ref_id', 'mate_pos', 'tlen', 'nf'):
            cat[k] += list(kw[k])
        cat['cf'] += [s['cf'] for s in c['scalars']]
        ro.append(rd['read_off'][1:] + ro[-1][-1]); cont_off.append(cont_off[-1] + len(rd['pos']))
        ss.append(c['slice']['alignment_start']); mr.append(1 if c['slice']['sequence_id'] == -2 else 0)
    got = R.host_series_pack(cat['flag'], cat['pos'], np.concatenate(ro), cont_off=cont_off, slice_start=ss, multi_ref=mr,
                             ap_delta=bool(same[0][0]['ap_delta']), names_included=bool(same[0][0]['read_names_included']),
                             cf=cat['cf'], ref_id=cat['ref_id'], read_group=cat['read_group'], names=cat['names'],
                             mate_ref_id=cat['mate_ref_id'], mate_pos=cat['mate_pos'], tlen=cat['tlen'], nf=cat['nf'], ctx=gpu_ctx)
    for ci, (c, rd, series) in enumerate(same):
        for nm in R.HOST_SERIES:
            assert got.series(ci, nm) == series.get(CID[nm], b''), (c['tag'], nm)
    # names for every record (CompressionHeader.readNamesIncluded) and absolute AP: against the oracle's restatement of the Writer
    c, rd, _ = conts[0]
    sc = [dict(flags=s['flags'], cf=s['cf'], seq_id=s['seq_id'], read_length=s['read_length'], start=int(p), read_group=s['read_group'],
               name=s['name'], mate_flags=s['mate_flags'], mate_seq_id=s['mate_seq_id'], mate_start=s['mate_start'], tlen=s['tlen'],
               nf=s['nf'], tl=0) for s, p in zip(c['scalars'], rd['pos'])]
    for apd, named in ((False, True), (True, True), (False, False)):
        want = orec.write_host_streams(sc, apd, c['slice']['sequence_id'] == -2, c['slice']['alignment_start'], capture_names=named)
        kw = _host_fields(c, rd)
        kw.update(ap_delta=apd, names_included=named)
        got = R.host_series_pack(cf=[s['cf'] for s in c['scalars']], ctx=gpu_ctx, **kw)
        for nm in R.HOST_SERIES:
            assert got.series(0, nm) == want[nm], (apd, named, nm)


def _cat(conts):
    cat = {}
    for k in ('bases', 'quals', 'cigar', 'pos', 'flag', 'mapq', 'rflags', 'ref_id'):
        cat[k] = np.concatenate([rd[k] for _, rd, _ in conts])
    ro, co, cont_off = [np.zeros(1, np.int64)], [np.zeros(1, np.int64)], [0]
    for _, rd, _ in conts:
        ro.append(rd['read_off'][1:] + ro[-1][-1]); co.append(rd['cigar_off'][1:] + co[-1][-1]); cont_off.append(cont_off[-1] + len(rd['pos']))
    cat['read_off'] = np.concatenate(ro); cat['cigar_off'] = np.concatenate(co)
    return cat, np.array(cont_off, np.uint64)


def test_slice_headers_and_digests_match_jar(R, G, gpu_ctx):
    """nofRecords, bases, sequenceId, alignment start / span and the six content-digest tags of the three slices the
    jar's ContainerFactory.buildSlice built -- as one batch, and one by one"""
    g, ref, conts = G
    cat, cont_off = _cat(conts)
    enc = R.features_encode(cat, ref, None, cont_off=cont_off, ctx=gpu_ctx)
    got = R.slice_headers(cat['bases'], cat['read_off'], cat['pos'], cat['flag'], enc.f_op, enc.f_val, enc.f_aux, enc.feat_off,
                          scores=cat['quals'], rflags=cat['rflags'], ref_id=cat['ref_id'], cont_off=cont_off, ref=ref, ctx=gpu_ctx)
    for h, (c, rd, _) in zip(got, conts):
        assert h['ref_md5'] == (bytes(16) if c['ref_md5'] is None else base64.b64decode(c['ref_md5'])), c['tag']      # Slice.setRefMD5
        for k in ('sequence_id', 'alignment_start', 'alignment_span', 'n_records', 'bases'):
            assert h[k] == c['slice'][k], (c['tag'], k, h[k], c['slice'][k])
        assert h['tags'] == {t: base64.b64decode(v) for t, v in c['slice_tags']}, c['tag']
    for c, rd, _ in conts:
        e1 = R.features_encode(rd, ref, None, ctx=gpu_ctx)
        h = R.slice_headers(rd['bases'], rd['read_off'], rd['pos'], rd['flag'], e1.f_op, e1.f_val, e1.f_aux, e1.feat_off,
                            scores=rd['quals'], rflags=rd['rflags'], ref_id=rd['ref_id'], ctx=gpu_ctx)[0]
        assert h['tags'] == {t: base64.b64decode(v) for t, v in c['slice_tags']} and h['alignment_span'] == c['slice']['alignment_span']


@pytest.mark.parametrize('spec', [None, '*8', 'm5', 'N40-D8', 'R8-N40'])
def test_slice_headers_parity_with_oracle(R, gpu_ctx, spec):
    """bigger slices, ragged read lengths (every byte alignment), quality models: explicit score arrays, and score arrays taken
    from the device result of the feature call (forced arrays in QS, sparse Q features otherwise)"""
    import ctypes
    from oracle import records as orec
    from cramtools_b200 import synth
    from tests.test_records_gpu import _concat_reads
    nc = 4
    ref = synth.reference(40000, 3)
    parts = [synth.reads(ref, 2500, 9, read_len=150, sub_rate=0.01, indel_frac=0.15, clip_frac=0.08), synth.reads(ref, 1500, 10, read_len=101, sub_rate=0.01),
             synth.reads(ref, 700, 11, read_len=37, indel_frac=0.2, clip_frac=0.1), synth.reads(ref, 1300, 12, read_len=75, indel_frac=0.1)]
    rd = _concat_reads(parts)
    n = len(rd['pos'])
    cont_off = np.linspace(0, n, nc + 1).astype(np.uint64)
    ref_id = np.zeros(n, np.int32)
    ref_id[int(cont_off[1]):int(cont_off[2])] = 2                      # slice 1: another single reference
    ref_id[int(cont_off[2]) + 7] = 1                                    # slice 2: multi-reference
    a3, b3 = int(cont_off[3]), int(cont_off[4])
    pos = rd['pos'].copy(); flag = rd['flag'].copy()
    pos[a3:b3] = 0; flag[a3:b3] |= 4; ref_id[a3:b3] = -1               # slice 3: nothing placed
    flag[5] |= 4                                                         # unmapped flag but placed: end counts as 0
    per_base = spec in ('N40-D8', 'R8-N40')
    rd = dict(rd, pos=pos, flag=flag, ref_id=ref_id)
    rfl = np.zeros(n, np.uint8); rfl[::97] = 1                           # some records without quality scores
    rd['rflags'] = rfl
    enc, D = R.features_encode_resident(rd, ref, spec, cont_off=cont_off, ctx=gpu_ctx)
    want, all_scores = [], []
    for c in range(nc):
        a, b = int(cont_off[c]), int(cont_off[c + 1])
        sub = dict(bases=rd['bases'][rd['read_off'][a]:rd['read_off'][b]], quals=rd['quals'][rd['read_off'][a]:rd['read_off'][b]],
                   read_off=rd['read_off'][a:b + 1] - rd['read_off'][a], cigar=rd['cigar'][rd['cigar_off'][a]:rd['cigar_off'][b]],
                   cigar_off=rd['cigar_off'][a:b + 1] - rd['cigar_off'][a], pos=rd['pos'][a:b], flag=rd['flag'][a:b], mapq=rd['mapq'][a:b],
                   rflags=rfl[a:b], ref_id=ref_id[a:b])
        ex = orec.extract(sub, ref)
        force, scores = orec.quality(sub, ref, spec)
        all_scores.append(scores)
        want.append(orec.slice_header(sub, ex, scores))
    scores = np.concatenate(all_scores)
    keys = ('sequence_id', 'alignment_start', 'alignment_span', 'n_records', 'bases', 'bd', 'sd')
    # every slice with its own reference sequence inside one buffer (ref_off / ref_len), the second one cut short so that the
    # span reaches past its end
    ref3 = np.concatenate([ref, ref[::-1].copy(), ref])
    roff = np.array([0, ref.size, 2 * ref.size, 0], np.uint64); rlen = np.array([ref.size, ref.size, ref.size, ref.size], np.uint64)
    got = R.slice_headers(rd['bases'], rd['read_off'], pos, flag, enc.f_op, enc.f_val, enc.f_aux, enc.feat_off, scores=scores, rflags=rfl,
                          ref_id=ref_id, cont_off=cont_off, ref=ref3, ref_off=roff, ref_len=rlen, ctx=gpu_ctx)
    for c in range(nc):
        assert {k: got[c][k] for k in keys} == {k: want[c][k] for k in keys}, (spec, c)
        seq = ref3[int(roff[c]):int(roff[c]) + int(rlen[c])]
        assert got[c]['ref_md5'] == orec.slice_ref_md5(seq, want[c]['sequence_id'], want[c]['alignment_start'], want[c]['alignment_span']), (spec, c)
    short = int(want[0]['alignment_start']) + int(want[0]['alignment_span']) // 2
    got_s = R.slice_headers(rd['bases'], rd['read_off'], pos, flag, enc.f_op, enc.f_val, enc.f_aux, enc.feat_off, scores=scores, rflags=rfl,
                            ref_id=ref_id, cont_off=cont_off, ref=ref3, ref_off=roff, ref_len=np.array([short, ref.size, ref.size, ref.size], np.uint64),
                            ctx=gpu_ctx)
    assert got_s[0]['ref_md5'] == orec.slice_ref_md5(ref3[:short], 0, want[0]['alignment_start'], want[0]['alignment_span'])
    assert [h['sequence_id'] for h in got] == [0, 2, -2, -1]
    # the same from the device-resident result of the feature call, everything else staged from the host
    got2 = R.slice_headers(rd['bases'], rd['read_off'], pos, flag, enc.f_op, enc.f_val, enc.f_aux, enc.feat_off, scores=None, rflags=rfl,
                           ref_id=ref_id, cont_off=cont_off, f_pos=enc.f_pos, force_qs=int(D.force_qs), read_qs_off=int(D.read_qs_off),
                           streams=int(D.streams), ctx=gpu_ctx)
    for c in range(nc):
        # (a record with the unmapped flag AND a position keeps, under a per-base model, sparse scores that Writer.write never
        # emits as features [Writer.java:180-186]: the feature arrays cannot give its SD term -- slice 0 holds that record)
        ks = [k for k in keys if not (per_base and c == 0 and k == 'sd')]
        assert {k: got2[c][k] for k in ks} == {k: want[c][k] for k in ks}, (spec, c, 'resident')


def test_mate_linking_matches_jar_and_feeds_the_writer(R, G, gpu_ctx):
    """cramgpu_mate_link against what the jar's CramSerilization.setMateInfo left in the records, then straight into
    cramgpu_host_series_pack: CF MF NS NP TS NF come out byte for byte as the jar's Writer wrote them -- with no mate decision
    taken on the host"""
    g, ref, conts = G
    for c, rd, series in conts:
        sc = c['scalars']
        enc = R.features_encode(rd, ref, None, ctx=gpu_ctx)
        cf, nf = R.mate_link(rd['flag'], rd['names'], rd['pos'], [s['mate_seq_id'] for s in sc], [s['mate_start'] for s in sc],
                             [s['tlen'] for s in sc], rd['read_off'], enc.f_op, enc.f_val, enc.f_aux, enc.feat_off, ref_id=rd['ref_id'],
                             coordinate_sorted=c['coordinate_sorted'], ctx=gpu_ctx)
        assert [int(x) for x in cf] == [s['cf'] & 6 for s in sc], c['tag']
        assert [int(x) for x in nf] == [s['nf'] for s in sc], c['tag']
        kw = _host_fields(c, rd)
        kw['nf'] = nf
        got = R.host_series_pack(mate_cf=cf, force_qs=enc.force_qs, rflags=rd['rflags'], ctx=gpu_ctx, **kw)
        for nm in R.HOST_SERIES:
            assert got.series(0, nm) == series.get(CID[nm], b''), (c['tag'], nm)


def test_mate_linking_parity_with_oracle_random_pairs(R, gpu_ctx):
    """thousands of templates over several containers: proper pairs, pairs split across containers, wrong TLEN / mate position /
    strand, three and four records of one name, secondary pairs sharing a primary's name, unpaired records, names that differ
    only in the last byte or in length"""
    from oracle import records as orec
    from cramtools_b200 import synth
    rng = np.random.default_rng(33)
    ref = synth.reference(60000, 5)
    rd = synth.reads(ref, 9000, 6, read_len=100, indel_frac=0.2, clip_frac=0.1)
    n = len(rd['pos'])
    ex = orec.extract(rd, ref)
    ends = orec.alignment_ends(rd, ex)
    flag = rd['flag'].astype(np.int64) & 0x10
    names = [b''] * n
    mate_ref = np.full(n, -1, np.int64); mate_pos = np.zeros(n, np.int64); tlen = np.zeros(n, np.int64)
    order = rng.permutation(n)
    i = 0
    t = 0
    while i + 4 <= n:
        kind = rng.integers(0, 10)
        nm = ('t%d' % t).encode() if kind != 9 else ('t%d' % (t // 2 * 2)).encode() + b'x' * int(rng.integers(0, 2))
        t += 1
        if kind == 0:                                   # unpaired
            names[order[i]] = nm; i += 1
            continue
        k = 3 if kind == 1 else 4 if kind == 2 else 2
        members = sorted(order[i:i + k]); i += k
        sec = 0x100 if kind == 3 else 0
        for j, m in enumerate(members):
            names[m] = nm
            flag[m] |= 1 | sec | (0x40 if j == 0 else 0x80)
        a, b = members[0], members[1]
        for x, y in ((a, b), (b, a)):
            mate_ref[x] = 0; mate_pos[x] = rd['pos'][y]
            if flag[y] & 0x10:
                flag[x] |= 0x20
        pa = ends[a] if flag[a] & 0x10 else rd['pos'][a]; pb = ends[b] if flag[b] & 0x10 else rd['pos'][b]
        tl = pb - pa + (1 if pb >= pa else -1)
        tlen[a], tlen[b] = tl, -tl
        if kind == 4: tlen[b] += 1                      # second TLEN off
        if kind == 5: mate_pos[a] += 1                  # mate position off
        if kind == 6: flag[a] ^= 0x20                   # mate strand off
        if kind == 7: flag[b] = (flag[b] & ~0x80) | 0x40      # two firsts
    cont_off = np.array([0, 2500, 2501, 6000, n], np.uint64)
    enc = R.features_encode(dict(rd, flag=(flag & 0xFFFF).astype(np.uint16)), ref, None, cont_off=cont_off, ctx=gpu_ctx)
    cf, nf = R.mate_link(flag, names, rd['pos'], mate_ref, mate_pos, tlen, rd['read_off'], enc.f_op, enc.f_val, enc.f_aux, enc.feat_off,
                         cont_off=cont_off, ctx=gpu_ctx)
    linked = 0
    for c in range(4):
        a, b = int(cont_off[c]), int(cont_off[c + 1])
        wcf, wnf = orec.set_mate_info(flag[a:b], names[a:b], np.zeros(b - a), rd['pos'][a:b], mate_ref[a:b], mate_pos[a:b], tlen[a:b], ends[a:b])
        assert (cf[a:b] == wcf).all() and (nf[a:b] == wnf).all(), c
        linked += int((wcf == 4).sum())
    assert linked > 500
    cf0, nf0 = R.mate_link(flag, names, rd['pos'], mate_ref, mate_pos, tlen, rd['read_off'], enc.f_op, enc.f_val, enc.f_aux, enc.feat_off,
                           cont_off=cont_off, coordinate_sorted=False, ctx=gpu_ctx)
    assert (cf0 == 2).all() and (nf0 == -1).all()
