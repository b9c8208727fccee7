"""CPU: rows W and P of the oracle against what the reference jar itself wrote and read (tests/golden/streams_jvm.json:
ContainerFactory.buildSlice -> Writer.write, DataReaderFactory.buildReader -> CramRecordReader.read, run in the mini-JVM
by oracle/make_golden_streams.py; only ExternalCompressor.compress was stubbed to keep the raw series)."""
import base64

import numpy as np
import pytest

from oracle import records as orec
from tests import golden_streams as gs

CID = orec.CONTENT_ID


@pytest.fixture(scope='module')
def G():
    return gs.load()


def test_series_to_compressor_map_matches_survey(G):
    """row R0: which external compressor CompressionHeaderFactory.build gives each series [:84-113]"""
    g, _, conts = G
    want = {'rans1': ['BA', 'QS', 'BF', 'CF', 'NS', 'RG', 'RL', 'TS'], 'rans0': ['AP', 'RI'],
            'gzip': ['BS', 'DL', 'FC', 'FN', 'FP', 'HC', 'IN', 'MF', 'MQ', 'NF', 'NP', 'PD', 'RN', 'RS', 'SC', 'TC', 'TL', 'TN']}
    for c, _, _ in conts:
        comp = c['compressors']
        for nm in want['rans1']:
            assert (comp[str(CID[nm])]['method'], comp[str(CID[nm])]['order']) == ('RANS', 'ONE'), nm
        for nm in want['rans0']:
            assert (comp[str(CID[nm])]['method'], comp[str(CID[nm])]['order']) == ('RANS', 'ZERO'), nm
        for nm in want['gzip']:
            assert comp[str(CID[nm])]['method'] == 'GZIP', nm
        assert base64.b64decode(c['core']) == b''            # every series is external: nothing in the core block


def _scalars(c, rd):
    """the record fields Writer.write reads, as the jar left them after createCramRecord + setMateInfo"""
    out = []
    for s, pos in zip(c['scalars'], rd['pos']):
        out.append(dict(flags=s['flags'], cf=s['cf'], seq_id=s['seq_id'], read_length=s['read_length'], start=int(pos),
                        read_group=s['read_group'], name=s['name'], mate_flags=s['mate_flags'], mate_seq_id=s['mate_seq_id'],
                        mate_start=s['mate_start'], tlen=s['tlen'], nf=s['nf'], tl=0))
    return out


def test_host_series_match_jar(G):
    """BF CF RI RL AP RG RN MF NS NP TS NF TL byte for byte"""
    g, _, conts = G
    for c, rd, series in conts:
        multi = c['slice']['sequence_id'] == -2
        got = orec.write_host_streams(_scalars(c, rd), c['ap_delta'], multi, c['slice']['alignment_start'],
                                      capture_names=bool(c['read_names_included']))
        for nm in orec.HOST_STREAMS:
            assert got[nm] == series.get(CID[nm], b''), (c['tag'], nm)


def test_host_series_c_restatement_matches_jar(G):
    """the C version (what bench.py's CPU baseline runs): BF / MF derived from the SAM flags inside"""
    g, _, conts = G
    for c, rd, series in conts:
        sc = c['scalars']
        got = orec.write_host_streams_c(rd['flag'], [s['cf'] for s in sc], [s['seq_id'] for s in sc], rd['read_off'], rd['pos'],
                                        [s['read_group'] for s in sc], [s['name'] for s in sc], [s['mate_seq_id'] for s in sc],
                                        [s['mate_start'] for s in sc], [s['tlen'] for s in sc], [s['nf'] for s in sc], None,
                                        c['ap_delta'], c['slice']['sequence_id'] == -2, c['slice']['alignment_start'],
                                        bool(c['read_names_included']))
        for nm in orec.HOST_STREAMS:
            assert got[nm] == series.get(CID[nm], b''), (c['tag'], nm)


def test_bit_flags_match_jar(G):
    """row F2, scalar half: BF and MF from the SAM flag"""
    g, _, conts = G
    for c, rd, _ in conts:
        for s, fl in zip(c['scalars'], rd['flag']):
            assert orec.cram_bit_flags(fl) == s['flags'], (c['tag'], s['name'], int(fl))
            assert orec.mate_bit_flags(fl) == s['mate_flags'], (c['tag'], s['name'], int(fl))


def test_feature_series_match_jar(G):
    """FN FC FP BS DL BA SC QS RS HC PD MQ of the C oracle's Writer restatement, byte for byte"""
    g, ref, conts = G
    for c, rd, series in conts:
        ex = orec.extract(rd, ref)
        force, scores = orec.quality(rd, ref, None)
        st = orec.write_streams(rd, ex, force, scores)
        assert ex.matrix == base64.b64decode(c['subst_matrix']), c['tag']
        for nm in orec.STREAMS:
            assert st[nm] == series.get(CID[nm], b''), (c['tag'], nm)


def test_reader_matches_jar(G):
    """row P: read_streams over the jar's own series gives what the jar's CramRecordReader read"""
    g, ref, conts = G
    for c, rd, series in conts:
        st = {nm: series.get(CID[nm], b'') for nm in orec.STREAMS}
        back = c['read_back']
        rl = [b['read_length'] for b in back]
        # the flags the host part of the reader hands over: CF bit 0 (stored score array), BF 0x4 (unmapped), CF bit 3 (unknown bases)
        rfl = np.array([(b['cf'] & 1) | (2 if b['flags'] & 4 else 0) | (4 if b['cf'] & 8 else 0) for b in back], np.uint8)
        rows, mapq, qs_off, ba_off = orec.read_streams(st, rl, rfl)
        for i, b in enumerate(back):
            want = [[f[0], f[1], f[4]] if f[0] == 'X' else f for f in (b['features'] or [])]
            got = [[f[0], f[1], base64.b64encode(f[2]).decode()] if f[0] == 'S' else f for f in rows[i]]
            assert got == want, (c['tag'], i, b['name'])
            if not b['flags'] & 4:
                assert mapq[i] == b['mapq'], (c['tag'], i)
            if b['quals'] is not None and b['cf'] & 1:
                assert st['QS'][qs_off[i]:qs_off[i] + rl[i]] == base64.b64decode(b['quals']), (c['tag'], i)
            if b['flags'] & 4 and b['bases'] is not None:
                assert st['BA'][ba_off[i]:ba_off[i] + rl[i]] == base64.b64decode(b['bases']), (c['tag'], i)
        assert set(c['bytes_left'].values()) == {0}            # and the jar's reader used up every byte of every series


def test_slice_header_and_digests_match_jar(G):
    """what ContainerFactory.buildSlice derives from the records -- nofRecords, bases, sequenceId, alignment
    start / span, and the six content-digest tags -- against the slices the jar built"""
    import hashlib
    import zlib
    g, ref, conts = G
    assert orec.SHA1_EMPTY == hashlib.sha1(b'').digest() and orec.SHA512_EMPTY == hashlib.sha512(b'').digest()
    assert orec.lib().rec_oracle_crc32(orec._p(np.frombuffer(b'123456789', np.uint8).copy(), orec.u8p), 9) == zlib.crc32(b'123456789') == 0xCBF43926
    seen = set()
    for c, rd, series in conts:
        ex = orec.extract(rd, ref)
        force, scores = orec.quality(rd, ref, None)
        h = orec.slice_header(rd, ex, scores)
        for k in ('sequence_id', 'alignment_start', 'alignment_span', 'n_records', 'bases'):
            assert h[k] == c['slice'][k], (c['tag'], k, h[k], c['slice'][k])
        want = {t: base64.b64decode(v) for t, v in c['slice_tags']}
        assert h['tags'] == want, c['tag']
        seen.add(h['sequence_id'])
    assert -2 in seen and 0 in seen          # a multi-reference slice and a single-reference one


def test_mate_linking_matches_jar(G):
    """CramSerilization.setMateInfo -- detached / mate-downstream bits of CF and recordsToNextFragment of every
    record, against what the jar's own setMateInfo left in the records (proper pairs, a pair with a wrong TLEN, a mate outside
    the container, a secondary pair, a triplet, unsorted header)"""
    g, ref, conts = G
    linked = 0
    for c, rd, series in conts:
        sc = c['scalars']
        ex = orec.extract(rd, ref)
        ends = orec.alignment_ends(rd, ex)
        cf, nf = orec.set_mate_info(rd['flag'], rd['names'], rd['ref_id'], rd['pos'], [s['mate_seq_id'] for s in sc],
                                    [s['mate_start'] for s in sc], [s['tlen'] for s in sc], ends, coordinate_sorted=c['coordinate_sorted'])
        assert [int(x) for x in cf] == [s['cf'] & 6 for s in sc], c['tag']
        assert [int(x) for x in nf] == [s['nf'] for s in sc], c['tag']
        linked += int((cf == 4).sum())
    assert linked == 10


def _feat_triples(feats):
    """golden feature json -> (op, pos, length) as cigar_from_features takes them"""
    out = []
    for f in feats or []:
        op = f[0]
        ln = len(base64.b64decode(f[2])) if op in ('S', 'I') else f[2] if op in ('D', 'N', 'H', 'P') else 1
        out.append((op, f[1], ln))
    return out


def test_cigar_rebuild_and_bam_records_match_jar(G):
    """Cram2SamRecordFactory.getCigar2 over the features the jar's reader read back, and the BAM record cramtools'
    BAMRecordView lays out for every record (bin, 4-bit bases, missing qualities as 0xFF, RG tag)"""
    g, ref, conts = G
    ops = set()
    for c, rd, series in conts:
        for i, (r, rb, sc) in enumerate(zip(c['reads'], c['read_back'], c['scalars'])):
            cig = orec.cigar_from_features(_feat_triples(rb['features']), rb['read_length'])
            txt = ''.join('%d%s' % e for e in cig)
            assert txt == rb['cigar2'], (c['tag'], i, txt, rb['cigar2'])
            ops.update(o for _, o in cig)
            unm = bool(r['flag'] & 4)
            paired = bool(r['flag'] & 1)
            rec = orec.bam_record(r['ref_id'], r['pos'], r['name'].encode('latin-1'), r['mapq'], r['flag'], [] if unm else cig,
                                  sc['mate_seq_id'] if paired else -1, sc['mate_start'] if paired else 0, sc['tlen'],
                                  base64.b64decode(r['bases']) if r['bases'] is not None else b'',
                                  base64.b64decode(r['quals']) if r['quals'] is not None else b'', base64.b64decode(rb['bam_tags']))
            assert rec == base64.b64decode(rb['bam']), (c['tag'], i)
            back = orec.bam_parse(rec)
            assert back['pos'] == r['pos'] and back['flag'] == r['flag'] and back['bin'] == rb['bam_bin']
            if r['bases'] is not None:
                assert back['bases'] == base64.b64decode(r['bases']).upper().replace(b'.', b'N')
    assert ops >= set('MIDNSHP')
    with pytest.raises(ValueError):
        orec.bam_record(0, 1, b'x', 0, 0, [(1, 'M')], -1, 0, 0, b'J', b'#')


def test_block_layer_matches_jar(G):
    """SURVEY 8f-5, the block layer: the external blocks of the rANS-coded series as the jar's Block.write(3, ...) wrote them"""
    from oracle import rans as orans
    g, ref, conts = G
    n = 0
    for c, rd, series in conts:
        for cid, blk in c['blocks_v3'].items():
            raw = series[int(cid)]
            comp = orans.compress(raw, blk['order'])
            got = orec.cram_block_v3(orec.BLOCK_METHOD['RANS'], orec.BLOCK_CONTENT['EXTERNAL'], int(cid), len(raw), comp)
            assert got == base64.b64decode(blk['block']), (c['tag'], cid)
            n += 1
    assert n >= 20


def test_slice_ref_md5_matches_jar(G):
    """Slice.refMD5 as the jar's Slice.setRefMD5 computed it for the golden slices (the span reaches past the reference end in
    two of them: the min() of [Slice.java:154] is exercised); not set for the multi-reference slice"""
    g, ref, conts = G
    seen = 0
    for c, rd, series in conts:
        s = c['slice']
        got = orec.slice_ref_md5(ref, s['sequence_id'], s['alignment_start'], s['alignment_span'])
        if c['ref_md5'] is None:
            assert s['sequence_id'] == -2 and got == bytes(16)
        else:
            assert got == base64.b64decode(c['ref_md5']), c['tag']
            seen += 1
    assert seen >= 2


def test_container_bytes_match_jar(G):
    """SURVEY 8f-5, the container part: ContainerIO.writeContainer of the three golden containers as the jar wrote them in the
    mini-JVM -- container header with CRC32, compression header block, slice header block with the digest tags, core and
    external blocks in the jar's HashMap order; LTF8 of a 6e9 record counter; the multi-reference slice"""
    g, ref, conts = G
    assert orec.ltf8(0) == b'\x00' and orec.ltf8(127) == b'\x7f' and orec.ltf8(128) == b'\x80\x80' and len(orec.ltf8(1 << 56)) == 9
    for c, rd, series in conts:
        hdr, grc, chb, ids, tags, core, ext = gs.container_parts(c, series)
        cv = c['container_v3']
        shc = orec.slice_header_content_v3(hdr['sequence_id'], hdr['alignment_start'], hdr['alignment_span'], hdr['n_records'], grc, ids,
                                           hdr['ref_md5'], tags)
        assert shc == base64.b64decode(cv['slice_header_content']), c['tag']
        whole = orec.container_v3(hdr, grc, chb, ids, hdr['ref_md5'], tags, (0, len(core), core), ext)
        want = base64.b64decode(cv['bytes'])
        assert whole == want, c['tag']
        assert int.from_bytes(want[:4], 'little') == cv['byte_size'] and cv['block_count'] == 3 + len(ids)
        assert ids == sorted(ids)          # HashMap<Integer, Block> with fewer keys than buckets iterates in key order


def test_pairing_restoration_matches_jar(G):
    """the inverse of setMateInfo: CramNormalizer.normalize's pairing passes run in the mini-JVM on the records the jar's reader
    read back -- links, mate fields and template sizes of the linked pairs, untouched detached records"""
    g, ref, conts = G
    links = 0
    for c, rd, series in conts:
        a = gs.read_back_arrays(c)
        got = orec.restore_mates(a['flag'], a['cf'], a['nf'], a['ref_id'], a['pos'], a['mate_flags'], a['mate_ref_id'], a['mate_pos'], a['tlen'], gs.read_back_ends(a))
        want = c['normalized_mates']
        for k, key in (('mate_flags', 'mate_flags'), ('mate_seq_id', 'mate_seq_id'), ('mate_pos', 'mate_start'), ('tlen', 'tlen'), ('nf', 'nf'),
                       ('next', 'next'), ('previous', 'previous'), ('pos', 'alignment_start')):
            assert [int(x) for x in got[k]] == [w[key] for w in want], (c['tag'], k)
        links += sum(1 for w in want if w['next'] >= 0)
    assert links == 10
