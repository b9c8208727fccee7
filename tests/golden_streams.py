"""helpers: load tests/golden/streams_jvm.json (per-series bytes of three slices written by the reference jar's own
Writer / ContainerFactory.buildSlice in the mini-JVM, and the records its CramRecordReader read back) into the SoA layout"""
import base64
import json
import os

import numpy as np

from cramtools_b200 import synth
from tests.golden_records import _parse_cigar

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def load():
    with open(os.path.join(ROOT, 'tests', 'golden', 'streams_jvm.json')) as f:
        g = json.load(f)
    ref = np.frombuffer(base64.b64decode(g['ref']), np.uint8)
    out = []
    for c in g['containers']:
        reads = c['reads']
        n = len(reads)
        bases, quals, cig = [], [], []
        read_off = np.zeros(n + 1, np.int64)
        cigar_off = np.zeros(n + 1, np.int64)
        rflags = np.zeros(n, np.uint8)
        for i, r in enumerate(reads):
            b = base64.b64decode(r['bases']) if r['bases'] is not None else b''
            q = base64.b64decode(r['quals']) if r['quals'] is not None else b'\xff' * len(b)
            if r['quals'] is None:
                rflags[i] |= 1
            if r['bases'] is None:
                rflags[i] |= 2
            bases.append(b); quals.append(q)
            read_off[i + 1] = read_off[i] + len(b)
            cg = synth.pack_cigar(_parse_cigar(r['cigar']))
            cig.append(cg)
            cigar_off[i + 1] = cigar_off[i] + len(cg)
        rd = dict(bases=np.frombuffer(b''.join(bases), np.uint8).copy(), quals=np.frombuffer(b''.join(quals), np.uint8).copy(),
                  read_off=read_off, cigar=np.concatenate(cig) if cig else np.zeros(0, np.uint32), cigar_off=cigar_off,
                  pos=np.array([r['pos'] for r in reads], np.int32), flag=np.array([r['flag'] for r in reads], np.uint16),
                  mapq=np.array([r['mapq'] for r in reads], np.uint8), ref_id=np.array([r['ref_id'] for r in reads], np.int32),
                  rflags=rflags, names=[r['name'] for r in reads])
        series = {int(k): base64.b64decode(v) for k, v in c['series'].items()}
        out.append((c, rd, series))
    return g, ref, out


def container_parts(c, series):
    """what a golden container is made of, for the container restatement / assembly: (slice header dict, global record counter,
    compression header bytes, content ids in the jar's map order, tags in the slice's order, core bytes, external blocks
    [(content id, method ordinal, raw length, bytes)] -- the bytes the stubbed compressor returned: the raw series)"""
    from oracle import records as orec
    cv = c['container_v3']
    s = c['slice']
    md5 = base64.b64decode(c['ref_md5']) if c['ref_md5'] else bytes(16)
    tags = [(t[0], base64.b64decode(t[1])) for t in c['slice_tags']]
    td = dict(tags)
    hdr = dict(sequence_id=s['sequence_id'], alignment_start=s['alignment_start'], alignment_span=max(s['alignment_span'], 0),
               n_records=s['n_records'], bases=s['bases'], ref_md5=md5,
               bd=int.from_bytes(td['BD'], 'little'), sd=int.from_bytes(td['SD'], 'little'))
    ids = cv['content_ids']
    ext = [(cid, orec.BLOCK_METHOD[c['compressors'][str(cid)]['method']], len(series[cid]), series[cid]) for cid in ids]
    return hdr, cv['global_record_counter'], base64.b64decode(cv['compression_header']), ids, tags, base64.b64decode(c['core']), ext


def read_back_arrays(c):
    """the records of a golden container as the jar's reader left them (ContainerParser's alignment starts included), as
    arrays for the pairing restoration: flag, cf, nf, ref_id, pos, mate_ref_id, mate_pos, tlen, read_len, and the read
    features that getAlignmentEnd looks at as (f_op, f_val, f_aux, feat_off)"""
    rb = c['read_back']
    n = len(rb)
    f_op, f_val, f_aux, feat_off = [], [], [], [0]
    for r in rb:
        for f in r['features'] or []:
            op = f[0]
            ln = len(base64.b64decode(f[2])) if op in ('S', 'I') else f[2] if op in ('D', 'N', 'H', 'P') else 0
            f_op.append(ord(op)); f_val.append(ln if op in ('D', 'N', 'H', 'P') else 0); f_aux.append(ln if op in ('S', 'I') else 0)
        feat_off.append(len(f_op))
    return dict(flag=np.array([r['flags'] for r in rb], np.uint16), cf=np.array([r['cf'] for r in rb], np.uint8),
                mate_flags=np.array([r['mate_flags'] for r in rb], np.uint8), nf=np.array([r['nf'] for r in rb], np.int32), ref_id=np.array([r['seq_id'] for r in rb], np.int32),
                pos=np.array([r['alignment_start'] for r in rb], np.int32), mate_ref_id=np.array([r['mate_seq_id'] for r in rb], np.int32),
                mate_pos=np.array([r['mate_start'] for r in rb], np.int32), tlen=np.array([r['tlen'] for r in rb], np.int32),
                read_off=np.concatenate([[0], np.cumsum([r['read_length'] for r in rb])]).astype(np.uint64),
                f_op=np.array(f_op, np.uint8), f_val=np.array(f_val, np.int32), f_aux=np.array(f_aux, np.int32),
                feat_off=np.array(feat_off, np.uint64), n=n)


def read_back_ends(a):
    """CramCompressionRecord.getAlignmentEnd per record of read_back_arrays [CramCompressionRecord.java:160-194]"""
    ends = []
    for r in range(a['n']):
        if a['flag'][r] & 4:
            ends.append(0); continue
        span = int(a['read_off'][r + 1] - a['read_off'][r])
        for f in range(int(a['feat_off'][r]), int(a['feat_off'][r + 1])):
            op = chr(a['f_op'][f])
            span += -1 if op == 'i' else -int(a['f_aux'][f]) if op == 'S' else int(a['f_val'][f]) if op == 'D' else 0
        ends.append(int(a['pos'][r]) + span - 1)
    return ends
