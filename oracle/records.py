"""ctypes front-end of oracle/records_oracle.c.  TEST INFRASTRUCTURE ONLY."""
import ctypes

import numpy as np

from . import rans as _rans

_c = ctypes
u8p = _c.POINTER(_c.c_uint8)
i32p = _c.POINTER(_c.c_int32)
i64p = _c.POINTER(_c.c_int64)
u32p = _c.POINTER(_c.c_uint32)
u16p = _c.POINTER(_c.c_uint16)


class RecFeatures(_c.Structure):
    _fields_ = [('n_feat', _c.c_int64), ('f_op', u8p), ('f_pos', i32p), ('f_val', i32p), ('f_aux', i32p),
                ('feat_off', i64p), ('payload', u8p), ('payload_len', _c.c_int64),
                ('subst_counts', (_c.c_int64 * 5) * 5), ('matrix', _c.c_uint8 * 5)]


class QsPolicy(_c.Structure):
    _fields_ = [('read_cat', _c.c_int32), ('read_param', _c.c_int32), ('n_base_cats', _c.c_int32),
                ('base_cat', _c.c_int32 * 8), ('treatment', _c.c_int32), ('treat_param', _c.c_int32)]


STREAMS = ['FN', 'FC', 'FP', 'BS', 'DL', 'BA', 'SC', 'QS', 'RS', 'HC', 'PD', 'MQ']

_done = False


def lib():
    global _done
    L = _rans.lib()
    if not _done:
        L.rec_oracle_normalize_base.restype = _c.c_uint8
        L.rec_oracle_normalize_base.argtypes = [_c.c_uint8]
        L.rec_oracle_itf8.restype = _c.c_int
        L.rec_oracle_itf8.argtypes = [_c.c_int32, u8p]
        L.rec_oracle_extract.restype = _c.c_int
        L.rec_oracle_extract.argtypes = [_c.c_int64, u8p, i64p, u32p, i64p, i32p, u16p, u8p, _c.c_int64, _c.POINTER(RecFeatures)]
        L.rec_oracle_free.argtypes = [_c.POINTER(RecFeatures)]
        L.rec_oracle_subst_matrix.argtypes = [(_c.c_int64 * 5) * 5, _c.c_uint8 * 5, (_c.c_uint8 * 5) * 5]
        L.rec_oracle_parse_policies.restype = _c.c_int
        L.rec_oracle_parse_policies.argtypes = [_c.c_char_p, _c.POINTER(QsPolicy), _c.c_int]
        L.rec_oracle_quality.restype = _c.c_int
        L.rec_oracle_quality.argtypes = [u8p, u8p, _c.c_int64, _c.c_int, u32p, _c.c_int64, _c.c_int32, _c.c_uint16,
                                         _c.c_uint8, u8p, _c.c_int64, _c.POINTER(QsPolicy), _c.c_int, u8p]
        L.rec_oracle_restore_bases.restype = _c.c_int
        L.rec_oracle_restore_bases.argtypes = [_c.c_int32, _c.c_int64, _c.c_int64, u8p, i32p, i32p, i32p, u8p, u8p,
                                               _c.c_int64, _c.c_int64, _c.c_uint8 * 5, u8p]
        L.rec_oracle_restore_quals.restype = _c.c_int
        L.rec_oracle_restore_quals.argtypes = [_c.c_int, _c.c_int64, _c.c_int64, u8p, i32p, i32p, i32p, u8p]
        L.rec_oracle_write_streams.argtypes = [_c.c_int64, u8p, i64p, u16p, u8p, i32p, _c.POINTER(RecFeatures), u8p, u8p,
                                               _c.POINTER(u8p), i64p]
        L.rec_oracle_write_host_streams.restype = None
        L.rec_oracle_write_host_streams.argtypes = [_c.c_int64, u16p, u8p, i32p, i64p, i32p, i32p, u8p, i64p, i32p, i32p, i32p, i32p,
                                                    i32p, _c.c_int, _c.c_int, _c.c_int32, _c.c_int, _c.POINTER(u8p), i64p]
        L.rec_oracle_crc32.restype = _c.c_uint32
        L.rec_oracle_crc32.argtypes = [u8p, _c.c_int64]
        L.rec_oracle_slice_header.restype = None
        L.rec_oracle_slice_header.argtypes = [_c.c_int64, u8p, u8p, i64p, u8p, i32p, i32p, u16p, u8p, i32p, i32p, i64p,
                                              _c.POINTER(SliceHeader)]
        L.rec_oracle_tracks_new.restype = _c.c_void_p
        L.rec_oracle_tracks_new.argtypes = [u8p, _c.c_int64, _c.c_int32]
        L.rec_oracle_tracks_free.argtypes = [_c.c_void_p]
        for nm in ('position', 'len', 'thrown'):
            getattr(L, 'rec_oracle_tracks_' + nm).restype = _c.c_int32
            getattr(L, 'rec_oracle_tracks_' + nm).argtypes = [_c.c_void_p]
        for nm in ('coverage', 'mismatches'):
            getattr(L, 'rec_oracle_tracks_' + nm).restype = _c.POINTER(_c.c_int16)
            getattr(L, 'rec_oracle_tracks_' + nm).argtypes = [_c.c_void_p]
        L.rec_oracle_tracks_ensure_range.argtypes = [_c.c_void_p, _c.c_int32, _c.c_int32]
        L.rec_oracle_tracks_update.argtypes = [_c.c_void_p, u8p, _c.c_int64, u32p, _c.c_int64, _c.c_int32]
        L.rec_oracle_span.restype = _c.c_int
        L.rec_oracle_span.argtypes = [_c.c_int64, i32p, u16p, u32p, _c.POINTER(_c.c_uint64), i32p, i32p]
        L.rec_oracle_quality_t.restype = _c.c_int
        L.rec_oracle_quality_t.argtypes = L.rec_oracle_quality.argtypes + [_c.c_void_p]
        _done = True
    return L


class SliceHeader(_c.Structure):
    """slice_header_t of records_oracle.c"""
    _fields_ = [('sequence_id', _c.c_int32), ('alignment_start', _c.c_int32), ('alignment_span', _c.c_int32),
                ('n_records', _c.c_uint32), ('bases', _c.c_uint64), ('bd', _c.c_uint32), ('sd', _c.c_uint32)]


# what MessageDigestHasher.asByteArray() yields in htsjdk 2.4.1 whatever the records hold: the digest of the empty message
SHA1_EMPTY = bytes.fromhex('da39a3ee5e6b4b0d3255bfef95601890afd80709')
SHA512_EMPTY = bytes.fromhex('cf83e1357eefb8bdf1542850d66d8007d620e4050b5715dc83f4a921d36ce9ce'
                             '47d0d13c5d85f2b0ff8318d2877eec2f63b931bd47417a81a538327af927da3e')


def slice_ref_md5(ref, sequence_id, alignment_start, alignment_span):
    """Slice.refMD5 as CramSerilization.convert leaves it (CramSerilization.java:184-197; Slice.setRefMD5 [Slice.java:145-159];
    SequenceUtil.calculateMD5 [SequenceUtil.java:962-966]): 16 zero bytes for a multi-reference slice (never set) and for an
    unplaced one, else MD5 of ref[alignmentStart - 1 : + min(span, len(ref) - alignmentStart + 1)]"""
    import hashlib
    if sequence_id < 0:
        return bytes(16)
    ref = bytes(ref)
    span = min(alignment_span, len(ref) - alignment_start + 1)
    return hashlib.md5(ref[alignment_start - 1:alignment_start - 1 + max(span, 0)]).digest()


def slice_header(rd, ex, scores):
    """ContainerFactory.buildSlice's header fields and BD / SD digests for one slice.
    rd needs ref_id besides the usual arrays; ex = extract(rd, ref); scores = quality(rd, ref, spec)[1].
    -> dict(sequence_id, alignment_start, alignment_span, n_records, bases, tags={BD, SD, B5, S5, B1, S1})"""
    import struct
    n = len(rd['pos'])
    one = np.zeros(1, np.uint8)
    bases = np.ascontiguousarray(rd['bases'], np.uint8) if len(rd['bases']) else one
    sc = np.ascontiguousarray(scores, np.uint8) if len(scores) else one
    rfl = np.ascontiguousarray(rd.get('rflags', np.zeros(n, np.uint8)), np.uint8)
    nf = ex.n_feat
    f_op = ex.f_op if nf else one
    f_val = ex.f_val if nf else np.zeros(1, np.int32)
    f_aux = ex.f_aux if nf else np.zeros(1, np.int32)
    H = SliceHeader()
    lib().rec_oracle_slice_header(n, _p(bases, u8p), _p(sc, u8p), _p(np.ascontiguousarray(rd['read_off'], np.int64), i64p), _p(rfl, u8p),
                                  _p(np.ascontiguousarray(rd['ref_id'], np.int32), i32p), _p(np.ascontiguousarray(rd['pos'], np.int32), i32p),
                                  _p(np.ascontiguousarray(rd['flag'], np.uint16), u16p), _p(np.ascontiguousarray(f_op), u8p),
                                  _p(np.ascontiguousarray(f_val), i32p), _p(np.ascontiguousarray(f_aux), i32p),
                                  _p(np.ascontiguousarray(ex.feat_off, np.int64), i64p), _c.byref(H))
    return dict(sequence_id=H.sequence_id, alignment_start=H.alignment_start, alignment_span=H.alignment_span, n_records=H.n_records,
                bases=H.bases, bd=H.bd, sd=H.sd,
                tags={'BD': struct.pack('<I', H.bd), 'SD': struct.pack('<I', H.sd), 'B5': SHA512_EMPTY, 'S5': SHA512_EMPTY,
                      'B1': SHA1_EMPTY, 'S1': SHA1_EMPTY})


def _p(a, t):
    return a.ctypes.data_as(t)


def itf8(v):
    buf = np.zeros(8, np.uint8)
    n = lib().rec_oracle_itf8(int(v), _p(buf, u8p))
    return buf[:n].tobytes()


def parse_policies(spec):
    """-> list of QsPolicy (sorted as the reference sorts), or raises ValueError"""
    if spec is None or spec == 'lossless':
        return None
    arr = (QsPolicy * 16)()
    n = lib().rec_oracle_parse_policies(spec.encode(), arr, 16)
    if n == -1:
        raise ValueError('Unknown read or base category')
    return arr, n


def subst_matrix(counts):
    """counts: dict 'AC' -> n (ref, read).  -> (5 matrix bytes, codes dict)"""
    C = ((_c.c_int64 * 5) * 5)()
    B = 'ACGTN'
    for k, v in counts.items():
        C[B.index(k[0])][B.index(k[1])] = v
    M = (_c.c_uint8 * 5)()
    K = ((_c.c_uint8 * 5) * 5)()
    lib().rec_oracle_subst_matrix(C, M, K)
    return bytes(M), {B[r] + B[b]: K[r][b] for r in range(5) for b in range(5) if r != b}


class Extracted(object):
    """numpy views (copied) of rec_features_t"""

    def __init__(self, rf, n_reads):
        n = rf.n_feat
        self.n_feat = n
        def arr(p, dt):
            return np.ctypeslib.as_array(p, (n,)).copy() if n and p else np.zeros(0, dt)          # no feature at all: NULL
        self.f_op = arr(rf.f_op, np.uint8)
        self.f_pos = arr(rf.f_pos, np.int32)
        self.f_val = arr(rf.f_val, np.int32)
        self.f_aux = arr(rf.f_aux, np.int32)
        self.feat_off = np.ctypeslib.as_array(rf.feat_off, (n_reads + 1,)).copy()
        pl = rf.payload_len
        self.payload = np.ctypeslib.as_array(rf.payload, (max(pl, 1),))[:pl].copy() if pl else np.zeros(0, np.uint8)
        self.matrix = bytes(rf.matrix)
        self.subst_counts = {'ACGTN'[r] + 'ACGTN'[b]: rf.subst_counts[r][b] for r in range(5) for b in range(5)
                             if rf.subst_counts[r][b]}


def extract(rd, ref):
    """rows F2-F5 for one container.  rd = SoA dict (cramtools_b200.synth.reads layout)."""
    n = len(rd['pos'])
    rf = RecFeatures()
    bases = np.ascontiguousarray(rd['bases'], np.uint8) if len(rd['bases']) else np.zeros(1, np.uint8)
    cigar = np.ascontiguousarray(rd['cigar'], np.uint32) if len(rd['cigar']) else np.zeros(1, np.uint32)
    ref = np.ascontiguousarray(ref, np.uint8)
    rc = lib().rec_oracle_extract(n, _p(bases, u8p), _p(np.ascontiguousarray(rd['read_off'], np.int64), i64p),
                                  _p(cigar, u32p), _p(np.ascontiguousarray(rd['cigar_off'], np.int64), i64p),
                                  _p(np.ascontiguousarray(rd['pos'], np.int32), i32p),
                                  _p(np.ascontiguousarray(rd['flag'], np.uint16), u16p), _p(ref, u8p), ref.size, _c.byref(rf))
    if rc:
        raise RuntimeError('Unsupported cigar operator')
    out = Extracted(rf, n)
    lib().rec_oracle_free(_c.byref(rf))
    return out


def quality(rd, ref, spec, tracks=None, positions=None):
    """row Q for every read.  -> (force_qs u8[n] (bit0 force, bit1 has array), scores u8[sum len] 0xFF = absent).
    tracks: a Tracks object after .container(rd) when the model has X / P selectors (positions: list that receives the
    window position after every record)"""
    n = len(rd['pos'])
    pol = parse_policies(spec)
    arr, npol = (None, -1) if pol is None else pol
    bases = np.ascontiguousarray(rd['bases'], np.uint8)
    quals = np.ascontiguousarray(rd['quals'], np.uint8)
    ref = np.ascontiguousarray(ref, np.uint8)
    cigar = np.ascontiguousarray(rd['cigar'], np.uint32)
    scores = np.full(max(bases.size, 1), 0xFF, np.uint8)
    force = np.zeros(n, np.uint8)
    rflags = rd.get('rflags', np.zeros(n, np.uint8))
    dummy = np.zeros(1, np.uint8)
    dummyc = np.zeros(1, np.uint32)
    for r in range(n):
        lo, hi = int(rd['read_off'][r]), int(rd['read_off'][r + 1])
        c0, c1 = int(rd['cigar_off'][r]), int(rd['cigar_off'][r + 1])
        rb = bases[lo:hi] if hi > lo else dummy
        q = quals[lo:hi] if hi > lo else dummy
        cg = cigar[c0:c1] if c1 > c0 else dummyc
        sc = scores[lo:hi] if hi > lo else dummy
        force[r] = lib().rec_oracle_quality_t(_p(rb, u8p), _p(q, u8p), hi - lo, 0 if (rflags[r] & 1) else 1, _p(cg, u32p),
                                              c1 - c0, int(rd['pos'][r]), int(rd['flag'][r]), int(rd['mapq'][r]),
                                              _p(ref, u8p), ref.size, arr, npol, _p(sc, u8p), None if tracks is None else tracks.h)
        if positions is not None:
            positions.append(tracks.position)
    return force, scores[:bases.size]


class Tracks(object):
    """ReferenceTracks of one reference sequence (ref/ReferenceTracks.java, restated in records_oracle.c): the coverage and
    mismatch counters the X and P selectors read, carried from container to container as CramSerilization.convert does"""
    THROWN = {0: None, 1: 'java/lang/ArrayIndexOutOfBoundsException', 2: 'java/lang/RuntimeException', 3: 'java/lang/RuntimeException'}

    def __init__(self, ref, window=1000000):
        self.ref = np.ascontiguousarray(ref, np.uint8)                  # kept alive: the C side holds the pointer
        self.h = lib().rec_oracle_tracks_new(_p(self.ref, u8p), self.ref.size, int(window))

    def __del__(self):
        if getattr(self, 'h', None):
            lib().rec_oracle_tracks_free(self.h); self.h = None

    position = property(lambda self: lib().rec_oracle_tracks_position(self.h))
    length = property(lambda self: lib().rec_oracle_tracks_len(self.h))
    thrown = property(lambda self: lib().rec_oracle_tracks_thrown(self.h))
    coverage = property(lambda self: np.ctypeslib.as_array(lib().rec_oracle_tracks_coverage(self.h), (self.length,)).copy())
    mismatches = property(lambda self: np.ctypeslib.as_array(lib().rec_oracle_tracks_mismatches(self.h), (self.length,)).copy())

    def container(self, rd):
        """what CramSerilization.convert does with the tracks before it looks at a record [:121-131]: the container's span on the
        sequence (getSpans), ensureRange, updateTracks.  rd: the container's reads (ref_id -1 = unplaced).  -> (start, span)"""
        n = len(rd['pos'])
        ref_id = rd.get('ref_id', np.zeros(n, np.int32))
        keep = np.flatnonzero(np.asarray(ref_id) != -1)
        pos = np.ascontiguousarray(np.asarray(rd['pos'])[keep], np.int32); flag = np.ascontiguousarray(np.asarray(rd['flag'])[keep], np.uint16)
        cigar = np.ascontiguousarray(rd['cigar'], np.uint32); bases = np.ascontiguousarray(rd['bases'], np.uint8)
        if cigar.size == 0:
            cigar = np.zeros(1, np.uint32)
        if bases.size == 0:
            bases = np.zeros(1, np.uint8)
        co = np.asarray(rd['cigar_off'])
        sub_off = np.zeros(keep.size + 1, np.uint64)
        pieces = []
        for k, r in enumerate(keep):
            pieces.append(cigar[int(co[r]):int(co[r + 1])]); sub_off[k + 1] = sub_off[k] + (int(co[r + 1]) - int(co[r]))
        sub = np.ascontiguousarray(np.concatenate(pieces) if pieces else np.zeros(1, np.uint32), np.uint32)
        if sub.size == 0:
            sub = np.zeros(1, np.uint32)
        st, sp = _c.c_int32(0), _c.c_int32(0)
        if not lib().rec_oracle_span(keep.size, _p(pos, i32p), _p(flag, u16p), _p(sub, u32p), sub_off.ctypes.data_as(_c.POINTER(_c.c_uint64)),
                                     _c.byref(st), _c.byref(sp)):
            return None
        lib().rec_oracle_tracks_ensure_range(self.h, st.value, sp.value)
        for r in keep:
            lo, hi = int(rd['read_off'][r]), int(rd['read_off'][r + 1])
            c0, c1 = int(co[r]), int(co[r + 1])
            rb = bases[lo:hi] if hi > lo else np.zeros(1, np.uint8)
            cg = cigar[c0:c1] if c1 > c0 else np.zeros(1, np.uint32)
            lib().rec_oracle_tracks_update(self.h, _p(np.ascontiguousarray(rb), u8p), hi - lo, _p(np.ascontiguousarray(cg), u32p), c1 - c0, int(rd['pos'][r]))
        return st.value, sp.value


def restore_bases(pos, rlen, ex, r, window, window_off0, matrix):
    """row N1 for read r of an Extracted feature set.  -> bytes or None (Java would throw)"""
    a, b = int(ex.feat_off[r]), int(ex.feat_off[r + 1])
    out = np.zeros(max(rlen, 1), np.uint8)
    window = np.ascontiguousarray(window, np.uint8)
    M = (_c.c_uint8 * 5)(*matrix)
    op = ex.f_op[a:b] if b > a else np.zeros(1, np.uint8)
    fp = ex.f_pos[a:b] if b > a else np.zeros(1, np.int32)
    fv = ex.f_val[a:b] if b > a else np.zeros(1, np.int32)
    fa = ex.f_aux[a:b] if b > a else np.zeros(1, np.int32)
    pl = ex.payload if ex.payload.size else np.zeros(1, np.uint8)
    rc = lib().rec_oracle_restore_bases(int(pos), rlen, b - a, _p(np.ascontiguousarray(op), u8p), _p(np.ascontiguousarray(fp), i32p),
                                        _p(np.ascontiguousarray(fv), i32p), _p(np.ascontiguousarray(fa), i32p), _p(pl, u8p),
                                        _p(window, u8p), window.size, window_off0, M, _p(out, u8p))
    return None if rc else out[:rlen].tobytes()


def restore_quals(force, rlen, scores, q_pos, q_val):
    """row N2.  scores: stored array (forced) ; q_pos/q_val: Q features otherwise.  -> bytes or '*'"""
    sc = np.array(scores, np.uint8).copy() if force else np.zeros(max(rlen, 1), np.uint8)
    if sc.size == 0:
        sc = np.zeros(1, np.uint8)
    n = len(q_pos)
    op = np.full(max(n, 1), ord('Q'), np.uint8)
    fp = np.array(list(q_pos) + [0] * (1 if n == 0 else 0), np.int32)
    fv = np.array(list(q_val) + [0] * (1 if n == 0 else 0), np.int32)
    fa = np.zeros(max(n, 1), np.int32)
    star = lib().rec_oracle_restore_quals(1 if force else 0, rlen, n, _p(op, u8p), _p(fp, i32p), _p(fv, i32p), _p(fa, i32p), _p(sc, u8p))
    return '*' if star else sc[:rlen].tobytes()


def write_streams(rd, ex, force, scores):
    """row W: -> dict series name -> bytes"""
    n = len(rd['pos'])
    total = int(rd['read_off'][-1])
    cap = 8 * total + 16 * int(ex.n_feat) + 16 * n + 64
    bufs = [np.zeros(cap, np.uint8) for _ in STREAMS]
    ptrs = (u8p * len(STREAMS))(*[_p(b, u8p) for b in bufs])
    lens = np.zeros(len(STREAMS), np.int64)
    rf = RecFeatures()
    keep = [np.ascontiguousarray(x) if len(x) else np.zeros(1, x.dtype) for x in (ex.f_op, ex.f_pos, ex.f_val, ex.f_aux, ex.feat_off, ex.payload)]
    rf.n_feat = ex.n_feat
    rf.f_op = _p(keep[0], u8p); rf.f_pos = _p(keep[1], i32p); rf.f_val = _p(keep[2], i32p); rf.f_aux = _p(keep[3], i32p)
    rf.feat_off = _p(keep[4], i64p); rf.payload = _p(keep[5], u8p); rf.payload_len = ex.payload.size
    bases = np.ascontiguousarray(rd['bases'], np.uint8) if total else np.zeros(1, np.uint8)
    sc = np.ascontiguousarray(scores, np.uint8) if total else np.zeros(1, np.uint8)
    lib().rec_oracle_write_streams(n, _p(bases, u8p), _p(np.ascontiguousarray(rd['read_off'], np.int64), i64p),
                                   _p(np.ascontiguousarray(rd['flag'], np.uint16), u16p),
                                   _p(np.ascontiguousarray(rd['mapq'], np.uint8), u8p),
                                   _p(np.ascontiguousarray(rd['pos'], np.int32), i32p), _c.byref(rf),
                                   _p(np.ascontiguousarray(force, np.uint8), u8p), _p(sc, u8p), ptrs, _p(lens, i64p))
    return {nm: bufs[i][:int(lens[i])].tobytes() for i, nm in enumerate(STREAMS)}


def read_streams(st, read_len, rflags):
    """row P: the feature part of CramRecordReader.read [encoding/reader/CramRecordReader.java:59-...], the mirror of
    Writer.write [Writer.java:180-258], restated for the series of write_streams (pure Python: small cases only).
    st: dict series name -> bytes; read_len[r]; rflags[r]: bit0 force-preserved score array, bit1 unmapped, bit2 unknown
    bases.  -> (rows, mapq, qs_off, ba_off): rows[r] = [[op, pos, value], ...] with value = int, or bytes for 'S';
    qs_off / ba_off = offset of the record's score array in QS / of an unmapped record's bases in BA (None = none).
    Parity: unpinned by the jar itself (its reader needs the whole codec factory); pinned indirectly -- reading back the
    series write_streams makes from the jar's feature lists must give those lists (tests/test_oracle_records.py)."""
    cur = {k: 0 for k in st}

    def byte(name):
        v = st[name][cur[name]]
        cur[name] += 1
        return v

    def itf8(name):                                  # ITF8.readUnsignedITF8 [ITF8.java:21-51]
        b1 = byte(name)
        if b1 < 0x80:
            v = b1
        elif b1 < 0xC0:
            v = ((b1 & 0x7F) << 8) | byte(name)
        elif b1 < 0xE0:
            v = ((b1 & 0x3F) << 16) | (byte(name) << 8)
            v |= byte(name)
        elif b1 < 0xF0:
            v = ((b1 & 0x1F) << 24) | (byte(name) << 16)
            v |= byte(name) << 8
            v |= byte(name)
        else:
            v = ((b1 & 0x0F) << 28) | (byte(name) << 20)
            v |= byte(name) << 12
            v |= byte(name) << 4
            v |= byte(name) & 0x0F
        return v - (1 << 32) if v >= (1 << 31) else v
    rows, mapq, qs_off, ba_off = [], [], [], []
    for r in range(len(rflags)):
        fl, rl = int(rflags[r]), int(read_len[r])
        row, mq, qo, bo = [], 0, None, None
        if not fl & 2:
            nf = itf8('FN')
            prev = 0
            for _ in range(nf):
                op = chr(byte('FC'))
                prev += itf8('FP')
                if op == 'X':
                    row.append([op, prev, byte('BS')])
                elif op == 'D':
                    row.append([op, prev, itf8('DL')])
                elif op == 'N':
                    row.append([op, prev, itf8('RS')])
                elif op == 'H':
                    row.append([op, prev, itf8('HC')])
                elif op == 'P':
                    row.append([op, prev, itf8('PD')])
                elif op == 'i':
                    row.append([op, prev, byte('BA')])
                elif op == 'Q':
                    row.append([op, prev, byte('QS')])
                elif op == 'S':
                    e = st['SC'].index(b'\t', cur['SC'])              # ByteArrayStopEncoding: bytes up to the stop byte
                    row.append([op, prev, bytes(st['SC'][cur['SC']:e])])
                    cur['SC'] = e + 1
                else:
                    raise ValueError('read feature operator %r' % op)
            mq = itf8('MQ')
        elif not fl & 4:
            bo = cur['BA']
            cur['BA'] += rl
        if fl & 1:
            qo = cur['QS']
            cur['QS'] += rl
        rows.append(row); mapq.append(mq); qs_off.append(qo); ba_off.append(bo)
    for k in st:
        assert cur[k] == len(st[k]), (k, cur[k], len(st[k]))      # every byte of every series accounted for
    return rows, mapq, qs_off, ba_off


# ---------------------------------------------------------------------------------------------- host-side series (row W, first half)
# EncodingKey.ordinal() = content id of the series' external block (htsjdk/samtools/cram/structure/EncodingKey)
CONTENT_ID = {'BF': 0, 'AP': 1, 'FP': 2, 'FC': 3, 'QS': 4, 'DL': 5, 'BA': 6, 'TN': 7, 'NF': 8, 'RL': 9, 'RG': 10, 'MQ': 11,
              'RN': 12, 'NP': 13, 'TS': 14, 'FN': 15, 'BS': 16, 'IN': 17, 'TC': 18, 'MF': 19, 'NS': 20, 'CF': 21, 'TL': 24,
              'RI': 25, 'RS': 26, 'SC': 27, 'HC': 28, 'PD': 29}
HOST_STREAMS = ['BF', 'CF', 'RI', 'RL', 'AP', 'RG', 'RN', 'MF', 'NS', 'NP', 'TS', 'NF', 'TL']

# CramCompressionRecord flag bits (htsjdk/samtools/cram/structure/CramCompressionRecord)
CF_FORCE_QS, CF_DETACHED, CF_MATE_DOWNSTREAM, CF_UNKNOWN_BASES = 1, 2, 4, 8
MF_NEG_STRAND, MF_UNMAPPED = 1, 2


def cram_bit_flags(sam_flag):
    """BF of Sam2CramRecordFactory.createCramRecord [:122-131,143-152]: the SAM flag without the two mate bits (they go to
    MF); for an unpaired read the proper-pair and first / last segment bits are dropped too."""
    f = int(sam_flag)
    if f & 0x1:
        return f & (0x1 | 0x2 | 0x4 | 0x10 | 0x40 | 0x80 | 0x100 | 0x200 | 0x400 | 0x800)
    return f & (0x4 | 0x10 | 0x100 | 0x200 | 0x400 | 0x800)


def mate_bit_flags(sam_flag):
    """MF [:114-115]: bit 0 mate on the reverse strand (SAM 0x20), bit 1 mate unmapped (SAM 0x8); 0 for an unpaired read"""
    f = int(sam_flag)
    if not f & 0x1:
        return 0
    return (MF_NEG_STRAND if f & 0x20 else 0) | (MF_UNMAPPED if f & 0x8 else 0)


def write_host_streams(sc, ap_delta, multi_ref, slice_start, capture_names=True):
    """The fields Writer.write emits before the read features [Writer.java:139-172], per record in this order:
    BF (ITF8), CF (byte), RI (ITF8, multi-reference slices only), RL, AP (delta to the previous record's start when the
    header says so -- ContainerFactory.buildSlice [:179-182], the first record against the slice start -- else the start
    itself), RG, RN (name + TAB), then MF (byte), NS, NP, TS for a detached record or NF for one whose mate follows, TL.
    sc: list of dicts with flags, cf, seq_id, read_length, start, read_group, name, mate_flags, mate_seq_id, mate_start,
    tlen, nf, tl.  -> dict series -> bytes"""
    out = {k: bytearray() for k in HOST_STREAMS}
    prev = slice_start
    for r in sc:
        out['BF'] += itf8(r['flags'])
        out['CF'].append(r['cf'] & 0xFF)
        if multi_ref:
            out['RI'] += itf8(r['seq_id'])
        out['RL'] += itf8(r['read_length'])
        if ap_delta:
            out['AP'] += itf8(r['start'] - prev)
            prev = r['start']
        else:
            out['AP'] += itf8(r['start'])
        out['RG'] += itf8(r['read_group'])
        if capture_names:
            out['RN'] += r['name'].encode('latin-1') + b'\t'
        if r['cf'] & CF_DETACHED:
            out['MF'].append(r['mate_flags'] & 0xFF)
            if not capture_names:
                out['RN'] += r['name'].encode('latin-1') + b'\t'
            out['NS'] += itf8(r['mate_seq_id'])
            out['NP'] += itf8(r['mate_start'])
            out['TS'] += itf8(r['tlen'])
        elif r['cf'] & CF_MATE_DOWNSTREAM:
            out['NF'] += itf8(r['nf'])
        out['TL'] += itf8(r.get('tl', 0))
    return {k: bytes(v) for k, v in out.items()}


def write_host_streams_c(flag, cf, ref_id, read_off, pos, read_group, names, mate_ref, mate_pos, tlen, nf, tl, ap_delta, multi_ref,
                         slice_start, capture_names):
    """the same through the C restatement (oracle/records_oracle.c: rec_oracle_write_host_streams), which is what the CPU
    baseline of bench.py runs; BF and MF come from the SAM flags there"""
    n = len(flag)

    def arr(x, dt):
        return None if x is None else np.ascontiguousarray(x, dt)

    def ptr(a, t):
        return None if a is None else a.ctypes.data_as(t)
    nb = None if names is None else [x.encode('latin-1') if isinstance(x, str) else bytes(x) for x in names]
    nm = None if nb is None else np.frombuffer(b''.join(nb) + b'\0', np.uint8).copy()
    no = None if nb is None else np.concatenate([[0], np.cumsum([len(x) for x in nb])]).astype(np.int64)
    a = dict(flag=arr(flag, np.uint16), cf=arr(cf, np.uint8), ref_id=arr(ref_id, np.int32), read_off=arr(read_off, np.int64),
             pos=arr(pos, np.int32), rg=arr(read_group, np.int32), mr=arr(mate_ref, np.int32), mp=arr(mate_pos, np.int32),
             ts=arr(tlen, np.int32), nf=arr(nf, np.int32), tl=arr(tl, np.int32))
    cap = 5 * n + 16
    bufs = [np.zeros((int(no[-1]) + n + 16) if (k == 6 and no is not None) else cap, np.uint8) for k in range(len(HOST_STREAMS))]
    ptrs = (u8p * len(HOST_STREAMS))(*[_p(b, u8p) for b in bufs])
    lens = np.zeros(len(HOST_STREAMS), np.int64)
    lib().rec_oracle_write_host_streams(n, ptr(a['flag'], u16p), ptr(a['cf'], u8p), ptr(a['ref_id'], i32p), ptr(a['read_off'], i64p),
                                        ptr(a['pos'], i32p), ptr(a['rg'], i32p), ptr(nm, u8p), ptr(no, i64p), ptr(a['mr'], i32p),
                                        ptr(a['mp'], i32p), ptr(a['ts'], i32p), ptr(a['nf'], i32p), ptr(a['tl'], i32p),
                                        1 if ap_delta else 0, 1 if multi_ref else 0, int(slice_start), 1 if capture_names else 0,
                                        ptrs, _p(lens, i64p))
    return {nmk: bufs[i][:int(lens[i])].tobytes() for i, nmk in enumerate(HOST_STREAMS)}


# ---------------------------------------------------------------------------------------------- mate linking
def alignment_ends(rd, ex):
    """CramCompressionRecord.getAlignmentEnd() [CramCompressionRecord.java:160-194] per record: 0 with the unmapped flag, else
    start + readLength - #'i' - len('S') + len('D') - 1 (reference skips not counted)"""
    n = len(rd['pos'])
    out = np.zeros(n, np.int64)
    rl = np.diff(np.asarray(rd['read_off'], np.int64))
    for r in range(n):
        if int(rd['flag'][r]) & 4:
            continue
        span = int(rl[r])
        for f in range(int(ex.feat_off[r]), int(ex.feat_off[r + 1])):
            op = chr(ex.f_op[f])
            if op == 'i':
                span -= 1
            elif op == 'S':
                span -= int(ex.f_aux[f])
            elif op == 'D':
                span += int(ex.f_val[f])
        out[r] = int(rd['pos'][r]) + span - 1
    return out


def set_mate_info(flag, names, ref_id, pos, mate_ref_id, mate_pos, tlen, aln_end, coordinate_sorted=True):
    """CramSerilization.setMateInfo / detach (src/main/java/htsjdk/samtools/cram/CramSerilization.java:258-373) for the records
    of one container, in order, with the record fields as Sam2CramRecordFactory.createCramRecord sets them
    [Sam2CramRecordFactory.java:112-146].  When the header is not coordinate sorted every record is detached [:175-179].
    -> (mate_cf u8: bit 1 (value 2) detached, bit 2 (value 4) mate downstream; nf i32: recordsToNextFragment, -1 = none)"""
    n = len(flag)
    flag = [int(x) for x in flag]
    det = [False] * n; down = [False] * n; nf = [-1] * n; nxt = [-1] * n; prv = [-1] * n
    if not coordinate_sorted:
        return np.full(n, 2, np.uint8), np.full(n, -1, np.int32)
    paired = [bool(f & 1) for f in flag]
    first = [paired[i] and bool(flag[i] & 0x40) for i in range(n)]
    last_seg = [paired[i] and not (flag[i] & 0x40) and bool(flag[i] & 0x80) for i in range(n)]
    neg = [bool(f & 0x10) for f in flag]
    mate_neg = [paired[i] and bool(flag[i] & 0x20) for i in range(n)]
    mseq = [int(mate_ref_id[i]) if paired[i] else -1 for i in range(n)]
    mstart = [int(mate_pos[i]) if paired[i] else 0 for i in range(n)]
    unm = [bool(f & 4) for f in flag]

    def detach(i):
        while i != -1:
            det[i] = True; down[i] = False; nf[i] = -1
            i = nxt[i]

    def insert_size(a, b):                                       # CramNormalizer.computeInsertSize [CramNormalizer.java:455-470]
        if unm[a] or unm[b] or int(ref_id[a]) != int(ref_id[b]):
            return 0
        pa = int(aln_end[a]) if neg[a] else int(pos[a])
        pb = int(aln_end[b]) if neg[b] else int(pos[b])
        return pb - pa + (1 if pb >= pa else -1)
    maps = ({}, {})
    for i in range(n):
        if not paired[i]:
            det[i] = True; down[i] = False; nf[i] = -1
            continue
        m = maps[1 if flag[i] & 0x100 else 0]
        key = bytes(names[i]) if not isinstance(names[i], str) else names[i].encode('latin-1')
        j = m.get(key)
        if j is None:
            m[key] = i
            continue
        if nxt[j] != -1 or prv[j] != -1:
            detach(i); detach(j)
            continue
        nf[j] = i - j - 1
        nxt[j] = i; prv[i] = j
        down[j] = True; down[i] = False; det[i] = False; det[j] = False
    for r in range(n):
        if nxt[r] == -1 or prv[r] != -1:
            continue
        last = r
        while nxt[last] != -1:
            last = nxt[last]
        if (first[r] and last_seg[last]) or (first[last] and last_seg[r]):
            tl = insert_size(r, last)
            if int(tlen[r]) == tl:
                l2 = nxt[r]
                while nxt[l2] != -1:
                    if int(tlen[l2]) != -tl:
                        break
                    l2 = nxt[l2]
                if int(tlen[l2]) != -tl:
                    detach(r)
            else:
                detach(r)
        else:
            detach(r)
        if (mseq[r] != int(ref_id[last]) or int(ref_id[r]) != mseq[last] or mate_neg[r] != neg[last] or mate_neg[last] != neg[r]
                or mstart[r] != int(pos[last]) or mstart[last] != int(pos[r])):
            detach(r)
    for m in maps:
        for i in m.values():
            if nxt[i] != -1:
                continue
            det[i] = True; down[i] = False; nf[i] = -1
    cf = np.array([(2 if det[i] else 0) | (4 if down[i] else 0) for i in range(n)], np.uint8)
    return cf, np.array(nf, np.int32)


# ---------------------------------------------------------------------------------------------- CIGAR rebuild + BAM records
_CONSUMES_READ = {'M': True, 'I': True, 'S': True, '=': True, 'X': True, 'D': False, 'N': False, 'H': False, 'P': False}


def cigar_from_features(features, read_length):
    """Cram2SamRecordFactory.getCigar2 [Cram2SamRecordFactory.java:119-225].  features: [(op char, 1-based position, length)] in
    record order, length = bases of an 'I' / 'S', the value of 'D' 'N' 'H' 'P' (ignored for 'i' 'X' 'B'); any other operator
    (quality scores) is skipped.  -> [(length, op char)]"""
    if not features:
        return [(read_length, 'M')]
    out = []
    last_op, last_len, last_pos = 'M', 0, 1
    for op, pos, ln in features:
        gap = pos - (last_pos + last_len)
        if gap > 0:
            if last_op != 'M':
                out.append((last_len, last_op)); last_pos += last_len; last_len = gap
            else:
                last_len += gap
            last_op = 'M'
        if op in ('I', 'S'):
            co, rl = ('I' if op == 'I' else 'S'), ln
        elif op == 'H':
            co, rl = 'H', ln
        elif op == 'i':
            co, rl = 'I', 1
        elif op == 'D':
            co, rl = 'D', ln
        elif op == 'N':
            co, rl = 'N', ln
        elif op == 'P':
            co, rl = 'P', ln
        elif op in ('X', 'B'):
            co, rl = 'M', 1
        else:
            continue
        if last_op != co:
            if last_len > 0:
                out.append((last_len, last_op))
            last_op, last_len, last_pos = co, rl, pos
        else:
            last_len += rl
        if not _CONSUMES_READ[co]:
            last_pos -= rl
    if last_op != 'M':
        out.append((last_len, last_op))
        if read_length >= last_pos + last_len:
            out.append((read_length - (last_len + last_pos) + 1, 'M'))
    elif read_length == 0 or read_length > last_pos - 1:
        out.append((last_len, 'M') if read_length == 0 else (read_length - last_pos + 1, 'M'))
    return out if out else [(read_length, 'M')]


def reg2bin(beg, end):
    """GenomicIndexUtil.regionToBin [GenomicIndexUtil.java:63-70]"""
    end -= 1
    for shift, base in ((14, 4681), (17, 585), (20, 73), (23, 9), (26, 1)):
        if beg >> shift == end >> shift:
            return base + (beg >> shift)
    return 0


def indexing_bin(pos, flag, cigar):
    """SAMRecord.computeIndexingBin [SAMRecord.java:1524-1531] with getAlignmentEnd [:591-598]; cigar = [(len, op)]"""
    beg = pos - 1
    end = 0 if flag & 4 else pos + sum(l for l, o in cigar if o in 'MDN=X') - 1
    if end <= 0:
        end = beg + 1
    return reg2bin(beg, end)


_NYBBLE = {c: i for i, c in enumerate('=ACMGRSVTWYHKDBN')}
_NYBBLE.update({c.lower(): i for c, i in list(_NYBBLE.items()) if c != '='})
_NYBBLE['.'] = 15


def bam_record(ref_id, pos, name, mapq, flag, cigar, mate_ref_id, mate_pos, tlen, bases, quals, tags=b''):
    """one BAM record as BAMRecordView lays it out (src/main/java/htsjdk/samtools/cram/encoding/reader/BAMRecordView.java:70-241):
    block_size refID pos-1 l_read_name mapq bin n_cigar flag l_seq next_refID next_pos-1 tlen name\\0 cigar seq(4 bit) qual tags;
    quals == b'' -> l_seq bytes 0xFF.  A base outside =ACMGRSVTWYHKDBN (any case) or '.': ValueError (IllegalArgumentException)."""
    import struct
    ops = 'MIDNSHP=X'
    body = struct.pack('<iiBBHHHiiii', ref_id, pos - 1, len(name) + 1, mapq & 0xFF, indexing_bin(pos, flag, cigar) & 0xFFFF, len(cigar) & 0xFFFF,
                       flag & 0xFFFF, len(bases), mate_ref_id, mate_pos - 1, tlen)
    body += bytes(name) + b'\0' + b''.join(struct.pack('<I', (l << 4) | ops.index(o)) for l, o in cigar)
    nyb = []
    for b in bytes(bases):
        if chr(b) not in _NYBBLE:
            raise ValueError('Bad  byte passed to charToCompressedBase: %d' % b)
        nyb.append(_NYBBLE[chr(b)])
    if len(nyb) & 1:
        nyb.append(0)
    body += bytes((nyb[i] << 4) | nyb[i + 1] for i in range(0, len(nyb), 2))
    body += bytes(quals) if len(quals) else b'\xff' * len(bases)
    body += bytes(tags)
    return struct.pack('<i', len(body)) + body


def bam_parse(rec):
    """the inverse, for one record with its block_size prefix -> dict (pos / mate_pos 1-based, bases ASCII upper case)"""
    import struct
    bs, ref, pos, lname, mapq, bin_, ncig, flag, lseq, mref, mpos, tlen = struct.unpack('<iiiBBHHHiiii', rec[:36])
    o = 36
    name = rec[o:o + lname - 1]; o += lname
    cig = [((w >> 4), 'MIDNSHP=X'[w & 15]) for w in struct.unpack('<%dI' % ncig, rec[o:o + 4 * ncig])]; o += 4 * ncig
    sq = rec[o:o + (lseq + 1) // 2]; o += (lseq + 1) // 2
    bases = bytes(ord('=ACMGRSVTWYHKDBN'[(sq[i >> 1] >> (0 if i & 1 else 4)) & 15]) for i in range(lseq))
    quals = rec[o:o + lseq]; o += lseq
    return dict(block_size=bs, ref_id=ref, pos=pos + 1, name=name, mapq=mapq, bin=bin_, flag=flag, cigar=cig, mate_ref_id=mref,
                mate_pos=mpos + 1, tlen=tlen, bases=bases, quals=quals, tags=rec[o:4 + bs])


# ---------------------------------------------------------------------------------------------- CRAM v3 blocks (the block layer)
BLOCK_METHOD = {'RAW': 0, 'GZIP': 1, 'BZIP2': 2, 'LZMA': 3, 'RANS': 4}          # BlockCompressionMethod.ordinal()
BLOCK_CONTENT = {'FILE_HEADER': 0, 'COMPRESSION_HEADER': 1, 'MAPPED_SLICE': 2, 'RESERVED': 3, 'EXTERNAL': 4, 'CORE': 5}


def cram_block_v3(method, content_type, content_id, raw_len, data):
    """Block.write(3, os) [htsjdk Block.java:272-294]: method, content type, ITF8 content id, ITF8 compressed size, ITF8 raw size,
    the (compressed) bytes, and the CRC32 of everything before it, little endian [CRC32OutputStream.java:50-51]"""
    import struct
    import zlib
    b = bytes([method, content_type]) + itf8(content_id) + itf8(len(data)) + itf8(raw_len) + bytes(data)
    return b + struct.pack('<I', zlib.crc32(b) & 0xFFFFFFFF)


# ---------------------------------------------------------------------------------------------- CRAM v3 container (the container part)
def ltf8(v):
    """LTF8.writeUnsignedLTF8 [htsjdk LTF8.java:104-190]: 1-9 bytes, the leading ones of the first byte count the bytes that follow"""
    v &= (1 << 64) - 1
    for n in range(1, 9):
        if v >> (7 * n) == 0:
            body = [(v >> (8 * k)) & 0xFF for k in range(n - 1, -1, -1)]
            body[0] |= (0xFF << (9 - n)) & 0xFF
            return bytes(body)
    return bytes([0xFF]) + bytes((v >> (8 * k)) & 0xFF for k in range(7, -1, -1))


def cram_array(ints):
    """CramArray.write(int[], os): ITF8 count, then every element as ITF8"""
    return itf8(len(ints)) + b''.join(itf8(int(v)) for v in ints)


def bam_tag_bytes(tag, value):
    """BinaryTagCodec.writeTag of a byte[] value [BinaryTagCodec.java:185-228, writeArray]: tag, 'B', 'c', int32 count, bytes"""
    import struct
    return tag.encode('latin-1') + b'Bc' + struct.pack('<i', len(value)) + bytes(value)


def slice_header_content_v3(sequence_id, alignment_start, alignment_span, n_records, global_record_counter, content_ids, ref_md5,
                            tags, embedded_ref_id=-1):
    """SliceIO.createSliceHeaderBlockContent(3, slice) [SliceIO.java:76-107] with nofBlocks as SliceIO.write sets it [:132]:
    1 (core) + the external blocks.  tags: [(name, bytes)] in the order of the slice's tag list (B1 S1 B5 S5 BD SD), or None"""
    b = itf8(sequence_id) + itf8(alignment_start) + itf8(alignment_span) + itf8(n_records) + ltf8(global_record_counter)
    b += itf8(1 + len(content_ids)) + cram_array(content_ids) + itf8(embedded_ref_id)
    b += bytes(ref_md5) if ref_md5 is not None else bytes(16)
    for name, val in (tags or []):
        b += bam_tag_bytes(name, val)
    return b


def container_header_v3(byte_size, sequence_id, alignment_start, alignment_span, n_records, global_record_counter, bases,
                        block_count, landmarks):
    """ContainerHeaderIO.writeContainerHeader(3, ...) [ContainerHeaderIO.java:69-84]: int32 (little endian) size of the blocks,
    ITF8 / LTF8 fields, the landmarks, and the CRC32 of those bytes, little endian"""
    import struct
    import zlib
    b = struct.pack('<i', byte_size) + itf8(sequence_id) + itf8(alignment_start) + itf8(alignment_span) + itf8(n_records)
    b += ltf8(global_record_counter) + ltf8(bases) + itf8(block_count) + cram_array(landmarks)
    return b + struct.pack('<I', zlib.crc32(b) & 0xFFFFFFFF)


def container_v3(slice_hdr, global_record_counter, compression_header, content_ids, ref_md5, tags, core, external):
    """ContainerIO.writeContainer [ContainerIO.java:152-209] of a container with ONE slice (how cramtools builds them: the
    records per slice equal the records per container): container header | compression header block | slice header block |
    core block | external blocks.  slice_hdr: dict with sequence_id, alignment_start, alignment_span, n_records, bases;
    core = (method, raw_len, data); external = [(content_id, method, raw_len, data)] in the order of content_ids.
    Container fields: ContainerFactory.buildContainer [:74-100] and calculateAlignmentBoundaries [:108-123]; blockCount and
    the landmark as writeContainer counts them [:183-197]."""
    RAW = BLOCK_METHOD['RAW']
    chb = cram_block_v3(RAW, BLOCK_CONTENT['COMPRESSION_HEADER'], 0, len(compression_header), compression_header)
    shc = slice_header_content_v3(slice_hdr['sequence_id'], slice_hdr['alignment_start'], slice_hdr['alignment_span'],
                                  slice_hdr['n_records'], global_record_counter, content_ids, ref_md5, tags)
    blocks = chb + cram_block_v3(RAW, BLOCK_CONTENT['MAPPED_SLICE'], 0, len(shc), shc)
    blocks += cram_block_v3(core[0], BLOCK_CONTENT['CORE'], 0, core[1], core[2])
    for cid, method, raw_len, data in external:
        blocks += cram_block_v3(method, BLOCK_CONTENT['EXTERNAL'], cid, raw_len, data)
    seq = slice_hdr['sequence_id']                       # the container starts at -1 and takes the slice's id unless that is -1
    if seq != -1:
        start, span = slice_hdr['alignment_start'], slice_hdr['alignment_span']       # min start, max(start + span) - min start
    else:
        start, span = -1, 0                              # Container's initial values [Container.java:28-30]
    hdr = container_header_v3(len(blocks), seq, start, span, slice_hdr['n_records'], global_record_counter, slice_hdr['bases'],
                              3 + len(external), [len(chb)])
    return hdr + blocks


# ---------------------------------------------------------------------------------------------- decode: restore pairing (inverse of set_mate_info)
def restore_mates(flag, cf, nf, seq_id, pos, mate_flags, mate_seq_id, mate_pos, tlen, aln_end):
    """The pairing part of CramNormalizer.normalize (src/main/java/htsjdk/samtools/cram/build/CramNormalizer.java:67-102, 236-270)
    for the records of one container as the reader left them, in order: next / previous links from the compression flags
    (cf: 2 = detached, 4 = mate downstream) and NF, then restoreMateInfo along every chain that starts at a record nobody
    points to -- setNextMate (mate start, mate sequence, and in the record's MATE flags [CramCompressionRecord.java:270-283] bit
    2 = mate unmapped, bit 1 = mate on the reverse strand: the SAM flag bits 0x8 / 0x20 are only derived from them when the
    SAMRecord is built) and the template size of the chain's first and last record (computeInsertSize [:455-470]).
    aln_end: CramCompressionRecord.getAlignmentEnd per record.
    -> dict(mate_flags, pos, mate_seq_id, mate_pos, tlen, nf, next, previous); raises IndexError where records.get() would throw"""
    n = len(flag)
    flag = [int(x) for x in flag]; pos = [int(x) for x in pos]; seq = [int(x) for x in seq_id]
    mseq = [int(x) for x in mate_seq_id]; mpos = [int(x) for x in mate_pos]; ts = [int(x) for x in tlen]; nf = [int(x) for x in nf]
    mf = [int(x) for x in mate_flags]
    nxt = [-1] * n; prv = [-1] * n
    end = [int(x) for x in aln_end]
    for i in range(n):
        if seq[i] == -1:
            if not (flag[i] & 4):
                end[i] -= pos[i]                                             # getAlignmentEnd is asked after the start was cleared
            pos[i] = 0                                                       # [:69-72]
    for i in range(n):
        if not (flag[i] & 1) or (int(cf[i]) & 2):                           # !isMultiFragment() || isDetached() [:81-87]
            nf[i] = -1; nxt[i] = -1; prv[i] = -1
            continue
        if int(cf[i]) & 4:                                                   # isHasMateDownStream() [:88-93]
            j = i + 1 + nf[i]
            if j < 0 or j >= n:
                raise IndexError('records.get(%d)' % j)
            nxt[i] = j; prv[j] = i

    def set_next_mate(r, m):                                                 # [:259-266]
        mpos[r] = pos[m]
        mf[r] = (mf[r] | 2) if flag[m] & 0x4 else (mf[r] & ~2)           # setMateUnmapped(next.isSegmentUnmapped())
        mf[r] = (mf[r] | 1) if flag[m] & 0x10 else (mf[r] & ~1)          # setMateNegativeStrand(next.isNegativeStrand())
        mseq[r] = seq[m]
        if mseq[r] == -1:
            mpos[r] = 0

    def insert_size(a, b):
        if (flag[a] & 4) or (flag[b] & 4) or seq[a] != seq[b]:
            return 0
        pa = end[a] if flag[a] & 0x10 else pos[a]
        pb = end[b] if flag[b] & 0x10 else pos[b]
        return pb - pa + (1 if pb >= pa else -1)
    for r in range(n):                                                       # [:95-101], restoreMateInfo [:236-257]
        if prv[r] != -1 or nxt[r] == -1:
            continue
        cur = r
        while nxt[cur] != -1:
            set_next_mate(cur, nxt[cur]); cur = nxt[cur]
        set_next_mate(cur, r)
        tl = insert_size(r, cur)
        ts[r] = tl; ts[cur] = -tl
    return dict(mate_flags=mf, pos=pos, mate_seq_id=mseq, mate_pos=mpos, tlen=ts, nf=nf, next=nxt, previous=prv)
