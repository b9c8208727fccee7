"""Host-side mirror of the reference's record-feature plug-in points, on top of the C ABI.

Mirrors, with the reference's names:
  QualityScorePreservation.lossyFromSpec / parsePolicies   (lossy/QualityScorePreservation.java:64,167-191)
  CramSerilization.convert's per-record loop               (cram/CramSerilization.java:144-166) -> features_encode
  CramNormalizer.restoreBases / restoreQualityScores       (cram/build/CramNormalizer.java:208-244,287-333) -> normalize_decode
Records are passed as structure-of-arrays dicts (the layout cramtools_b200.synth.reads produces), which is what the
Java shim of INTEGRATION.md builds from List<SAMRecord> before calling the same exported symbols.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import CramGpuError

u8p, u16p, u32p, i32p, i64p, u64p = (C.POINTER(t) for t in (C.c_uint8, C.c_uint16, C.c_uint32, C.c_int32, C.c_int64, C.c_uint64))

SERIES = ['FN', 'FC', 'FP', 'BS', 'DL', 'BA', 'SC', 'QS', 'RS', 'HC', 'PD', 'MQ']
FEATURES_ON_DEVICE_ONLY = 0x80000000      # cramgpu_features_encode_resident: host_series flag
SERIES_CONTENT_ID = {'FN': 15, 'FC': 3, 'FP': 2, 'BS': 16, 'DL': 5, 'BA': 6, 'SC': 27, 'QS': 4, 'RS': 26, 'HC': 28, 'PD': 29, 'MQ': 11}
NONE = np.uint64(0xFFFFFFFFFFFFFFFF)


class QsPolicy(C.Structure):
    _fields_ = [('read_cat', C.c_int32), ('read_param', C.c_int32), ('n_base_cats', C.c_int32),
                ('base_cat', C.c_int32 * 8), ('treatment', C.c_int32), ('treat_param', C.c_int32)]


class Reads(C.Structure):
    _fields_ = [('n_reads', C.c_uint64), ('n_containers', C.c_uint32), ('cont_off', u64p), ('ref_off', u64p), ('ref_len', u64p),
                ('ref', C.c_void_p), ('ref_total', C.c_uint64), ('bases', C.c_void_p), ('quals', C.c_void_p), ('read_off', C.c_void_p),
                ('cigar', C.c_void_p), ('cigar_off', C.c_void_p), ('pos', C.c_void_p), ('flag', C.c_void_p), ('mapq', C.c_void_p),
                ('rflags', C.c_void_p)]


class Features(C.Structure):
    _fields_ = [('n_feat', C.c_uint64), ('f_op', C.c_void_p), ('f_pos', C.c_void_p), ('f_val', C.c_void_p), ('f_aux', C.c_void_p),
                ('feat_off', C.c_void_p), ('payload', C.c_void_p), ('payload_len', C.c_uint64), ('force_qs', C.c_void_p),
                ('matrix', C.c_void_p), ('streams', C.c_void_p), ('streams_len', C.c_uint64), ('stream_off', C.c_void_p),
                ('stream_len', C.c_void_p), ('read_qs_off', C.c_void_p), ('read_ba_off', C.c_void_p)]


class NormIn(C.Structure):
    _fields_ = [('n_reads', C.c_uint64), ('n_containers', C.c_uint32), ('cont_off', u64p), ('win_off', u64p), ('win_len', u64p),
                ('win_start0', i64p), ('ref', C.c_void_p), ('ref_total', C.c_uint64), ('matrix', C.c_void_p), ('pos', C.c_void_p),
                ('read_off', C.c_void_p), ('rflags', C.c_void_p), ('feat_off', C.c_void_p), ('f_op', C.c_void_p), ('f_pos', C.c_void_p),
                ('f_val', C.c_void_p), ('f_aux', C.c_void_p), ('payload', C.c_void_p), ('qs', C.c_void_p), ('read_qs_off', C.c_void_p),
                ('ba', C.c_void_p), ('read_ba_off', C.c_void_p), ('payload_len', C.c_uint64)]


class Series(C.Structure):
    _fields_ = [('n_reads', C.c_uint64), ('n_containers', C.c_uint32), ('cont_off', u64p), ('streams', C.c_void_p),
                ('streams_len', C.c_uint64), ('stream_off', u64p), ('stream_len', u64p), ('read_off', C.c_void_p),
                ('rflags', C.c_void_p)]


HOST_SERIES = ['BF', 'CF', 'RI', 'RL', 'AP', 'RG', 'RN', 'MF', 'NS', 'NP', 'TS', 'NF', 'TL']
HOST_SERIES_CONTENT_ID = {'BF': 0, 'CF': 21, 'RI': 25, 'RL': 9, 'AP': 1, 'RG': 10, 'RN': 12, 'MF': 19, 'NS': 20, 'NP': 13, 'TS': 14,
                          'NF': 8, 'TL': 24}


class HostFields(C.Structure):
    _fields_ = [('n_reads', C.c_uint64), ('n_containers', C.c_uint32), ('cont_off', u64p), ('slice_start', i32p), ('multi_ref', u8p),
                ('ap_delta', C.c_uint32), ('names_included', C.c_uint32), ('flag', C.c_void_p), ('cf', C.c_void_p),
                ('mate_cf', C.c_void_p), ('force_qs', C.c_void_p), ('rflags', C.c_void_p), ('ref_id', C.c_void_p), ('pos', C.c_void_p),
                ('read_off', C.c_void_p), ('read_group', C.c_void_p), ('names', C.c_void_p), ('name_off', C.c_void_p),
                ('mate_ref_id', C.c_void_p), ('mate_pos', C.c_void_p), ('tlen', C.c_void_p), ('nf', C.c_void_p), ('tl', C.c_void_p)]


class HostSeries(C.Structure):
    _fields_ = [('streams', C.c_void_p), ('streams_len', C.c_uint64), ('stream_off', C.c_void_p), ('stream_len', C.c_void_p)]


_lib.SIGNATURES.update({
    'cramgpu_host_series_pack': (C.c_int, [C.c_void_p, C.POINTER(HostFields), C.POINTER(HostSeries)]),
    'cramgpu_host_series_pack_resident': (C.c_int, [C.c_void_p, C.POINTER(HostFields), C.c_uint32, C.POINTER(HostSeries),
                                                    C.POINTER(HostSeries)]),
    'cramgpu_host_series_pack_dev': (C.c_int, [C.c_void_p, C.POINTER(HostFields), C.POINTER(HostSeries), u64p, u64p, C.c_void_p]),
    'cramgpu_features_unpack': (C.c_int, [C.c_void_p, C.POINTER(Series), C.POINTER(Features), u8p, i32p]),
    'cramgpu_features_unpack_resident': (C.c_int, [C.c_void_p, C.POINTER(Series), C.POINTER(Features), u8p, i32p]),
    'cramgpu_features_unpack_dev': (C.c_int, [C.c_void_p, C.POINTER(Series), C.POINTER(Features), C.POINTER(C.c_void_p), i32p,
                                              C.c_void_p]),
    'cramgpu_parse_lossy_spec': (C.c_int, [C.c_char_p, C.POINTER(QsPolicy), C.c_int]),
    'cramgpu_features_encode': (C.c_int, [C.c_void_p, C.POINTER(Reads), C.POINTER(QsPolicy), C.c_int, C.POINTER(Features)]),
    'cramgpu_features_encode_dev': (C.c_int, [C.c_void_p, C.POINTER(Reads), C.POINTER(QsPolicy), C.c_int, C.POINTER(Features),
                                              u64p, u64p, C.c_void_p]),
    'cramgpu_features_encode_resident': (C.c_int, [C.c_void_p, C.POINTER(Reads), C.POINTER(QsPolicy), C.c_int, C.c_uint32,
                                                   C.POINTER(Features), C.POINTER(Features)]),
    'cramgpu_normalize_decode': (C.c_int, [C.c_void_p, C.POINTER(NormIn), u8p, u8p, u8p, i32p]),
    'cramgpu_normalize_decode_dev': (C.c_int, [C.c_void_p, C.POINTER(NormIn), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                               C.c_void_p]),
})


class QualityScorePreservation(object):
    """lossy/QualityScorePreservation.java: spec -> policy list (sorted DROP, BIN, PRESERVE)"""

    def __init__(self, policies, n):
        self.policies, self.n = policies, n

    @staticmethod
    def lossyFromSpec(spec):
        lib = _lib.load()
        arr = (QsPolicy * 16)()
        n = lib.cramgpu_parse_lossy_spec(spec.encode(), arr, 16)
        if n == _lib.ERR_ARG:
            raise CramGpuError(n, 'Unknown read or base category')
        if n == _lib.ERR_UNSUPPORTED:
            raise CramGpuError(n, 'The quality score lossy model requires reference tracks.')
        return QualityScorePreservation(arr, n)

    def getPreservationPolicies(self):
        return [self.policies[i] for i in range(self.n)]


def _policy_args(spec):
    """None / 'lossless' -> lossless; else parse"""
    if spec is None or spec == 'lossless':
        return None, -1
    q = spec if isinstance(spec, QualityScorePreservation) else QualityScorePreservation.lossyFromSpec(spec)
    return q.policies, q.n


def _vp(a):
    return C.c_void_p(a.ctypes.data) if a is not None else None


def _arr(x, dt, minlen=1):
    if isinstance(x, (bytes, bytearray)):
        x = np.frombuffer(bytes(x), dtype=np.uint8)
    a = np.ascontiguousarray(x, dtype=dt)
    if a.size < minlen:
        a = np.concatenate([a, np.zeros(minlen - a.size, dt)])
    return a


def _view(ptr, n, dt):
    n = int(n)
    if n == 0 or not ptr:
        return np.zeros(0, dt)
    buf = (C.c_uint8 * (n * np.dtype(dt).itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dt, count=n).copy()


class EncodedFeatures(object):
    """host copy of a cramgpu_features result"""

    def __init__(self, f, n_reads, n_containers):
        nf = int(f.n_feat)
        self.n_feat = nf
        self.f_op = _view(f.f_op, nf, np.uint8); self.f_pos = _view(f.f_pos, nf, np.int32)
        self.f_val = _view(f.f_val, nf, np.int32); self.f_aux = _view(f.f_aux, nf, np.int32)
        self.feat_off = _view(f.feat_off, n_reads + 1, np.uint64)
        if self.feat_off.size == 0:
            self.feat_off = np.zeros(n_reads + 1, np.uint64)
        self.payload = _view(f.payload, f.payload_len, np.uint8)
        self.force_qs = _view(f.force_qs, n_reads, np.uint8)
        def table(ptr, cols, dt):
            v = _view(ptr, cols * n_containers, dt)
            return v.reshape(n_containers, cols) if v.size else np.zeros((n_containers, cols), dt)
        self.matrix = table(f.matrix, 8, np.uint8)
        self.streams = _view(f.streams, f.streams_len, np.uint8)
        self.stream_off = table(f.stream_off, 12, np.uint64)
        self.stream_len = table(f.stream_len, 12, np.uint64)
        self.read_qs_off = _view(f.read_qs_off, n_reads, np.uint64)
        self.read_ba_off = _view(f.read_ba_off, n_reads, np.uint64)

    def series(self, container, name):
        k = SERIES.index(name)
        o, ln = int(self.stream_off[container, k]), int(self.stream_len[container, k])
        return self.streams[o:o + ln].tobytes()


def _reads_struct(rd, ref, cont_off, ref_off, ref_len, keep):
    n = len(rd['pos'])
    cont_off = _arr([0, n] if cont_off is None else cont_off, np.uint64)
    nc = cont_off.size - 1
    ref = _arr(ref, np.uint8)
    ref_off = _arr(np.zeros(nc) if ref_off is None else ref_off, np.uint64)
    ref_len = _arr(np.full(nc, ref.size) if ref_len is None else ref_len, np.uint64)
    arrs = dict(bases=_arr(rd['bases'], np.uint8), quals=_arr(rd['quals'], np.uint8), read_off=_arr(rd['read_off'], np.uint64),
                cigar=_arr(rd['cigar'], np.uint32), cigar_off=_arr(rd['cigar_off'], np.uint64), pos=_arr(rd['pos'], np.int32),
                flag=_arr(rd['flag'], np.uint16), mapq=_arr(rd['mapq'], np.uint8),
                rflags=_arr(rd['rflags'], np.uint8) if rd.get('rflags') is not None else None)
    keep.extend([cont_off, ref_off, ref_len, ref] + list(arrs.values()))
    R = Reads()
    R.n_reads = n; R.n_containers = nc
    R.cont_off = cont_off.ctypes.data_as(u64p); R.ref_off = ref_off.ctypes.data_as(u64p); R.ref_len = ref_len.ctypes.data_as(u64p)
    R.ref = _vp(ref); R.ref_total = ref.size
    for k, v in arrs.items():
        setattr(R, k, _vp(v))
    return R, n, nc


def features_encode(rd, ref, spec=None, cont_off=None, ref_off=None, ref_len=None, ctx=None, allow_bin_fault=False):
    """Feature extraction + quality preservation + series packing through cramgpu_features_encode (host buffers).
    rd: SoA dict; ref: reference bytes; spec: None/'lossless' or a lossy model string such as '*8', 'm5', 'N40-D8'.
    A BIN treatment that meets a score of 110..255 fails the call with CRAMGPU_ERR_UNSUPPORTED (the reference throws
    ArrayIndexOutOfBoundsException, Binning.java:63-71); allow_bin_fault=True returns the results anyway, bit 7 of
    force_qs marking the records."""
    ctx = ctx or _lib.default_context()
    keep = []
    R, n, nc = _reads_struct(rd, ref, cont_off, ref_off, ref_len, keep)
    pol, npol = _policy_args(spec)
    F = Features()
    rc = ctx.lib.cramgpu_features_encode(ctx.handle, C.byref(R), pol, npol, C.byref(F))
    if not (allow_bin_fault and rc == _lib.ERR_UNSUPPORTED):
        ctx.check(rc)
    return EncodedFeatures(F, n, nc)


def features_encode_resident(rd, ref, spec=None, host_series=None, cont_off=None, ref_off=None, ref_len=None, ctx=None,
                             features_on_device_only=False):
    """cramgpu_features_encode_resident: as features_encode, but the results also stay on the device and only the series
    named in host_series (default: all but QS and BA) come back to the host.  -> (EncodedFeatures, Features with DEVICE
    pointers); the device result is valid until the next record call on ctx."""
    ctx = ctx or _lib.default_context()
    keep = []
    R, n, nc = _reads_struct(rd, ref, cont_off, ref_off, ref_len, keep)
    pol, npol = _policy_args(spec)
    if host_series is None:
        host_series = [s for s in SERIES if s not in ('QS', 'BA')]
    mask = 0
    for name in host_series:
        mask |= 1 << SERIES.index(name)
    if features_on_device_only:
        mask |= FEATURES_ON_DEVICE_ONLY
    F, D = Features(), Features()
    ctx.check(ctx.lib.cramgpu_features_encode_resident(ctx.handle, C.byref(R), pol, npol, mask, C.byref(F), C.byref(D)))
    return EncodedFeatures(F, n, nc), D


class UnpackedFeatures(object):
    """host copy of a cramgpu_features_unpack result"""

    def __init__(self, f, n_reads, streams, mapq, status):
        nf = int(f.n_feat)
        self.n_feat = nf
        self.f_op = _view(f.f_op, nf, np.uint8); self.f_pos = _view(f.f_pos, nf, np.int32)
        self.f_val = _view(f.f_val, nf, np.int32); self.f_aux = _view(f.f_aux, nf, np.int32)
        self.feat_off = _view(f.feat_off, n_reads + 1, np.uint64)
        self.payload = _view(f.payload, f.payload_len, np.uint8)
        self.read_qs_off = _view(f.read_qs_off, n_reads, np.uint64)
        self.read_ba_off = _view(f.read_ba_off, n_reads, np.uint64)
        self.streams, self.mapq, self.status = streams, mapq, status


def features_unpack(streams, stream_off, stream_len, read_off, rflags, cont_off=None, ctx=None, check=True):
    """CramRecordReader.read's feature part through cramgpu_features_unpack (host buffers): decoded series -> feature arrays.
    stream_off / stream_len: [n_containers][12]; rflags: CRAMGPU_NF_* per read.  -> UnpackedFeatures"""
    ctx = ctx or _lib.default_context()
    n = len(rflags)
    cont_off = _arr([0, n] if cont_off is None else cont_off, np.uint64)
    nc = cont_off.size - 1
    streams = _arr(streams, np.uint8)
    so = _arr(np.asarray(stream_off, np.uint64).reshape(nc, 12), np.uint64); sl = _arr(np.asarray(stream_len, np.uint64).reshape(nc, 12), np.uint64)
    ro = _arr(read_off, np.uint64); rf = _arr(rflags, np.uint8)
    I = Series()
    I.n_reads = n; I.n_containers = nc
    I.cont_off = cont_off.ctypes.data_as(u64p); I.streams = _vp(streams); I.streams_len = streams.size
    I.stream_off = so.ctypes.data_as(u64p); I.stream_len = sl.ctypes.data_as(u64p); I.read_off = _vp(ro); I.rflags = _vp(rf)
    F = Features()
    mapq = np.zeros(max(n, 1), np.uint8); status = np.zeros(max(nc, 1), np.int32)
    rc = ctx.lib.cramgpu_features_unpack(ctx.handle, C.byref(I), C.byref(F), mapq.ctypes.data_as(u8p), status.ctypes.data_as(i32p))
    if rc != _lib.OK and (check or rc not in (_lib.ERR_CORRUPT, _lib.ERR_UNSUPPORTED)):
        ctx.check(rc)
    return UnpackedFeatures(F, n, streams, mapq[:n], status[:nc])


def normalize_decode(pos, read_off, rflags, enc, window, window_start0, matrix, cont_off=None, win_off=None, win_len=None,
                     qs=None, read_qs_off=None, ba=None, read_ba_off=None, ctx=None, check=True):
    """CramNormalizer base + quality restoration through cramgpu_normalize_decode (host buffers).
    enc supplies the feature arrays (f_op, f_pos, f_val, f_aux, feat_off, payload); window/window_start0 describe the
    reference region(s).  Returns (bases u8, quals u8, qual_star u8, status i32)."""
    ctx = ctx or _lib.default_context()
    n = len(pos)
    cont_off = _arr([0, n] if cont_off is None else cont_off, np.uint64)
    nc = cont_off.size - 1
    window = _arr(window, np.uint8)
    win_off = _arr(np.zeros(nc) if win_off is None else win_off, np.uint64)
    win_len = _arr(np.full(nc, window.size) if win_len is None else win_len, np.uint64)
    ws0 = _arr(np.full(nc, window_start0) if np.isscalar(window_start0) else window_start0, np.int64)
    matrix = _arr(np.asarray(matrix, np.uint8).reshape(nc, -1)[:, :8] if np.asarray(matrix).size == 8 * nc else
                  np.pad(np.asarray(matrix, np.uint8).reshape(nc, -1), ((0, 0), (0, 8 - np.asarray(matrix).size // nc))), np.uint8)
    a = dict(pos=_arr(pos, np.int32), read_off=_arr(read_off, np.uint64), rflags=_arr(rflags, np.uint8),
             feat_off=_arr(enc.feat_off, np.uint64), f_op=_arr(enc.f_op, np.uint8), f_pos=_arr(enc.f_pos, np.int32),
             f_val=_arr(enc.f_val, np.int32), f_aux=_arr(enc.f_aux, np.int32), payload=_arr(enc.payload, np.uint8),
             qs=_arr(qs if qs is not None else [], np.uint8), ba=_arr(ba if ba is not None else [], np.uint8),
             read_qs_off=_arr(read_qs_off, np.uint64) if read_qs_off is not None else None,
             read_ba_off=_arr(read_ba_off, np.uint64) if read_ba_off is not None else None)
    I = NormIn()
    I.n_reads = n; I.n_containers = nc
    I.cont_off = cont_off.ctypes.data_as(u64p); I.win_off = win_off.ctypes.data_as(u64p); I.win_len = win_len.ctypes.data_as(u64p)
    I.win_start0 = ws0.ctypes.data_as(i64p); I.ref = _vp(window); I.ref_total = window.size; I.matrix = _vp(matrix)
    for k, v in a.items():
        setattr(I, k, _vp(v))
    I.payload_len = a['payload'].size
    total = int(a['read_off'][n])
    bases = np.zeros(max(total, 1), np.uint8); quals = np.zeros(max(total, 1), np.uint8)
    star = np.zeros(max(n, 1), np.uint8); status = np.zeros(max(n, 1), np.int32)
    rc = ctx.lib.cramgpu_normalize_decode(ctx.handle, C.byref(I), bases.ctypes.data_as(u8p), quals.ctypes.data_as(u8p),
                                          star.ctypes.data_as(u8p), status.ctypes.data_as(i32p))
    if rc != _lib.OK and (check or rc != _lib.ERR_CORRUPT):
        ctx.check(rc)
    return bases[:total], quals[:total], star[:n], status[:n]


class PackedHostSeries(object):
    """result of host_series_pack: series(c, name) -> bytes"""

    def __init__(self, hs, nc):
        self.streams = _view(hs.streams, hs.streams_len, np.uint8)
        self.stream_off = _view(hs.stream_off, nc * len(HOST_SERIES), np.uint64).reshape(nc, len(HOST_SERIES))
        self.stream_len = _view(hs.stream_len, nc * len(HOST_SERIES), np.uint64).reshape(nc, len(HOST_SERIES))

    def series(self, c, name):
        k = HOST_SERIES.index(name)
        o, ln = int(self.stream_off[c, k]), int(self.stream_len[c, k])
        return self.streams[o:o + ln].tobytes()


def host_series_pack(flag, pos, read_off, cont_off=None, slice_start=None, multi_ref=None, ap_delta=True, names_included=False,
                     cf=None, mate_cf=None, force_qs=None, rflags=None, ref_id=None, read_group=None, names=None,
                     mate_ref_id=None, mate_pos=None, tlen=None, nf=None, tl=None, ctx=None):
    """The fields Writer.write emits in front of the read features [Writer.java:139-172], through cramgpu_host_series_pack
    (host buffers).  names: list of bytes / str.  -> PackedHostSeries"""
    ctx = ctx or _lib.default_context()
    n = len(flag)
    cont_off = _arr([0, n] if cont_off is None else cont_off, np.uint64)
    nc = cont_off.size - 1
    keep = dict(flag=_arr(flag, np.uint16), pos=_arr(pos, np.int32), read_off=_arr(read_off, np.uint64))
    opt = dict(cf=(cf, np.uint8), mate_cf=(mate_cf, np.uint8), force_qs=(force_qs, np.uint8), rflags=(rflags, np.uint8),
               ref_id=(ref_id, np.int32), read_group=(read_group, np.int32), mate_ref_id=(mate_ref_id, np.int32),
               mate_pos=(mate_pos, np.int32), tlen=(tlen, np.int32), nf=(nf, np.int32), tl=(tl, np.int32))
    for k, (v, dt) in opt.items():
        keep[k] = None if v is None else _arr(v, dt)
    if names is not None:
        nb = [x.encode('latin-1') if isinstance(x, str) else bytes(x) for x in names]
        keep['names'] = _arr(b''.join(nb), np.uint8)
        keep['name_off'] = np.concatenate([[0], np.cumsum([len(x) for x in nb])]).astype(np.uint64)
    else:
        keep['names'] = keep['name_off'] = None
    ss = _arr(np.zeros(nc) if slice_start is None else slice_start, np.int32)
    mr = None if multi_ref is None else _arr(multi_ref, np.uint8)
    I = HostFields()
    I.n_reads = n; I.n_containers = nc
    I.cont_off = cont_off.ctypes.data_as(u64p); I.slice_start = ss.ctypes.data_as(i32p)
    I.multi_ref = mr.ctypes.data_as(u8p) if mr is not None else None
    I.ap_delta = 1 if ap_delta else 0; I.names_included = 1 if names_included else 0
    for k, v in keep.items():
        setattr(I, k, _vp(v))
    H = HostSeries()
    ctx.check(ctx.lib.cramgpu_host_series_pack(ctx.handle, C.byref(I), C.byref(H)))
    return PackedHostSeries(H, nc)


# ---------------------------------------------------------------------------------------------- slice headers and digests
class SliceIn(C.Structure):
    _fields_ = [('n_reads', C.c_uint64), ('n_containers', C.c_uint32), ('cont_off', u64p), ('bases', C.c_void_p), ('scores', C.c_void_p),
                ('read_off', C.c_void_p), ('rflags', C.c_void_p), ('ref_id', C.c_void_p), ('pos', C.c_void_p), ('flag', C.c_void_p),
                ('f_op', C.c_void_p), ('f_pos', C.c_void_p), ('f_val', C.c_void_p), ('f_aux', C.c_void_p), ('feat_off', C.c_void_p),
                ('force_qs', C.c_void_p), ('read_qs_off', C.c_void_p), ('streams', C.c_void_p),
                ('ref', C.c_void_p), ('ref_total', C.c_uint64), ('ref_off', u64p), ('ref_len', u64p)]


class SliceHeader(C.Structure):
    _fields_ = [('sequence_id', C.c_int32), ('alignment_start', C.c_int32), ('alignment_span', C.c_int32), ('n_records', C.c_uint32),
                ('bases', C.c_uint64), ('bd', C.c_uint32), ('sd', C.c_uint32), ('ref_md5', C.c_uint8 * 16)]


_lib.SIGNATURES.update({
    'cramgpu_slice_headers': (C.c_int, [C.c_void_p, C.POINTER(SliceIn), C.POINTER(SliceHeader)]),
    'cramgpu_slice_headers_dev': (C.c_int, [C.c_void_p, C.POINTER(SliceIn), C.c_void_p, C.c_void_p]),
    'cramgpu_slice_sha_tags': (None, [u8p, u8p]),
})


def slice_sha_tags():
    """(B1 = S1, B5 = S5): the constant SHA tags htsjdk 2.4.1 writes (include/cramgpu.h)"""
    lib = _lib.load()
    a, b = np.zeros(20, np.uint8), np.zeros(64, np.uint8)
    lib.cramgpu_slice_sha_tags(a.ctypes.data_as(u8p), b.ctypes.data_as(u8p))
    return a.tobytes(), b.tobytes()


def slice_headers(bases, read_off, pos, flag, f_op, f_val, f_aux, feat_off, scores=None, rflags=None, ref_id=None, cont_off=None,
                  f_pos=None, force_qs=None, read_qs_off=None, streams=None, ref=None, ref_off=None, ref_len=None, ctx=None):
    """ContainerFactory.buildSlice's header fields and BD / SD digests [ContainerFactory.java:138-176] for every slice of a
    batch, through cramgpu_slice_headers (host buffers; `streams` / force_qs / read_qs_off: device pointers of a feature
    call when scores is None).  -> list of dicts with sequence_id, alignment_start, alignment_span, n_records, bases, tags"""
    import struct
    ctx = ctx or _lib.default_context()
    n = len(pos)
    cont_off = _arr([0, n] if cont_off is None else cont_off, np.uint64)
    nc = cont_off.size - 1
    keep = dict(bases=_arr(bases, np.uint8), read_off=_arr(read_off, np.uint64), pos=_arr(pos, np.int32), flag=_arr(flag, np.uint16),
                f_op=_arr(f_op, np.uint8), f_val=_arr(f_val, np.int32), f_aux=_arr(f_aux, np.int32), feat_off=_arr(feat_off, np.uint64),
                scores=None if scores is None else _arr(scores, np.uint8), rflags=None if rflags is None else _arr(rflags, np.uint8),
                ref_id=None if ref_id is None else _arr(ref_id, np.int32), f_pos=None if f_pos is None else _arr(f_pos, np.int32))
    I = SliceIn()
    I.n_reads = n; I.n_containers = nc; I.cont_off = cont_off.ctypes.data_as(u64p)
    for k, v in keep.items():
        setattr(I, k, _vp(v))
    for k, v in (('force_qs', force_qs), ('read_qs_off', read_qs_off), ('streams', streams)):
        setattr(I, k, v if (v is None or isinstance(v, int)) else _vp(v))
    if ref is not None:
        refa = _arr(ref, np.uint8)
        roff = _arr(np.zeros(nc) if ref_off is None else ref_off, np.uint64); rlen = _arr(np.full(nc, refa.size) if ref_len is None else ref_len, np.uint64)
        I.ref = _vp(refa); I.ref_total = refa.size; I.ref_off = roff.ctypes.data_as(u64p); I.ref_len = rlen.ctypes.data_as(u64p)
    out = (SliceHeader * max(nc, 1))()
    ctx.check(ctx.lib.cramgpu_slice_headers(ctx.handle, C.byref(I), out))
    s1, s5 = slice_sha_tags()
    return [dict(sequence_id=h.sequence_id, alignment_start=h.alignment_start, alignment_span=h.alignment_span, n_records=h.n_records,
                 bases=h.bases, bd=h.bd, sd=h.sd, ref_md5=bytes(h.ref_md5),
                 tags={'BD': struct.pack('<I', h.bd), 'SD': struct.pack('<I', h.sd), 'B5': s5, 'S5': s5, 'B1': s1, 'S1': s1})
            for h in out[:nc]]


# ---------------------------------------------------------------------------------------------- mate linking
class MateIn(C.Structure):
    _fields_ = [('n_reads', C.c_uint64), ('n_containers', C.c_uint32), ('cont_off', u64p), ('coordinate_sorted', C.c_uint32),
                ('reserved', C.c_uint32), ('flag', C.c_void_p), ('names', C.c_void_p), ('name_off', C.c_void_p), ('ref_id', C.c_void_p),
                ('pos', C.c_void_p), ('mate_ref_id', C.c_void_p), ('mate_pos', C.c_void_p), ('tlen', C.c_void_p), ('read_off', C.c_void_p),
                ('f_op', C.c_void_p), ('f_val', C.c_void_p), ('f_aux', C.c_void_p), ('feat_off', C.c_void_p)]


_lib.SIGNATURES.update({
    'cramgpu_mate_link': (C.c_int, [C.c_void_p, C.POINTER(MateIn), C.c_void_p, C.c_void_p]),
    'cramgpu_mate_link_dev': (C.c_int, [C.c_void_p, C.POINTER(MateIn), C.c_void_p, C.c_void_p, C.c_void_p]),
})


def mate_link(flag, names, pos, mate_ref_id, mate_pos, tlen, read_off, f_op, f_val, f_aux, feat_off, ref_id=None, cont_off=None,
              coordinate_sorted=True, ctx=None):
    """CramSerilization.setMateInfo (CramSerilization.java:258-373) through cramgpu_mate_link (host buffers).
    names: list of bytes / str.  -> (mate_cf u8[n]: 2 = detached, 4 = mate downstream, 0 = linked second; nf i32[n])"""
    ctx = ctx or _lib.default_context()
    n = len(flag)
    cont_off = _arr([0, n] if cont_off is None else cont_off, np.uint64)
    nb = [x.encode('latin-1') if isinstance(x, str) else bytes(x) for x in names]
    keep = dict(flag=_arr(flag, np.uint16), names=_arr(b''.join(nb), np.uint8),
                name_off=np.concatenate([[0], np.cumsum([len(x) for x in nb])]).astype(np.uint64),
                ref_id=None if ref_id is None else _arr(ref_id, np.int32), pos=_arr(pos, np.int32), mate_ref_id=_arr(mate_ref_id, np.int32),
                mate_pos=_arr(mate_pos, np.int32), tlen=_arr(tlen, np.int32), read_off=_arr(read_off, np.uint64), f_op=_arr(f_op, np.uint8),
                f_val=_arr(f_val, np.int32), f_aux=_arr(f_aux, np.int32), feat_off=_arr(feat_off, np.uint64))
    I = MateIn()
    I.n_reads = n; I.n_containers = cont_off.size - 1; I.cont_off = cont_off.ctypes.data_as(u64p)
    I.coordinate_sorted = 1 if coordinate_sorted else 0
    for k, v in keep.items():
        setattr(I, k, _vp(v))
    cf = np.zeros(max(n, 1), np.uint8); nf = np.zeros(max(n, 1), np.int32)
    ctx.check(ctx.lib.cramgpu_mate_link(ctx.handle, C.byref(I), _vp(cf), _vp(nf)))
    return cf[:n], nf[:n]


# ---------------------------------------------------------------------------------------------- CIGAR rebuild + BAM records (device resident)
class BamRecords(C.Structure):
    _fields_ = [('n_reads', C.c_uint64), ('ref_id', C.c_void_p), ('pos', C.c_void_p), ('mate_ref_id', C.c_void_p), ('mate_pos', C.c_void_p),
                ('tlen', C.c_void_p), ('flag', C.c_void_p), ('bin', C.c_void_p), ('mapq', C.c_void_p), ('rflags', C.c_void_p),
                ('names', C.c_void_p), ('name_off', C.c_void_p), ('cigar', C.c_void_p), ('cigar_off', C.c_void_p), ('bases', C.c_void_p),
                ('quals', C.c_void_p), ('read_off', C.c_void_p), ('tags', C.c_void_p), ('tag_off', C.c_void_p),
                ('names_len', C.c_uint64), ('cigar_len', C.c_uint64), ('bases_len', C.c_uint64), ('tags_len', C.c_uint64)]


_lib.SIGNATURES.update({
    'cramgpu_cigar_rebuild_dev': (C.c_int, [C.c_void_p, C.c_uint64] + [C.c_void_p] * 7 + [C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                                                                         C.POINTER(C.c_uint64), C.c_void_p]),
    'cramgpu_bam_unpack_dev': (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(BamRecords), C.c_void_p, C.c_void_p]),
    'cramgpu_bam_pack_dev': (C.c_int, [C.c_void_p, C.POINTER(BamRecords), C.POINTER(C.c_void_p), C.POINTER(C.c_uint64), C.POINTER(C.c_void_p),
                                       C.c_void_p, C.c_void_p]),
})


_lib.SIGNATURES.update({
    'cramgpu_blocks_frame_dev': (C.c_int, [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, i32p, u8p, u8p,
                                           C.POINTER(C.c_void_p), C.POINTER(C.c_uint64), C.POINTER(C.c_void_p), C.c_void_p]),
})


# ---------------------------------------------------------------------------------------------- CRAM v3 containers (device resident)
CONTAINER_TAGS = 1


class ContainerIn(C.Structure):
    _fields_ = [('n_containers', C.c_uint32), ('flags', C.c_uint32), ('d_slices', C.c_void_p), ('global_record_counter', u64p),
                ('comp_header', C.c_void_p), ('comp_header_off', u64p), ('block_first', C.POINTER(C.c_uint32)), ('nblocks', C.c_uint32),
                ('d_comp', C.c_void_p), ('d_comp_off', C.c_void_p), ('d_comp_len', C.c_void_p), ('d_raw_len', C.c_void_p),
                ('content_id', i32p), ('method', u8p), ('content_type', u8p)]


_lib.SIGNATURES.update({
    'cramgpu_containers_assemble_dev': (C.c_int, [C.c_void_p, C.POINTER(ContainerIn), C.POINTER(C.c_void_p), C.POINTER(C.c_uint64),
                                                  C.POINTER(C.c_void_p), C.c_void_p]),
})


def slice_header_array(headers):
    """list of dicts (as slice_headers returns) -> numpy bytes of a cramgpu_slice_header array (to be placed in device memory)"""
    arr = (SliceHeader * max(len(headers), 1))()
    for h, d in zip(arr, headers):
        h.sequence_id = d['sequence_id']; h.alignment_start = d['alignment_start']; h.alignment_span = d['alignment_span']
        h.n_records = d['n_records']; h.bases = d['bases']; h.bd = d.get('bd', 0); h.sd = d.get('sd', 0)
        md5 = d.get('ref_md5') or bytes(16)
        for i in range(16):
            h.ref_md5[i] = md5[i]
    return np.frombuffer(bytes(arr), np.uint8)[:C.sizeof(SliceHeader) * len(headers)].copy()


def containers_assemble(d_slices, global_record_counter, comp_headers, block_first, d_comp, d_comp_off, d_comp_len, d_raw_len,
                        content_id, method, content_type, tags=True, ctx=None, stream=None):
    """ContainerIO.writeContainer for one-slice containers through cramgpu_containers_assemble_dev.  d_*: device pointers (ints);
    comp_headers: list of bytes (CompressionHeader.toByteArray() per container).  -> (device pointer, total bytes, device
    pointer of the [n + 1] container offsets); the result lives in the context until the next framing call"""
    ctx = ctx or _lib.default_context()
    n = len(comp_headers)
    grc = _arr(global_record_counter, np.uint64)
    ch = _arr(np.frombuffer(b''.join(comp_headers), np.uint8) if comp_headers else np.zeros(0, np.uint8), np.uint8)
    if ch.size == 0:
        ch = np.zeros(1, np.uint8)
    cho = np.concatenate([[0], np.cumsum([len(x) for x in comp_headers])]).astype(np.uint64)
    bf = _arr(block_first, np.uint32)
    cid = _arr(content_id, np.int32); m = _arr(method, np.uint8); t = _arr(content_type, np.uint8)
    I = ContainerIn()
    I.n_containers = n; I.flags = CONTAINER_TAGS if tags else 0; I.d_slices = d_slices
    I.global_record_counter = grc.ctypes.data_as(u64p); I.comp_header = _vp(ch); I.comp_header_off = cho.ctypes.data_as(u64p)
    I.block_first = bf.ctypes.data_as(C.POINTER(C.c_uint32)); I.nblocks = cid.size
    I.d_comp = d_comp; I.d_comp_off = d_comp_off; I.d_comp_len = d_comp_len; I.d_raw_len = d_raw_len
    I.content_id = cid.ctypes.data_as(i32p); I.method = m.ctypes.data_as(u8p); I.content_type = t.ctypes.data_as(u8p)
    out, olen, ooff = C.c_void_p(), C.c_uint64(), C.c_void_p()
    ctx.check(ctx.lib.cramgpu_containers_assemble_dev(ctx.handle, C.byref(I), C.byref(out), C.byref(olen), C.byref(ooff), stream))
    return out.value, olen.value, ooff.value


# ---------------------------------------------------------------------------------------------- coverage / mismatch tracks (selectors X and P)
_lib.SIGNATURES.update({
    'cramgpu_tracks_create': (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32, C.POINTER(C.c_void_p)]),
    'cramgpu_tracks_destroy': (None, [C.c_void_p, C.c_void_p]),
    'cramgpu_tracks_position': (C.c_int64, [C.c_void_p]),
    'cramgpu_tracks_update_dev': (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(Reads), C.POINTER(QsPolicy), C.c_int, i32p, C.c_void_p]),
})


class ReferenceTracks(object):
    """ref/ReferenceTracks.java for one sequence, on the device: carried from call to call like the reference carries its
    object from container to container.  update(rd, spec, cont_off) = what CramSerilization.convert does with the tracks for a
    batch of containers (span, ensureRange, updateTracks) plus the selector bytes the next features_encode call on the same
    context consumes."""

    def __init__(self, ref, window=0, ctx=None):
        self.ctx = ctx or _lib.default_context()
        ref = _arr(ref, np.uint8)
        h = C.c_void_p()
        self.ctx.check(self.ctx.lib.cramgpu_tracks_create(self.ctx.handle, _vp(ref), ref.size, int(window), C.byref(h)))
        self.h = h

    def close(self):
        if getattr(self, 'h', None):
            self.ctx.lib.cramgpu_tracks_destroy(self.ctx.handle, self.h); self.h = None

    __del__ = close

    @property
    def position(self):
        return int(self.ctx.lib.cramgpu_tracks_position(self.h))

    def update(self, rd, spec, cont_off=None):
        """-> per-container status list; raises CramGpuError (with .container_status) where the call fails"""
        import torch
        ctx = self.ctx
        n = len(rd['pos'])
        cont_off = _arr([0, n] if cont_off is None else cont_off, np.uint64)
        nc = cont_off.size - 1
        dev = 'cuda:%d' % ctx.device if hasattr(ctx, 'device') else 'cuda:0'
        host = dict(bases=_arr(rd['bases'], np.uint8), read_off=_arr(rd['read_off'], np.uint64), cigar=_arr(rd['cigar'], np.uint32),
                    cigar_off=_arr(rd['cigar_off'], np.uint64), pos=_arr(rd['pos'], np.int32), flag=_arr(rd['flag'], np.uint16))
        t = {k: torch.from_numpy(v.view(np.uint8).copy()).to(dev) for k, v in host.items()}
        torch.cuda.synchronize()
        R = Reads()
        R.n_reads = n; R.n_containers = nc; R.cont_off = cont_off.ctypes.data_as(u64p)
        for k, v in t.items():
            setattr(R, k, v.data_ptr())
        pol, npol = _policy_args(spec)
        status = np.zeros(max(nc, 1), np.int32)
        rc = ctx.lib.cramgpu_tracks_update_dev(ctx.handle, self.h, C.byref(R), pol, npol, status.ctypes.data_as(i32p), None)
        ctx.sync()
        self.container_status = status[:nc].tolist()
        try:
            ctx.check(rc)
        except CramGpuError as e:
            e.container_status = self.container_status
            raise
        return self.container_status


# ---------------------------------------------------------------------------------------------- decode: restore the pairing
class MateRestore(C.Structure):
    _fields_ = [('n_reads', C.c_uint64), ('n_containers', C.c_uint32), ('cont_off', u64p), ('flag', C.c_void_p), ('cf', C.c_void_p),
                ('mate_flags', C.c_void_p), ('nf', C.c_void_p), ('ref_id', C.c_void_p), ('pos', C.c_void_p), ('mate_ref_id', C.c_void_p), ('mate_pos', C.c_void_p),
                ('tlen', C.c_void_p), ('read_off', C.c_void_p), ('feat_off', C.c_void_p), ('f_op', C.c_void_p), ('f_val', C.c_void_p),
                ('f_aux', C.c_void_p)]


_lib.SIGNATURES.update({
    'cramgpu_mate_restore_dev': (C.c_int, [C.c_void_p, C.POINTER(MateRestore), C.c_void_p, C.c_void_p]),
})


def mate_restore(flag, cf, mate_flags, nf, ref_id, pos, mate_ref_id, mate_pos, tlen, read_off, f_op, f_val, f_aux, feat_off, cont_off=None, ctx=None):
    """The pairing part of CramNormalizer.normalize through cramgpu_mate_restore_dev (arrays staged to the device with torch).
    -> dict(mate_flags, nf, pos, mate_ref_id, mate_pos, tlen, status)"""
    import torch
    ctx = ctx or _lib.default_context()
    n = len(flag)
    cont_off = _arr([0, n] if cont_off is None else cont_off, np.uint64)
    nc = cont_off.size - 1
    dev = 'cuda:%d' % ctx.device
    host = dict(flag=_arr(flag, np.uint16), cf=_arr(cf, np.uint8), mate_flags=_arr(mate_flags, np.uint8), nf=_arr(nf, np.int32), ref_id=_arr(ref_id, np.int32), pos=_arr(pos, np.int32),
                mate_ref_id=_arr(mate_ref_id, np.int32), mate_pos=_arr(mate_pos, np.int32), tlen=_arr(tlen, np.int32),
                read_off=_arr(read_off, np.uint64), feat_off=_arr(feat_off, np.uint64), f_op=_arr(f_op, np.uint8), f_val=_arr(f_val, np.int32),
                f_aux=_arr(f_aux, np.int32))
    t = {k: torch.from_numpy(v.view(np.uint8).copy()).to(dev) for k, v in host.items()}
    st = torch.zeros(4 * max(nc, 1), dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()
    I = MateRestore()
    I.n_reads = n; I.n_containers = nc; I.cont_off = cont_off.ctypes.data_as(u64p)
    for k, v in t.items():
        setattr(I, k, v.data_ptr())
    ctx.check(ctx.lib.cramgpu_mate_restore_dev(ctx.handle, C.byref(I), st.data_ptr(), None))
    ctx.sync()
    out = {k: t[k].cpu().numpy().view(host[k].dtype)[:n].copy() for k in ('mate_flags', 'nf', 'pos', 'mate_ref_id', 'mate_pos', 'tlen')}
    out['status'] = st.cpu().numpy().view(np.int32)[:nc].copy()
    return out
