"""Seeded synthetic inputs of SURVEY.md section 8d (numpy only; no reference code involved).

Used by the tests, by oracle/make_golden.py and by bench.py so that the oracle, the
mini-JVM run of the reference jar and the GPU all see identical bytes.
"""
import numpy as np

ACGT = np.frombuffer(b'ACGT', dtype=np.uint8)


def reference(n, seed):
    """uniform random upper-case ACGT contig"""
    rng = np.random.default_rng(seed)
    return ACGT[rng.integers(0, 4, size=n, dtype=np.uint8)]


def qualities(n_reads, read_len, seed):
    """Illumina-like phred scores in [2, 41]: first-order Markov chain per read with a
    position dependent downward drift (high at 5', decaying at 3').  Returns (n_reads, read_len) u8."""
    rng = np.random.default_rng(seed)
    q = np.empty((n_reads, read_len), dtype=np.int16)
    cur = rng.choice(np.array([41, 40, 39, 38, 37, 35, 32], dtype=np.int16),
                     p=[0.45, 0.2, 0.12, 0.08, 0.07, 0.05, 0.03], size=n_reads)
    steps = np.array([-9, -4, -2, -1, 0, 1, 2, 4], dtype=np.int16)
    for i in range(read_len):
        q[:, i] = cur
        drift = 0.10 + 0.25 * (i / max(1, read_len - 1))
        p = np.array([0.01 + 0.03 * drift, 0.02 + 0.1 * drift, 0.05 + 0.2 * drift, 0.08 + 0.3 * drift,
                      0.0, 0.12, 0.05, 0.02])
        p[4] = 1.0 - p.sum()
        cur = np.clip(cur + steps[rng.choice(8, p=p, size=n_reads)], 2, 41).astype(np.int16)
    return q.astype(np.uint8)


def qs_stream(n, seed, read_len=150):
    """QS-like byte stream: raw phred bytes of consecutive reads (what Writer puts in series QS)"""
    nr = (n + read_len - 1) // read_len
    return np.ascontiguousarray(qualities(nr, read_len, seed).reshape(-1)[:n])


def ba_stream(n, seed):
    """BA-like byte stream: uniform ACGT with 1 % N"""
    rng = np.random.default_rng(seed)
    out = ACGT[rng.integers(0, 4, size=n, dtype=np.uint8)]
    out[rng.random(n) < 0.01] = ord('N')
    return out


def read_names(n_reads, seed):
    """list of bytes names SIM:1:FC01:<lane>:<tile>:<x>:<y>"""
    rng = np.random.default_rng(seed)
    lane = rng.integers(1, 9, size=n_reads)
    tile = rng.integers(1101, 2229, size=n_reads)
    x = rng.integers(1000, 30000, size=n_reads)
    y = rng.integers(1000, 200000, size=n_reads)
    return [b'SIM:1:FC01:%d:%d:%d:%d' % (lane[i], tile[i], x[i], y[i]) for i in range(n_reads)]


def rn_stream(n, seed):
    """RN-like byte stream: names each followed by TAB (ByteArrayStopEncoding)"""
    names = read_names(n // 24 + 2, seed)
    blob = b'\t'.join(names) + b'\t'
    return np.frombuffer(blob[:n], dtype=np.uint8).copy()


def itf8_stream(n, seed):
    """ITF8-series-like stream (BF/CF/RL/AP...): small ints, mostly one byte, a few multi-byte"""
    rng = np.random.default_rng(seed)
    out = rng.choice(np.array([0, 16, 83, 99, 147, 163, 150, 1, 3], dtype=np.uint8), size=n,
                     p=[0.3, 0.3, 0.08, 0.08, 0.08, 0.08, 0.04, 0.02, 0.02])
    return out


STREAM_KINDS = {'qs': qs_stream, 'ba': ba_stream, 'rn': rn_stream, 'itf8': itf8_stream}


# ---------------------------------------------------------------------------------------------- reads

CIGAR_OPS = 'MIDNSHP=X'          # BAM op codes 0..8


def pack_cigar(ops):
    """[(len, 'M'), ...] -> BAM-packed uint32 (len << 4 | op)"""
    return np.array([(ln << 4) | CIGAR_OPS.index(op) for ln, op in ops], dtype=np.uint32)


def reads(ref, n_reads, seed, read_len=150, sub_rate=0.001, indel_frac=0.05, clip_frac=0.02,
          unmapped_tail=0, contig_start=0, contig_end=None, with_names=True):
    """Simulated coordinate-sorted single-end reads over `ref`: per-base substitution rate
    `sub_rate`; `indel_frac` of the reads carry one indel (50/50 I/D, length 1-3, not in the first/last 10
    bases); `clip_frac` carry one soft clip of 5-20 bases at either end; CIGAR uses M; MAPQ 60 w.p. 0.9 else
    uniform 0-59; forward/reverse 50/50; an optional unmapped tail.  Returns a dict of flat numpy arrays
    (structure of arrays, the layout the C ABI takes):
      bases, quals  u8 [sum len]      read_off i64 [n+1]
      cigar u32 BAM-packed            cigar_off i64 [n+1]
      pos i32 [n] (1-based, 0 = unmapped)   flag u16 [n]   mapq u8 [n]   ref_id i32 [n]   names (list of bytes)
    """
    rng = np.random.default_rng(seed)
    L = read_len
    n = n_reads
    hi = (len(ref) if contig_end is None else contig_end) - L - 8
    start = np.sort(rng.integers(contig_start, hi, size=n)).astype(np.int64)       # 0-based
    kind = np.zeros(n, dtype=np.int8)           # 0 plain, 1 insertion, 2 deletion, 3 left clip, 4 right clip
    u = rng.random(n)
    kind[u < indel_frac / 2] = 1
    kind[(u >= indel_frac / 2) & (u < indel_frac)] = 2
    kind[(u >= indel_frac) & (u < indel_frac + clip_frac / 2)] = 3
    kind[(u >= indel_frac + clip_frac / 2) & (u < indel_frac + clip_frac)] = 4
    ev_len = np.where(kind <= 2, rng.integers(1, 4, size=n), rng.integers(5, 21, size=n)).astype(np.int64)
    ev_pos = rng.integers(10, L - 13, size=n).astype(np.int64)
    j = np.arange(L, dtype=np.int64)[None, :]
    p = ev_pos[:, None]
    ln = ev_len[:, None]
    k = kind[:, None]
    shift = np.zeros((n, L), dtype=np.int64)
    shift = np.where((k == 1) & (j >= p), -np.minimum(j - p + 1, ln), shift)     # insertion: ref lags behind
    shift = np.where((k == 2) & (j >= p), ln, shift)                             # deletion: ref jumps ahead
    shift = np.where(k == 3, -ln, shift)                                         # left clip: aligned part starts at ln
    idx = np.clip(start[:, None] + j + shift, 0, len(ref) - 1)
    bases = ref[idx]
    random_bases = ACGT[rng.integers(0, 4, size=(n, L), dtype=np.uint8)]
    novel = ((k == 1) & (j >= p) & (j < p + ln)) | ((k == 3) & (j < ln)) | ((k == 4) & (j >= L - ln))
    bases = np.where(novel, random_bases, bases)
    sub = (rng.random((n, L)) < sub_rate) & ~novel
    other = ACGT[(np.searchsorted(ACGT, np.minimum(bases, ord('T'))) + rng.integers(1, 4, size=(n, L))) % 4]
    bases = np.where(sub, other, bases).astype(np.uint8)
    quals = qualities(n, L, seed + 17)
    mapq = np.where(rng.random(n) < 0.9, 60, rng.integers(0, 60, size=n)).astype(np.uint8)
    flag = np.where(rng.random(n) < 0.5, 0, 16).astype(np.uint16)
    pos = (start + 1).astype(np.int32)
    # CIGARs
    M, I, D, S = 0, 1, 2, 4
    cig = []
    ncig = np.where(kind == 0, 1, np.where(kind <= 2, 3, 2))
    cigar_off = np.zeros(n + 1, dtype=np.int64)
    cigar_off[1:] = np.cumsum(ncig)
    cigar = np.zeros(int(cigar_off[-1]), dtype=np.uint32)
    o = cigar_off[:-1]
    m0 = kind == 0
    cigar[o[m0]] = (L << 4) | M
    m1 = kind == 1
    cigar[o[m1]] = (ev_pos[m1] << 4) | M
    cigar[o[m1] + 1] = (ev_len[m1] << 4) | I
    cigar[o[m1] + 2] = ((L - ev_pos[m1] - ev_len[m1]) << 4) | M
    m2 = kind == 2
    cigar[o[m2]] = (ev_pos[m2] << 4) | M
    cigar[o[m2] + 1] = (ev_len[m2] << 4) | D
    cigar[o[m2] + 2] = ((L - ev_pos[m2]) << 4) | M
    m3 = kind == 3
    cigar[o[m3]] = (ev_len[m3] << 4) | S
    cigar[o[m3] + 1] = ((L - ev_len[m3]) << 4) | M
    m4 = kind == 4
    cigar[o[m4]] = ((L - ev_len[m4]) << 4) | M
    cigar[o[m4] + 1] = (ev_len[m4] << 4) | S
    ref_id = np.zeros(n, dtype=np.int32)
    if unmapped_tail:
        t = unmapped_tail
        ub = ACGT[rng.integers(0, 4, size=(t, L), dtype=np.uint8)]
        bases = np.concatenate([bases, ub])
        quals = np.concatenate([quals, qualities(t, L, seed + 23)])
        mapq = np.concatenate([mapq, np.zeros(t, np.uint8)])
        flag = np.concatenate([flag, np.full(t, 4, np.uint16)])
        pos = np.concatenate([pos, np.zeros(t, np.int32)])
        ref_id = np.concatenate([ref_id, np.full(t, -1, np.int32)])
        cigar_off = np.concatenate([cigar_off, np.full(t, cigar_off[-1], np.int64)])
        n += t
    read_off = np.arange(n + 1, dtype=np.int64) * L
    out = dict(bases=np.ascontiguousarray(bases.reshape(-1)), quals=np.ascontiguousarray(quals.reshape(-1)),
               read_off=read_off, cigar=cigar, cigar_off=cigar_off, pos=pos, flag=flag, mapq=mapq, ref_id=ref_id)
    if with_names:
        out['names'] = read_names(n, seed + 5)
    return out


def cigar_of(rd, i):
    """[(len, op char)] of read i"""
    c = rd['cigar'][rd['cigar_off'][i]:rd['cigar_off'][i + 1]]
    return [(int(x >> 4), CIGAR_OPS[int(x & 15)]) for x in c]
