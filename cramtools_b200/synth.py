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
