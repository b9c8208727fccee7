"""helpers shared by the record-path tests: load tests/golden/records_jvm.json into the SoA layout"""
import base64
import json
import os

import numpy as np

from cramtools_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _parse_cigar(s):
    out, num = [], ''
    for ch in s:
        if ch.isdigit():
            num += ch
        else:
            out.append((int(num), ch)); num = ''
    return out


def load():
    with open(os.path.join(ROOT, 'tests', 'golden', 'records_jvm.json')) as f:
        g = json.load(f)
    ref = np.frombuffer(base64.b64decode(g['ref']), np.uint8)
    reads = g['reads']
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
        bases.append(b); quals.append(q)
        read_off[i + 1] = read_off[i] + len(b)
        c = synth.pack_cigar(_parse_cigar(r['cigar']))
        cig.append(c)
        cigar_off[i + 1] = cigar_off[i] + len(c)
    rd = dict(bases=np.frombuffer(b''.join(bases), np.uint8).copy(), quals=np.frombuffer(b''.join(quals), np.uint8).copy(),
              read_off=read_off, cigar=np.concatenate(cig) if cig else np.zeros(0, np.uint32), cigar_off=cigar_off,
              pos=np.array([r['pos'] for r in reads], np.int32), flag=np.array([r['flag'] for r in reads], np.uint16),
              mapq=np.array([r['mapq'] for r in reads], np.uint8), ref_id=np.array([r['ref_id'] for r in reads], np.int32),
              rflags=rflags, names=[r['name'] for r in reads])
    return g, ref, rd


def features_as_lists(f_op, f_pos, f_val, f_aux, feat_off, payload, n_reads, coded=True):
    """feature SoA -> the golden file's per-read list form"""
    out = []
    for r in range(n_reads):
        row = []
        for i in range(int(feat_off[r]), int(feat_off[r + 1])):
            op = chr(int(f_op[i]))
            if op == 'X':
                row.append(['X', int(f_pos[i]), int(f_aux[i]) & 0xFF, (int(f_aux[i]) >> 8) & 0xFF, int(f_val[i]) if coded else -1])
            elif op == 'S':
                seq = bytes(payload[int(f_val[i]):int(f_val[i]) + int(f_aux[i])])
                row.append(['S', int(f_pos[i]), base64.b64encode(seq).decode()])
            else:
                row.append([op, int(f_pos[i]), int(f_val[i])])
        out.append(row)
    return out
