"""Loader of tests/golden/tracks_jvm.json (the X / P selectors run container after container in the mini-JVM)."""
import base64
import json
import os

import numpy as np

from cramtools_b200 import synth
from tests.golden_records import _parse_cigar

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def load():
    """-> (golden dict, ref u8, containers: list of read dicts (structure of arrays, one per container))"""
    with open(os.path.join(ROOT, 'tests', 'golden', 'tracks_jvm.json')) as f:
        g = json.load(f)
    ref = np.frombuffer(base64.b64decode(g['ref']), np.uint8)
    conts = []
    for reads in g['containers']:
        n = len(reads)
        bases, quals, cig = [], [], []
        read_off = np.zeros(n + 1, np.int64); cigar_off = np.zeros(n + 1, np.int64)
        rflags = np.zeros(n, np.uint8)
        for i, r in enumerate(reads):
            b = base64.b64decode(r['bases'])
            q = base64.b64decode(r['quals']) if r['quals'] is not None else b'\xff' * len(b)
            if r['quals'] is None:
                rflags[i] |= 1
            bases.append(b); quals.append(q)
            read_off[i + 1] = read_off[i] + len(b)
            cg = synth.pack_cigar(_parse_cigar(r['cigar']))
            cig.append(cg); cigar_off[i + 1] = cigar_off[i] + len(cg)
        conts.append(dict(bases=np.frombuffer(b''.join(bases), np.uint8).copy(), quals=np.frombuffer(b''.join(quals), np.uint8).copy(),
                          read_off=read_off, cigar=np.concatenate(cig), cigar_off=cigar_off,
                          pos=np.array([r['pos'] for r in reads], np.int32), flag=np.array([r['flag'] for r in reads], np.uint16),
                          mapq=np.array([r['mapq'] for r in reads], np.uint8), ref_id=np.array([r['ref_id'] for r in reads], np.int32),
                          rflags=rflags))
    return g, ref, conts
